#!/usr/bin/env python
"""Selected raw metrics of one kernel from an `ncu --page raw --csv` export.
  python tools/ncu_metric.py RAW.csv KERNEL_PREFIX SUBSTR [SUBSTR ...]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ki = hdr.index('Kernel Name')
sel = [r for r in rows[2:] if r[ki].replace('void ', '').startswith(sys.argv[2])]
r = sel[min(1, len(sel) - 1)]
for i, h in enumerate(hdr):
    if any(t in h for t in sys.argv[3:]):
        print(f"{h:90s} {r[i]:>16s} {units[i]}")
