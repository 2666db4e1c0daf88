#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import collections, csv, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
tot, cnt = collections.defaultdict(float), collections.Counter()
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    n = row['Kernel Name'].split('(')[0].replace('hs::', '')
    v = float(row['Metric Value'].replace(',', ''))
    u = row['Metric Unit']
    v = v / 1e3 if u == 'ns' else (v * 1e3 if u == 'ms' else v)
    tot[n] += v
    cnt[n] += 1
s = sum(tot.values())
print(f"total {s:.1f} us over {sum(cnt.values())} launches")
for n in sorted(tot, key=lambda n: -tot[n]):
    print(f"{n[:60]:60s} {cnt[n]:4d} x {tot[n] / cnt[n]:9.1f} us = {tot[n]:9.1f} us {100 * tot[n] / s:5.1f}%")
