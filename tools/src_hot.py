#!/usr/bin/env python
"""Hot SASS of one kernel from `ncu --page source --csv`: top stall-sample instructions and an opcode histogram."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
data = [r for r in rows[2:] if len(r) == len(hdr)]
ia, iss, ie = hdr.index('Source'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed')
num = lambda x: int(x) if x.strip().isdigit() else 0
print('samples', sum(num(r[iss]) for r in data), 'warp-instr', sum(num(r[ie]) for r in data), 'sass', len(data))
top = sorted(range(len(data)), key=lambda i: -num(data[i][iss]))[:int(sys.argv[2]) if len(sys.argv) > 2 else 30]
for i in sorted(top):
    r = data[i]
    print(i, r[iss], r[ie], r[ia][:100])
h, hs = collections.Counter(), collections.Counter()
for r in data:
    op = r[ia].strip().split()
    if not op:
        continue
    o = (op[1] if op[0].startswith('@') and len(op) > 1 else op[0]).split('.')[0]
    h[o] += num(r[ie]); hs[o] += num(r[iss])
print([(o, c, hs[o]) for o, c in h.most_common(20)])
