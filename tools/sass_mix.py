#!/usr/bin/env python
"""Instruction mix (by opcode) and stall-sample totals of an `ncu --page source --csv` dump (first kernel instance)."""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
data = []
for r in rows[2:]:
    if len(r) != len(hdr) or not r[0].startswith('0x'):
        break
    data.append(r)
ix = {h: i for i, h in enumerate(hdr)}
mix = collections.Counter(); smp = collections.Counter()
tot = 0; tots = 0
for r in data:
    src = r[ix['Source']].strip()
    m = re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_.]+)', src)
    op = m.group(2).split('.')[0] if m else src[:10]
    n = int(r[ix['Instructions Executed']]); s = int(r[ix['# Samples']])
    mix[op] += n; smp[op] += s; tot += n; tots += s
print(f"sass lines {len(data)}  warp-inst {tot}  samples {tots}")
for op, n in mix.most_common(28):
    print(f"  {op:10s} {100*n/tot:6.2f}% inst  {100*smp[op]/max(tots,1):6.2f}% samples")
st = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {h: sum(int(r[ix[h]]) for r in data) for h in st}
print('stalls:', ', '.join(f"{h[6:]}={100*v/max(tots,1):.1f}%" for h, v in sorted(agg.items(), key=lambda x: -x[1]) if v))
