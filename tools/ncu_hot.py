#!/usr/bin/env python
"""Top SASS instructions by warp-stall samples for one kernel of an `ncu --page source --csv` dump.
usage: ncu_hot.py dump.csv <kernel substring> [occurrence=0] [top=25]"""
import csv, sys
path, pat = sys.argv[1], sys.argv[2]
occ = int(sys.argv[3]) if len(sys.argv) > 3 else 0
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
rows = list(csv.reader(open(path)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
sel = [i for i in starts if pat in rows[i][1]]
i0 = sel[occ]; i1 = min([s for s in starts if s > i0] + [len(rows)])
hdr = rows[i0 + 1]; ix = {h: k for k, h in enumerate(hdr)}
body = [r for r in rows[i0 + 2:i1] if len(r) == len(hdr)]
S = "Warp Stall Sampling (All Samples)"
tot = sum(float(r[ix[S]] or 0) for r in body)
print(f"{rows[i0][1][:100]}\n{len(body)} instructions, {tot:.0f} samples")
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: sum(float(r[ix[h]] or 0) for r in body) for h in stall_cols}
print("stall mix: " + ", ".join(f"{h[6:]}={100*v/max(tot,1):.1f}%" for h, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
order = sorted(range(len(body)), key=lambda k: -float(body[k][ix[S]] or 0))[:top]
for k in sorted(order):
    r = body[k]
    st = sorted(((h[6:], float(r[ix[h]] or 0)) for h in stall_cols), key=lambda x: -x[1])[:2]
    print(f"{k:5d} {100*float(r[ix[S]] or 0)/max(tot,1):5.1f}%  exec {r[ix['Instructions Executed']]:>9s}  {r[ix['Source']][:70]:70s} {st[0][0]}:{st[0][1]:.0f} {st[1][0]}:{st[1][1]:.0f}")
