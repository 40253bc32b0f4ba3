#!/usr/bin/env python
"""Hot loop of one kernel as SASS with per-instruction execution counts, from `ncu -i X.ncu-rep --page source --csv`.
usage: ncu_sass_hot.py dump.csv <kernel substring> [min executed per launch, default 1e6]
Ends with the opcode mix (share of executed warp instructions) and the checks the round-1 verdict asked for:
packed fp32 (FFMA2/FMUL2/FADD2), MUFU per kind, local-memory traffic (LDL/STL: spills)."""
import collections, csv, sys
path, pat = sys.argv[1], sys.argv[2]
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 1e6
rows = list(csv.reader(open(path)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
i0 = [i for i in starts if pat in rows[i][1]][0]
i1 = min([s for s in starts if s > i0] + [len(rows)])
hdr = rows[i0 + 1]; ix = {h: k for k, h in enumerate(hdr)}
body = [r for r in rows[i0 + 2:i1] if len(r) == len(hdr)]
print("# " + rows[i0][1][:150])
print(f"# {len(body)} SASS instructions; listed: those executed >= {thr:.0f} times in the profiled launch (warp-level counts)")
mix, tot = collections.Counter(), 0.0
for k, r in enumerate(body):
    ex = float(r[ix["Instructions Executed"]] or 0)
    src = r[ix["Source"]].split()
    op = src[1] if src and src[0].startswith("@") else (src[0] if src else "")
    mix[op.split(".")[0] + ("." + op.split(".")[1] if op.startswith("MUFU") and "." in op else "")] += ex
    tot += ex
    if ex >= thr:
        print(f"{k:5d} {ex / 1e6:8.2f}M  {r[ix['Source']]}")
print(f"# executed warp instructions: {tot / 1e6:.1f} M; opcode mix:")
for op, v in mix.most_common(24):
    print(f"#   {op:12s} {v / 1e6:8.1f} M {100 * v / tot:5.1f} %")
print(f"# local memory (spills): LDL {mix.get('LDL', 0) / 1e6:.2f} M, STL {mix.get('STL', 0) / 1e6:.2f} M")
