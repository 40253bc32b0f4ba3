#!/usr/bin/env python
"""Hot SASS lines of an `ncu --page source --csv` dump: share of samples / instructions + main stall reasons."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr) and r[0].startswith('0x')]
ix = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ix['# Samples']]) for r in data)
totinst = sum(int(r[ix['Instructions Executed']]) for r in data)
print('total samples', tot, 'warp-inst', totinst, 'sass lines', len(data))
for k, r in enumerate(data):
    s = int(r[ix['# Samples']]); n = int(r[ix['Instructions Executed']])
    if s > tot * thr or n > totinst * thr * 1.5:
        print(f"{k:5d} {100*s/tot:5.2f}%smp {100*n/totinst:5.2f}%inst thr {r[ix['Avg. Threads Executed']]:>5s} w={r[ix['stall_wait']]:>5s} br={r[ix['stall_branch_resolving']]:>4s} ssb={r[ix['stall_short_sb']]:>4s} lsb={r[ix['stall_long_sb']]:>4s} ns={r[ix['stall_not_selected']]:>4s} | {r[ix['Source']].strip()[:80]}")
