#!/usr/bin/env python
"""Summary of an ncu launch list (`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
--clock-control none --csv --log-file launches.csv <command>`): per kernel name the number of launches, the average
device time and its share of the total.  Times under ncu are cold-cache and serialised: compare SHARES.
usage: ncu_launches.py launches.csv [title] > profiles/rNN_launches.txt"""
import collections, csv, re, sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10]
hdr = next(r for r in rows if "Kernel Name" in r)
ix = {h: i for i, h in enumerate(hdr)}
t, n, dr, dw = collections.defaultdict(float), collections.Counter(), collections.defaultdict(float), collections.defaultdict(float)
scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
ids = set()
for r in rows:
    if r is hdr or len(r) != len(hdr) or r[ix["ID"]] == "ID":
        continue
    name = re.sub(r"\(.*$", "", r[ix["Kernel Name"]])[:110]
    v = float(r[ix["Metric Value"]].replace(",", "")) * scale.get(r[ix["Metric Unit"]], 1.0)
    m = r[ix["Metric Name"]]
    if m == "gpu__time_duration.sum":
        t[name] += v
        n[name] += 1
        ids.add(r[ix["ID"]])
    elif m == "dram__bytes_read.sum":
        dr[name] += v
    elif m == "dram__bytes_write.sum":
        dw[name] += v
tot = sum(t.values())
print(f"# {sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]}")
print("# per-kernel device time over the captured launches (cold-cache, serialised under ncu: compare shares)")
print(f"# {len(ids)} launches, total {tot:.1f} us")
for name in sorted(t, key=lambda k: -t[k]):
    print(f"{t[name] / n[name]:10.1f} us avg  x{n[name]:5d}  {100 * t[name] / tot:5.1f}%  dram r/w {dr[name] / n[name] / 1e6:8.1f} /{dw[name] / n[name] / 1e6:8.1f} MB  {name}")
