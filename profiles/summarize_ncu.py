#!/usr/bin/env python
"""Summarise ncu outputs brought back in gpurun_out/ into small text files under profiles/.

  python profiles/summarize_ncu.py launches gpurun_out/launches.csv  > profiles/rNN_launches.txt
  python profiles/summarize_ncu.py raw      gpurun_out/raw.csv       > profiles/rNN_kernels.txt
(raw.csv = `ncu -i X.ncu-rep --page raw --csv`)
"""
import collections
import csv
import re
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_xu.sum',
        'sm__inst_executed_pipe_lsu.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'lts__t_sectors_op_red.sum', 'lts__t_sectors_op_atom.sum',
        'sm__cycles_elapsed.max', 'smsp__cycles_active.avg']


def launches(path):
    lines = [l for l in open(path) if l.startswith('"')]
    agg, tot = collections.OrderedDict(), 0.0
    scale = {'byte': 1e-6, 'Kbyte': 1e-3, 'Mbyte': 1.0, 'Gbyte': 1e3}
    for row in csv.DictReader(lines):
        v = float(row['Metric Value'].replace(',', ''))
        u = row['Metric Unit']
        k = re.sub(r'\(.*', '', row['Kernel Name'])[:100]
        a = agg.setdefault(k, [0, 0.0, 0.0])
        if row['Metric Name'] == 'gpu__time_duration.sum':
            v = v / 1e3 if u in ('ns', 'nsecond') else (v * 1e3 if u in ('ms', 'msecond') else v)
            a[0] += 1; a[1] += v; tot += v
        elif row['Metric Name'].startswith('dram__bytes'):
            a[2] += v * scale.get(u, 1e-6)
    print(f"# per-kernel device time over the captured launches (cold-cache, serialised: compare shares); DRAM MB = read + write per launch\n# total {tot:.1f} us")
    for k, (n, t, mb) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{t / n:10.1f} us avg  x{n:4d}  {100 * t / tot:5.1f}%  {mb / max(n, 1):9.1f} MB  {k}")


def raw(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    for d in data:
        print('---- ' + d[idx['Kernel Name']][:110])
        for k in KEYS:
            if k in idx:
                print(f"  {k:66s} {d[idx[k]][:24]:>24s} {units[idx[k]]}")


if __name__ == '__main__':
    {'launches': launches, 'raw': raw}[sys.argv[1]](sys.argv[2])
