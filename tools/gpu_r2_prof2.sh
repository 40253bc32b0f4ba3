#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/one_view.py 2"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'radix_scatter|radix_count|radix_colscan|scan_reduce|scan_apply' -s 17 -c 17 \
    -o gpurun_out/prof_r02b $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"; ls -la gpurun_out/*.ncu-rep 2>/dev/null
