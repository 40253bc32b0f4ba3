#!/bin/bash
# Per-launch device times of 3 views (ncu launch list) + one --set full capture of the hot kernels of the second view.
mkdir -p gpurun_out
CMD="python tools/one_view.py 3"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
tail -1 gpurun_out/plain.log
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'radix_scatter|radix_count|radix_colscan|render_bwd|render_fwd|emit_keys|preprocess_fwd|preprocess_bwd|scan_apply' -s 21 -c 21 \
    -o gpurun_out/prof_r02 $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"; ls -la gpurun_out/*.ncu-rep 2>/dev/null
