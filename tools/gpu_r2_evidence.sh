#!/bin/bash
# ncu evidence of the shipped build: launch list of the bench command, then one --set full capture of one view's kernels.
# Each ncu pass runs only after the same command has exited 0 without it.
mkdir -p gpurun_out
BENCH="python bench.py --steps 3 --warmup 3 --no-train-iter --no-cpu-baseline"
timeout 300 $BENCH > gpurun_out/ev_bench_plain.json 2> gpurun_out/ev_bench_plain.err || { echo "plain bench failed"; tail -5 gpurun_out/ev_bench_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 \
    --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
echo "launch list exit $?"; wc -l gpurun_out/launches.csv
CMD="python tools/one_view.py 3"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
tail -1 gpurun_out/plain.log
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'radix_scatter|radix_count|radix_colscan|render_bwd|render_fwd|emit_keys|preprocess_fwd|preprocess_bwd|scan_spine|compact_active|zero_fill' -s 26 -c 26 \
    -o gpurun_out/prof_r02_final $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"; ls -la gpurun_out/*.ncu-rep 2>/dev/null
