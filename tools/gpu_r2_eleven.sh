#!/bin/bash
mkdir -p gpurun_out
export B200SPLAT_LIB=$PWD/3dgs_study_b200/variants/libb200splat_lean4.so
for combo in "0 0" "1 0" "0 1" "1 1"; do
  set -- $combo
  export B200_EXP_STATS_SIDE=$1 B200_EXP_CLEAN=$2
  timeout 300 python bench.py --steps 80 --warmup 10 --no-train-iter --no-cpu-baseline > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err; echo "stats_side=$1 clean=$2 exit $?"; tail -2 gpurun_out/bench_x.err
  python - <<P
import json
try:
    d=json.load(open('gpurun_out/bench_x.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'])
except Exception as e: print('bench parse', e)
P
done
export B200_EXP_STATS_SIDE=1 B200_EXP_CLEAN=1
BENCH_TRACE=1 timeout 300 python bench.py --steps 30 --warmup 8 --no-train-iter --no-cpu-baseline > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err; echo "trace exit $?"
tail -45 gpurun_out/timeline_n1_r0.txt | cut -c1-125
