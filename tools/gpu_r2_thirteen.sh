#!/bin/bash
mkdir -p gpurun_out
for combo in "0 0" "1 0" "0 1" "1 1"; do
  set -- $combo
  export B200_EXP_NOSTATUS=$1 B200_EXP_NOCLEAR=$2
  BENCH_TRACE=1 BENCH_TRACE_MIN_US=0 timeout 300 python bench.py --steps 60 --warmup 10 --no-train-iter --no-cpu-baseline > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err; echo "nostatus=$1 noclear=$2 exit $?"; tail -2 gpurun_out/bench_x.err
  python - <<P
import json
try:
    d=json.load(open('gpurun_out/bench_x.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'])
except Exception as e: print('bench parse', e)
P
  grep -B3 -A1 "render_bwd" gpurun_out/timeline_n1_r0.txt | tail -12 | cut -c1-110
done
