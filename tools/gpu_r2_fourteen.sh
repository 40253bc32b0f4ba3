#!/bin/bash
mkdir -p gpurun_out
for q in 8 2; do
  export B200_EXP_ZERO_Q=$q
  BENCH_TRACE=1 BENCH_TRACE_MIN_US=0 timeout 300 python bench.py --steps 60 --warmup 10 --no-train-iter --no-cpu-baseline > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err; echo "zero blocks per 4 SMs=$q exit $?"; tail -1 gpurun_out/bench_x.err
  python - <<P
import json
try:
    d=json.load(open('gpurun_out/bench_x.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'])
except Exception as e: print('bench parse', e)
P
  grep -B3 -A6 "render_bwd" gpurun_out/timeline_n1_r0.txt | tail -11 | cut -c1-110
done
