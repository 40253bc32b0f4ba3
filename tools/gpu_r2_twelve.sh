#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_scale.py -q -m gpu -p no:cacheprovider -x > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/tests.log
for combo in "0 0" "1 1"; do
  set -- $combo
  export B200_EXP_STATS_SIDE=$1 B200_EXP_CLEAN=$2
  timeout 300 python bench.py --steps 80 --warmup 10 --no-train-iter --no-cpu-baseline > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err; echo "stats_side=$1 clean=$2 exit $?"; tail -2 gpurun_out/bench_x.err
  python - <<P
import json
try:
    d=json.load(open('gpurun_out/bench_x.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'], d['stages_ms'])
except Exception as e: print('bench parse', e)
P
done
BENCH_TRACE=1 BENCH_TRACE_MIN_US=0 timeout 300 python bench.py --steps 30 --warmup 8 --no-train-iter --no-cpu-baseline > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err; echo "trace exit $?"
tail -75 gpurun_out/timeline_n1_r0.txt | cut -c1-118
