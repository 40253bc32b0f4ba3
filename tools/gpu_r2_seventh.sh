#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_dropin.py -q -m gpu -p no:cacheprovider -x > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -4 gpurun_out/tests.log
BENCH_TRACE=1 timeout 300 python bench.py --steps 40 --warmup 8 --no-train-iter --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; tail -3 gpurun_out/bench.err
python - <<P
import json
try:
    d=json.load(open('gpurun_out/bench.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'], d['stages_ms'])
except Exception as e: print('bench parse', e)
P
tail -40 gpurun_out/timeline_n1_r0.txt | cut -c1-125
