#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -p no:cacheprovider -x -k "graphed or factorized or capacity" > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -15 gpurun_out/tests.log
for flag in "" "--no-graph"; do
timeout 300 python bench.py --steps 40 --warmup 8 --no-train-iter --no-cpu-baseline $flag > gpurun_out/bench$flag.json 2> gpurun_out/bench$flag.err; echo "bench $flag exit $?"; tail -3 gpurun_out/bench$flag.err
python - <<P
import json
try:
    d=json.load(open('gpurun_out/bench$flag.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'], d['step_submission'][:40])
except Exception as e: print('bench parse', e)
P
done
