#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_widen.py -q -m gpu -p no:cacheprovider -x > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -4 gpurun_out/tests.log
timeout 600 python tools/bench_configs.py all > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs exit $?"; cat gpurun_out/configs.jsonl | cut -c1-400
timeout 300 python tools/bench_configs.py sweep 3000000 translucent > gpurun_out/configs_translucent.jsonl 2>> gpurun_out/configs.err; cat gpurun_out/configs_translucent.jsonl | cut -c1-400
timeout 300 python bench.py --steps 30 --warmup 5 --no-train-iter --no-cpu-baseline --preset translucent > gpurun_out/bench_translucent.json 2> gpurun_out/bench_translucent.err; echo "bench translucent exit $?"
python - <<'P'
import json
try:
    d=json.load(open('gpurun_out/bench_translucent.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['stages_ms'], d['workload_detail'])
except Exception as e: print('bench parse', e)
P
timeout 200 python tools/bench_config4.py > gpurun_out/config4_n1.jsonl 2> gpurun_out/config4.err; cat gpurun_out/config4_n1.jsonl
