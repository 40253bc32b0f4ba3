#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_scale.py tests/test_gpu_parity.py -q -m gpu -p no:cacheprovider -x > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -5 gpurun_out/tests.log
CMD="python tools/one_view.py 3"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list exit $?"
timeout 300 python bench.py --steps 30 --warmup 5 --no-train-iter --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python - <<'P'
import json
try:
    d=json.load(open('gpurun_out/bench.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'], d['stages_ms'])
except Exception as e: print('bench parse', e)
P
