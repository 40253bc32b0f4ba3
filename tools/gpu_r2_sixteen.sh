#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -p no:cacheprovider -x > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -4 gpurun_out/tests.log
for prio in 0; do
export B200_EXP_PRIO=$prio
BENCH_TRACE=1 BENCH_TRACE_MIN_US=0 timeout 300 python bench.py --steps 80 --warmup 10 --no-train-iter --no-cpu-baseline > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err; echo "prio=$prio bench exit $?"; tail -1 gpurun_out/bench_x.err
python - <<P
import json
try:
    d=json.load(open('gpurun_out/bench_x.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'])
except Exception as e: print('bench parse', e)
P
tail -60 gpurun_out/timeline_n1_r0.txt | grep -v "radix_\|scan_spine\|emit_keys\|tile_order" | tail -22 | cut -c1-112
done
