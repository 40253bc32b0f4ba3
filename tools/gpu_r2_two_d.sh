#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -q -m gpu -p no:cacheprovider -x -k "packed or exchange or graphed or two_rank" > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/tests.log
BENCH_TRACE=1 BENCH_TRACE_MIN_US=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 60 --warmup 10 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench2 exit $?"
python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_n2.json').read().strip().splitlines()[-1])
    print(d['value'], d['ms_per_step'], d['e2e']['value'])
except Exception as e: print('bench parse', e)
P
grep "accumulate_packed\|barrier" gpurun_out/timeline_n2_r0.txt | tail -8 | cut -c1-110
