#!/bin/bash
# usage: bash tools/gpu_r2_scale.sh N   (under gpurun --gpus N): exchange check + bench at N ranks (+ timeline of rank 0)
N=${1:-8}
mkdir -p gpurun_out
PORT=29611
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT tools/check_exchange.py > gpurun_out/exchange_n$N.txt 2>&1; echo "exchange check exit $?"; tail -3 gpurun_out/exchange_n$N.txt
BENCH_TRACE=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((PORT+1)) bench.py --gpus $N --steps 50 --warmup 10 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench exit $?"
python - <<P
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_n$N.json') if l.startswith('{')][-1])
    print('N=$N value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'e2e+img', d['e2e_with_image_readback']['value'])
except Exception as e: print('parse', e)
P
tail -3 gpurun_out/bench_n$N.err
