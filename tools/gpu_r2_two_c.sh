#!/bin/bash
# 2 GPUs: the whole GPU suite (incl. the 2-rank NCCL exchange test), the exchange check, N=2 bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/tests2.log 2>&1; echo "tests exit $?"; tail -5 gpurun_out/tests2.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/check_exchange.py > gpurun_out/exchange_check_n2.txt 2>&1; echo "check exit $?"; tail -6 gpurun_out/exchange_check_n2.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 60 --warmup 10 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench2 exit $?"; tail -2 gpurun_out/bench_n2.err
python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_n2.json').read().strip().splitlines()[-1])
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('exchange'), d.get('step_submission'))
except Exception as e: print('bench parse', e)
P
