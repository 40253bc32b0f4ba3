#!/bin/bash
mkdir -p gpurun_out
PORT=29811
for ex in auto packed; do
  PORT=$((PORT+1))
  BENCH_STEP_TIMES=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 2 --steps 40 --warmup 8 --exchange $ex --no-train-iter > gpurun_out/b2_$ex.out 2> gpurun_out/b2_$ex.err
  echo "== $ex rc=$?"; grep "step ms" gpurun_out/b2_$ex.err | cut -c1-400
  python - <<P
import json
d=json.loads([l for l in open('gpurun_out/b2_$ex.out') if l.startswith('{')][-1])
print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])
P
done
