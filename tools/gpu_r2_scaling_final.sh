#!/bin/bash
# Final scaling session of round 2 (one 8-GPU box): bench.py at N = 8, 4, 2 as the driver launches it, N = 8 timeline.
mkdir -p gpurun_out
PORT=29911
run() { local name=$1 n=$2; shift 2; PORT=$((PORT+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT "$@" > gpurun_out/$name.out 2> gpurun_out/$name.err
  echo "== $name rc=$?"; grep '^{' gpurun_out/$name.out | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d.get('value'), d.get('ms_per_step'), 'e2e', d.get('e2e',{}).get('value'), 'img', d.get('e2e_with_image_readback',{}).get('value'), d.get('step_submission','')[:30])"; grep "step ms" gpurun_out/$name.err | cut -c1-300; }
BENCH_TRACE=1 BENCH_STEP_TIMES=1 run scale_n8 8 bench.py --gpus 8 --steps 40 --warmup 8
BENCH_STEP_TIMES=1 run scale_n4 4 bench.py --gpus 4 --steps 40 --warmup 8
BENCH_STEP_TIMES=1 run scale_n2 2 bench.py --gpus 2 --steps 40 --warmup 8
BENCH_STEP_TIMES=1 run scale_n8_nograph 8 bench.py --gpus 8 --steps 40 --warmup 8 --no-graph
run ref_n8 8 bench.py --impl reference --gpus 8 --steps 3 --warmup 2
