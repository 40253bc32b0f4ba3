#!/bin/bash
# 8-GPU box: bench.py at N = 8 as the driver launches it (with the rank-0 timeline), configs[3] at N = 8
mkdir -p gpurun_out
PORT=29931
run() { local name=$1 n=$2; shift 2; PORT=$((PORT+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT "$@" > gpurun_out/$name.out 2> gpurun_out/$name.err
  echo "== $name rc=$?"; grep '^{' gpurun_out/$name.out | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d.get('value'), d.get('ms_per_step'), 'e2e', d.get('e2e',{}).get('value'), 'img', d.get('e2e_with_image_readback',{}).get('value'), d.get('step_submission','')[:30], d.get('views_per_s'))"; }
BENCH_TRACE=1 BENCH_TRACE_MIN_US=0 run scale_n8 8 bench.py --gpus 8 --steps 100 --warmup 10
grep "accumulate_packed\|barrier\|preprocess_fwd" gpurun_out/timeline_n8_r0.txt | tail -10 | cut -c1-110
run c4_n8 8 tools/bench_config4.py
