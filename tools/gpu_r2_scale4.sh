#!/bin/bash
mkdir -p gpurun_out
PORT=29951
run() { local name=$1 n=$2; shift 2; PORT=$((PORT+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT "$@" > gpurun_out/$name.out 2> gpurun_out/$name.err
  echo "== $name rc=$?"; grep '^{' gpurun_out/$name.out | tail -1 | cut -c1-200; }
run scale_n4 4 bench.py --gpus 4 --steps 100 --warmup 10
run c4_n4 4 tools/bench_config4.py
run scale_n2 2 bench.py --gpus 2 --steps 100 --warmup 10
run c4_n2 2 tools/bench_config4.py
