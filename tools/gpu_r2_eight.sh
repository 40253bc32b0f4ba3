#!/bin/bash
# One 8-GPU session: scaling of the bench (N = 2, 4, 8), configs[3] (N = 2, 4, 8), both opacity presets, packed vs factorized.
mkdir -p gpurun_out
PORT=29711
run() { # name nproc args...
  local name=$1 n=$2; shift 2
  PORT=$((PORT+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT "$@" > gpurun_out/$name.out 2> gpurun_out/$name.err
  echo "== $name rc=$?"; grep '^{' gpurun_out/$name.out | tail -1 | cut -c1-330
}
run exchange_n8 8 tools/check_exchange.py; tail -1 gpurun_out/exchange_n8.out
for n in 8 4 2; do
  BENCH_TRACE=$([ $n = 8 ] && echo 1) run bench_n$n $n bench.py --gpus $n --steps 40 --warmup 8
done
for n in 8 4 2; do run config4_n$n $n tools/bench_config4.py; done
run bench_translucent_packed_n8 8 bench.py --gpus 8 --steps 30 --warmup 6 --preset translucent --exchange packed
run bench_translucent_factorized_n8 8 bench.py --gpus 8 --steps 30 --warmup 6 --preset translucent --exchange factorized
run bench_opaque_factorized_n8 8 bench.py --gpus 8 --steps 30 --warmup 6 --exchange factorized
run bench_ref_n8 8 bench.py --impl reference --gpus 8 --steps 3 --warmup 2
