#!/bin/bash
# Round 2, first GPU session: the widened parity suite on the round-1 kernels, the peaks microbenchmarks, a bench line.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt
./tools/b200_peaks > gpurun_out/peaks.json 2> gpurun_out/peaks.err; echo "peaks exit $?"; cat gpurun_out/peaks.json
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider --durations=15 > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -40 gpurun_out/tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 400 python bench.py --steps 50 --warmup 10 --no-train-iter > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; head -c 1500 gpurun_out/bench.json; echo; tail -5 gpurun_out/bench.err
