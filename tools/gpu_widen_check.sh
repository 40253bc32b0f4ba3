#!/bin/bash
# One short GPU session for the 8f pieces: parity tests, kernel timings next to the reference's own code on
# the same box, the full training iteration.  Every step has its own timeout; outputs land in gpurun_out/.
mkdir -p gpurun_out
rm -f gpurun_out/widen_observed.jsonl
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/widen_gpu.txt 2>&1
timeout 240 python -m pytest tests/test_gpu_widen.py -q -m gpu -p no:cacheprovider > gpurun_out/widen_tests.log 2>&1
echo "exit $?" >> gpurun_out/widen_tests.log
tail -15 gpurun_out/widen_tests.log
timeout 150 python tools/bench_next.py > gpurun_out/bench_next.log 2>&1
echo "exit $?" >> gpurun_out/bench_next.log
tail -6 gpurun_out/bench_next.log
timeout 150 python tools/train_iter_bench.py > gpurun_out/train_iter.log 2>&1
echo "exit $?" >> gpurun_out/train_iter.log
tail -4 gpurun_out/train_iter.log
