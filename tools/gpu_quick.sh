#!/bin/bash
# short GPU session: widened-row parity tests and their benches
mkdir -p gpurun_out; rm -f gpurun_out/widen_observed.jsonl
timeout 240 python -m pytest tests/test_gpu_widen.py -q -m gpu -p no:cacheprovider > gpurun_out/widen_tests.log 2>&1; echo "exit $?" >> gpurun_out/widen_tests.log; tail -5 gpurun_out/widen_tests.log
timeout 150 python tools/bench_next.py > gpurun_out/bench_next.log 2>&1; echo "exit $?" >> gpurun_out/bench_next.log; head -c 900 gpurun_out/bench_next.log
