#!/bin/bash
# end-of-round verification on one B200: the whole GPU test suite, smoke(), a short bench.py (with the training iteration)
mkdir -p gpurun_out; rm -f gpurun_out/widen_observed.jsonl
timeout 300 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/final_gpu_tests.log 2>&1; echo "exit $?" >> gpurun_out/final_gpu_tests.log; tail -4 gpurun_out/final_gpu_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "exit $?" >> gpurun_out/final_smoke.log; tail -3 gpurun_out/final_smoke.log
timeout 300 python bench.py --steps 50 --warmup 10 --no-cpu-baseline > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench exit $?"; head -c 1500 gpurun_out/final_bench.json; tail -3 gpurun_out/final_bench.err
timeout 100 python tools/bench_next.py > gpurun_out/bench_next.log 2>&1; echo "exit $?" >> gpurun_out/bench_next.log
