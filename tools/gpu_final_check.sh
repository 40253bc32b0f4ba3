#!/bin/bash
# end-of-round verification on one B200: the whole GPU test suite, smoke(), a short bench.py, the translucent sweep point
mkdir -p gpurun_out; rm -f gpurun_out/widen_observed.jsonl
timeout 600 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/final_gpu_tests.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/final_gpu_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/final_smoke.log
timeout 300 python bench.py --steps 50 --warmup 10 --no-cpu-baseline --no-train-iter > gpurun_out/final_bench_short.json 2> gpurun_out/final_bench.err; echo "bench exit $?"; cut -c1-260 gpurun_out/final_bench_short.json
timeout 300 python tools/bench_configs.py sweep 3000000 translucent > gpurun_out/cfg_translucent.jsonl 2>/dev/null; cut -c1-420 gpurun_out/cfg_translucent.jsonl
