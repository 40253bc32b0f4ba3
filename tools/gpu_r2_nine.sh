#!/bin/bash
# the no-allocation entry point alone in a fresh process, then the whole GPU suite (it runs last), smoke, the bench lines
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider -x > gpurun_out/noalloc_alone.log 2>&1; echo "noalloc alone exit $?"; tail -3 gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider -x --durations=8 > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -16 gpurun_out/tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench.err; echo "bench exit $?"; tail -3 gpurun_out/bench.err; cat gpurun_out/bench_n1.json | cut -c1-1500
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_n1.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; tail -3 gpurun_out/bench_ref.err; cat gpurun_out/bench_ref_n1.json | cut -c1-800
