#!/bin/bash
# One GPU session that re-establishes every number of a round on a fresh B200 (about 3 minutes of box time):
#   1. the whole GPU test suite and smoke()            -> gpurun_out/tests.log, smoke.log
#   2. bench.py at N=1 (headline + e2e + roofline + training iteration)  -> gpurun_out/bench.json
#   3. widened-row benches                              -> gpurun_out/bench_next.jsonl
#   4. ncu launch list of bench.py (device time + DRAM bytes per launch; follows the plain run of the same command)
# Usage:  gpurun --timeout 900 -- 'bash tools/gpu_round_start.sh'
# then:   python profiles/summarize_ncu.py launches gpurun_out/launches.csv > profiles/rNN_launches.txt
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
timeout 300 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 400 python bench.py --steps 100 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; head -c 600 gpurun_out/bench.json; echo
timeout 150 python tools/bench_next.py > gpurun_out/bench_next.log 2>&1; echo "bench_next exit $?"
CMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-train-iter"
timeout 200 $CMD > gpurun_out/bench_plain.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/bench_ncu.log 2>&1
echo "launch list exit $?"
