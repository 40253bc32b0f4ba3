#!/bin/bash
# two more points of BASELINE configs[2] (densification sweep) with the final build
mkdir -p gpurun_out
timeout 24 python tools/bench_configs.py sweep 1000000,6000000 > gpurun_out/config3_more_final.jsonl 2> gpurun_out/config3_final.err; echo "sweep exit $?"; cat gpurun_out/config3_more_final.jsonl | cut -c1-400
