#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'accumulate_packed' -s 4 -c 6 --csv --log-file gpurun_out/acc.csv python tools/bench_config4.py --steps 3 --warmup 2 > /dev/null 2>&1
python tools/ncu_launches.py gpurun_out/acc.csv acc
