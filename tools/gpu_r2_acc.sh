#!/bin/bash
mkdir -p gpurun_out
for b in 4 5 3; do
export B200_EXP_ACC_BPS=$b
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'accumulate_packed' -s 4 -c 4 --csv --log-file gpurun_out/acc.csv python tools/bench_config4.py --steps 2 --warmup 2 > /dev/null 2>&1
echo "blocks per SM $b"; python tools/ncu_launches.py gpurun_out/acc.csv acc | tail -2
done
