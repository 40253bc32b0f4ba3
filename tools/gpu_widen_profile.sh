#!/bin/bash
# ncu evidence for the widened rows: launch list (device time, DRAM bytes) of one fused training iteration, and one
# --set full capture of each new kernel.  Each ncu run follows a plain run of the same command that exited 0.
mkdir -p gpurun_out
CMD="python tools/train_iter_bench.py --arms fused --iters 2 --warmup 3"
timeout 200 $CMD > gpurun_out/train_iter_plain.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 900 \
    --csv --log-file gpurun_out/train_iter_launches.csv $CMD > gpurun_out/train_iter_ncu.log 2>&1
echo "launch list exit $?"
CMD2="python tools/next_kernels_once.py"
timeout 200 $CMD2 > gpurun_out/next_kernels_plain.log 2>&1 &&
timeout 500 ncu --set full --clock-control none --import-source on \
    -k regex:'l1_ssim_fwd_kernel|l1_ssim_bwd_kernel|adam_kernel|knn_search_kernel|activate_fwd_kernel|activate_bwd_kernel' \
    -s 2 -c 10 -o gpurun_out/prof_next $CMD2 > gpurun_out/next_kernels_ncu.log 2>&1
echo "full capture exit $?"
tail -2 gpurun_out/next_kernels_plain.log gpurun_out/next_kernels_ncu.log
ls -la gpurun_out/*.ncu-rep
