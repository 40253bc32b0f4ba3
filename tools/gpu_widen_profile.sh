#!/bin/bash
# ncu --set full capture of the (rewritten) loss kernels; the plain run of the same command comes first.
mkdir -p gpurun_out
CMD2="python tools/next_kernels_once.py"
timeout 200 $CMD2 > gpurun_out/next_kernels_plain.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'l1_ssim_fwd_kernel|l1_ssim_bwd_kernel|l1_ssim_loss_kernel' \
    -s 3 -c 3 -o gpurun_out/prof_loss_v2 -f $CMD2 > gpurun_out/next_kernels_ncu.log 2>&1
echo "full capture exit $?"
ls -la gpurun_out/*.ncu-rep
