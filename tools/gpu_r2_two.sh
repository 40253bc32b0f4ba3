#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi.py -q -m gpu -p no:cacheprovider -x > gpurun_out/tests_multi.log 2>&1; echo "multi test exit $?"; tail -5 gpurun_out/tests_multi.log
bash tools/gpu_r2_scale.sh 2
