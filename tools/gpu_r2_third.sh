#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_scale.py tests/test_gpu_parity.py tests/test_gpu_dropin.py -q -m gpu -p no:cacheprovider -x > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -12 gpurun_out/tests.log
bash tools/gpu_r2_prof.sh
