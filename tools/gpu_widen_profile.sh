#!/bin/bash
# knn parity re-check + ncu launch list (device time, DRAM bytes) of one fused training iteration.
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_widen.py -q -m gpu -k knn -p no:cacheprovider > gpurun_out/widen_knn.log 2>&1
echo "exit $?" >> gpurun_out/widen_knn.log
tail -4 gpurun_out/widen_knn.log
CMD="python tools/train_iter_bench.py --arms fused --iters 2 --warmup 3"
timeout 200 $CMD > gpurun_out/train_iter_plain.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 900 \
    --csv --log-file gpurun_out/train_iter_launches.csv $CMD > gpurun_out/train_iter_ncu.log 2>&1
echo "ncu exit $?"
tail -3 gpurun_out/train_iter_plain.log
wc -l gpurun_out/train_iter_launches.csv
