#!/bin/bash
# Builds the COMPARATOR library 3dgs_study_b200/variants/libb200splat_classic.so (baseline-algorithm variant, see
# csrc/classic/binning_classic.cu).  Never loaded unless B200SPLAT_LIB points at it (`bench.py --impl classic`).
set -e
cd "$(dirname "$0")/../3dgs_study_b200/csrc"
OUT=../variants; mkdir -p $OUT/obj_classic
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fvisibility=hidden --expt-relaxed-constexpr -ccbin /usr/bin/g++"
$NV -c api.cu -o $OUT/obj_classic/api.o &
$NV -c preprocess.cu -o $OUT/obj_classic/preprocess.o &
$NV -c classic/binning_classic.cu -o $OUT/obj_classic/binning_classic.o &
$NV -c classic/render_classic.cu -o $OUT/obj_classic/render_classic.o &
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libb200splat_classic.so $OUT/obj_classic/*.o -Xlinker --version-script=exports.map
echo built $OUT/libb200splat_classic.so
