#!/bin/bash
# tools/build_variant.sh NAME "-DFLAG=..": builds 3dgs_study_b200/variants/libb200splat_NAME.so with extra nvcc flags
# (kernel-variant experiments; B200SPLAT_LIB=<that file> selects it).  next_steps / ply_io come from the product build.
set -e
cd "$(dirname "$0")/../3dgs_study_b200/csrc"
NAME=$1; shift
OUT=../variants; mkdir -p $OUT/obj_$NAME
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v --expt-relaxed-constexpr -ccbin /usr/bin/g++ $*"
for f in api preprocess binning render_tiles; do rm -f $OUT/obj_$NAME/$f.o;
  $NV -c $f.cu -o $OUT/obj_$NAME/$f.o 2> $OUT/obj_$NAME/$f.log &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libb200splat_$NAME.so $OUT/obj_$NAME/*.o next_steps.o ply_io.o -Xlinker --version-script=exports.map -lpthread
grep -h "registers\|spill" $OUT/obj_$NAME/binning.log | paste - - | grep -i "64 reg\|7[0-9] reg\|8[0-9] reg" | head -4
