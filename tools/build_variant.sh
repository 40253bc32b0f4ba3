#!/bin/bash
# tools/build_variant.sh NAME "-DFLAG=..": builds 3dgs_study_b200/variants/libb200splat_NAME.so with extra nvcc flags
set -e
cd "$(dirname "$0")/../3dgs_study_b200/csrc"
NAME=$1; shift
OUT=../variants; mkdir -p $OUT/obj_$NAME
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v --expt-relaxed-constexpr -ccbin /usr/bin/g++ $*"
for f in api preprocess binning render_tiles; do rm -f $OUT/obj_$NAME/$f.o;
  $NV -c $f.cu -o $OUT/obj_$NAME/$f.o 2> $OUT/obj_$NAME/$f.log &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libb200splat_$NAME.so $OUT/obj_$NAME/*.o -Xlinker --version-script=exports.map
grep -h "registers\|spill" $OUT/obj_$NAME/render_tiles.log | paste - - | awk '{print $0}' | head -8
