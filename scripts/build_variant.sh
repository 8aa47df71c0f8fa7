#!/bin/bash
# A/B builds: [FMAD=true] scripts/build_variant.sh NAME "-DSOME_MACRO ..." -> opencl_pathtracer_b200/_lib_NAME/ (select with PTC_LIB_DIR)
set -e
NAME=$1; shift
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/opencl_pathtracer_b200/_lib_$NAME
mkdir -p $OUT
cp $ROOT/opencl_pathtracer_b200/_lib/libpathtracer_host.so $ROOT/opencl_pathtracer_b200/_lib/libpathtracer_scene.so $OUT/
OBJS=""
FMAD=${FMAD:-false}
for f in pt_kernels pt_wide_build pt_lbvh pathtracer_cuda pt_multi pt_static; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=$FMAD -std=c++17 "$@" -Xcompiler -fPIC,-fopenmp,-ffp-contract=off,-fvisibility=hidden \
       -I $ROOT/include -I $ROOT/opencl_pathtracer_b200/csrc -c $ROOT/opencl_pathtracer_b200/csrc/$f.cu -o $OUT/$f.o &
  OBJS="$OBJS $OUT/$f.o"
done
wait
nvcc -shared -o $OUT/libpathtracer_cuda.so $OBJS -Xcompiler -fopenmp -lgomp -ldl 2>/dev/null
rm -f $OBJS
ls -la $OUT
