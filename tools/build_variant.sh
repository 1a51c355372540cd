#!/bin/bash
# Build a tuning variant of libvofront.so into variants/NAME/ (git-ignored, travels with gpurun):
#   tools/build_variant.sh NAME [-DVO_CTAS=24 ...]
set -e
cd "$(dirname "$0")/.."
NAME=$1; shift
OUT=variants/$NAME; mkdir -p $OUT
CS=acs_visual_odometry_b200/csrc
for f in api stencil nms_select freak match stages; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC "$@" -c $CS/$f.cu -o $OUT/$f.o &
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $OUT/libvofront.so $OUT/*.o
rm -f $OUT/*.o
echo $OUT/libvofront.so
