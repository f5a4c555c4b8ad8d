#!/bin/bash
# usage: tools/build_variant.sh NAME -DHOGB200_HIST_WARPS=6 ...   -> gpurun_out/variants/libhogb200_NAME.so
# (experiments only; select with HOGB200_LIB=<path> -- gpurun_out/ itself is not shipped, so variants live in build_variants/)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build_variants/$name
for f in resize.cu hist.cu svm.cu group.cu hog_host.cu hog_storage.cu tbd_host.cpp group_ms.cpp; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-O2 "$@" -c opencv_b200/csrc/$f -o build_variants/$name/${f%.*}.o &
done
wait
/usr/local/cuda/bin/nvcc -shared -o build_variants/libhogb200_$name.so build_variants/$name/*.o -Xcompiler -fPIC -cudart static
echo build_variants/libhogb200_$name.so
