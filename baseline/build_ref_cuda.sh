#!/bin/bash
# On-box COMPARATOR (never linked into libhogb200.so): the reference's own CUDA HOG kernels K0, K2-K9
# (modules/cudaobjdetect/src/cuda/hog.cu:101-821) compiled for sm_100a, so that they can be timed on the same B200
# beside ours.  SURVEY.md 2.3: hog.cu does not compile with CUDA 12 because of its texture-reference resize (K1,
# hog.cu:828-885); everything above line 824 does once opencv2/core/cuda/common.hpp's bindTexture helper (lines 99-103) is
# left out.  This script DERIVES the masked copies from /root/reference into the git-ignored baseline/_ref/ (no reference
# source enters the repository) and builds baseline/_ref/ref_cuda_bench from them and baseline/ref_cuda_driver.cu.
set -e
cd "$(dirname "$0")/.."
REF=${REF:-/root/reference}
OUT=baseline/_ref
[ -f "$REF/modules/cudaobjdetect/src/cuda/hog.cu" ] || { echo "no reference sources: keeping the prebuilt $OUT/ref_cuda_bench"; exit 0; }
[ -f oracle/_ref/libcvref.so ] || make -s -j8 -C oracle ref
mkdir -p $OUT/shadow/opencv2/core/cuda
head -824 "$REF/modules/cudaobjdetect/src/cuda/hog.cu" > $OUT/hog_masked.cu
printf '    } // namespace hog\n}}}\n#endif\n' >> $OUT/hog_masked.cu
sed '99,103d' "$REF/modules/core/include/opencv2/core/cuda/common.hpp" > $OUT/shadow/opencv2/core/cuda/common.hpp
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
INC="-I$OUT/shadow -Ioracle/refcfg -I$REF/modules/core/include -I$REF/modules/cudaobjdetect/include"
$NVCC -std=c++14 -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo $INC -DHAVE_CUDA -w -c $OUT/hog_masked.cu -o $OUT/hog_masked.o
$NVCC -std=c++14 -gencode arch=compute_100a,code=sm_100a -O3 $INC -DHAVE_CUDA -w -c baseline/ref_cuda_driver.cu -o $OUT/ref_cuda_driver.o
$NVCC -o $OUT/ref_cuda_bench $OUT/ref_cuda_driver.o $OUT/hog_masked.o -Loracle/_ref -lcvref -Xlinker -rpath -Xlinker '$ORIGIN/../../oracle/_ref' -cudart static
echo $OUT/ref_cuda_bench
