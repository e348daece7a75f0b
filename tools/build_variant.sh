#!/bin/bash
# Build an experimental variant of librt.so with extra nvcc flags:  tools/build_variant.sh NAME "-DRT_X=1 ..."
# Output: raytracer_b200/csrc/variants/librt_NAME.so (git-ignored; select it with RT_LIBRT_PATH).
set -e
NAME=$1; shift
cd "$(dirname "$0")/../raytracer_b200/csrc"
mkdir -p variants/$NAME
for f in rt_api rt_build rt_frame rt_restir rt_raster rt_noalloc; do
  FM="-fmad=false"
  if [ $f = rt_restir ]; then FM="--use_fast_math"; fi   # as the Makefile builds it
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo $FM -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v "$@" -c $f.cu -o variants/$NAME/$f.o 2> variants/$NAME/$f.ptxas.log &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o variants/librt_$NAME.so variants/$NAME/*.o -lcudart_static -lpthread -ldl -lrt
grep -A2 "wf_extend_kernelILb0" variants/$NAME/rt_frame.ptxas.log | grep Used | sed "s/^/$NAME extend<0>: /"
