#!/bin/bash
# A/B builds of the CUDA library for tuning experiments: tools/build_variant.sh NAME "-DHRT_KZ_MODE=1 ..."
# -> hobbyraytracer_b200/csrc/variants/libhrt_b200_NAME.so (select with HRT_PRODUCT_SO=...; not the shipped product)
set -e
NAME=$1; DEFS=$2
HERE=$(cd "$(dirname "$0")/../hobbyraytracer_b200/csrc" && pwd)
OUT=$HERE/variants; mkdir -p $OUT/obj_$NAME
NVFLAGS="-gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math,-pthread $DEFS"
for f in hrt_kernels.cu hrt_api.cu hrt_host.cpp hrt_builder.cpp; do
  /usr/local/cuda/bin/nvcc $NVFLAGS -x cu -c $HERE/$f -o $OUT/obj_$NAME/${f%.*}.o &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libhrt_b200_$NAME.so $OUT/obj_$NAME/*.o -lcudart_static -lpthread -ldl -lrt
echo built $OUT/libhrt_b200_$NAME.so
