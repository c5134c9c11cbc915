#!/bin/bash
# A/B builds of the CUDA library for tuning experiments: tools/build_variant.sh NAME "-DHRT_WIDE_PHASE_MIN=12 ..."
# -> hobbyraytracer_b200/csrc/variants/libhrt_b200_NAME.so (select with HRT_PRODUCT_SO=...; not the shipped product).
# Only hrt_kernels.cu is recompiled with the extra flags; the other objects are the product's (run `make` in csrc first).
set -e
NAME=$1; DEFS=$2
HERE=$(cd "$(dirname "$0")/../hobbyraytracer_b200/csrc" && pwd)
OUT=$HERE/variants; mkdir -p $OUT
NVFLAGS="-gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math,-pthread -Xptxas -v $DEFS"
/usr/local/cuda/bin/nvcc $NVFLAGS -c $HERE/hrt_kernels.cu -o $OUT/hrt_kernels_$NAME.o 2> $OUT/$NAME.ptxas.log
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libhrt_b200_$NAME.so $OUT/hrt_kernels_$NAME.o $HERE/hrt_bvh_gpu.o $HERE/hrt_api.o $HERE/hrt_host.o $HERE/hrt_builder.o -lcudart_static -lpthread -ldl -lrt
rm -f $OUT/hrt_kernels_$NAME.o
echo built $OUT/libhrt_b200_$NAME.so
