#!/bin/bash
O=gpurun_out; mkdir -p $O
V=hobbyraytracer_b200/csrc/variants
python tools/ab_raybatch.py hobbyraytracer_b200/csrc/libhrt_b200.so $V/libhrt_b200_p8b4.so $V/libhrt_b200_p12b4.so $V/libhrt_b200_p16b4.so $V/libhrt_b200_p12b5.so $V/libhrt_b200_p12b6.so $V/libhrt_b200_p12r28.so $V/libhrt_b200_p16r30.so 2>&1 | tee $O/r2c_ab.log
HRT_DEBUG_BUILD=1 python -m pytest tests/test_gpu_lbvh.py -x -q -k two_million -s 2>&1 | tail -15 | tee $O/r2c_two_million.log
