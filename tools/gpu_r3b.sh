#!/bin/bash
# r3b: the sanitizer case against the product and against the bounds-checked variant (compute-sanitizer is closed on this pool)
O=gpurun_out; mkdir -p $O
python tools/sanitize_case.py > $O/r3b_sanitize_case_product.log 2>&1; tail -2 $O/r3b_sanitize_case_product.log
HRT_PRODUCT_SO=hobbyraytracer_b200/csrc/variants/libhrt_b200_bounds.so python tools/sanitize_case.py > $O/r3b_sanitize_case_bounds.log 2>&1; tail -4 $O/r3b_sanitize_case_bounds.log
HRT_PRODUCT_SO=$PWD/hobbyraytracer_b200/csrc/variants/libhrt_b200_bounds.so python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_progressive.py -m gpu -x -q > $O/r3b_pytest_bounds.log 2>&1; tail -3 $O/r3b_pytest_bounds.log
