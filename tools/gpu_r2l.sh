#!/bin/bash
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2l_pytest.log 2>&1; tail -12 $O/r2l_pytest.log
bash tools/gpu_sanitize.sh memcheck
