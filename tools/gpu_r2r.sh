#!/bin/bash
O=gpurun_out; mkdir -p $O
HRT_DEBUG_BUILD=1 python -m pytest tests -m gpu -x -q -s > $O/r2r_pytest_s.log 2>&1; tail -4 $O/r2r_pytest_s.log; grep "device build of shape" $O/r2r_pytest_s.log | awk '$9 > 500000' | head -20
