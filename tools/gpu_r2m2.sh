#!/bin/bash
O=gpurun_out; mkdir -p $O
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/r2m_pytest_noalloc.log 2>&1; tail -4 $O/r2m_pytest_noalloc.log
python -m pytest tests -m gpu -x -q > $O/r2m_pytest.log 2>&1; tail -5 $O/r2m_pytest.log
