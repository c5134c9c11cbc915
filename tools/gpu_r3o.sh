#!/bin/bash
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_resolution or chunked" --durations=3 2>&1 | tail -8
