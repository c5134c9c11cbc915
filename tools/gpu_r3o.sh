#!/bin/bash
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "other_configs or full_resolution" --durations=5 2>&1 | tail -12
