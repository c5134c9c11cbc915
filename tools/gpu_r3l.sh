#!/bin/bash
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r3l_pytest_gpu.log 2>&1; tail -3 $O/r3l_pytest_gpu.log
python bench.py --no-cpu --workload raybatch --steps 3 --warmup 2 > $O/r3l_raybatch.json 2> $O/r3l_raybatch.err; tail -c 300 $O/r3l_raybatch.err
python -c "
import json; d=json.load(open('gpurun_out/r3l_raybatch.json')); print('raybatch', round(d['value']), 'e2e', round(d['e2e']['value']), d['parity'])"
