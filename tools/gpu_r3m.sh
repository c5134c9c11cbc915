#!/bin/bash
# r3m: final state: gpu suite, default bench (with its CPU arm), and the ncu launch list of the same command
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q -rs > $O/r3m_pytest_gpu.log 2>&1; tail -5 $O/r3m_pytest_gpu.log
python bench.py --steps 3 --warmup 3 > $O/r3m_bench.json 2> $O/r3m_bench.err; tail -c 200 $O/r3m_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $O/r3m_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > $O/ncu_launches.log 2>&1
python tools/launch_share.py $O/r3m_launches.csv | head -8
python -c "
import json; d=json.load(open('gpurun_out/r3m_bench.json')); print('bench', round(d['value']), 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],3), 'launches', d['gpu_launches'], d['kernel_ms'])"
