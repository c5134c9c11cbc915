#!/bin/bash
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2j_pytest.log 2>&1; tail -15 $O/r2j_pytest.log
python bench.py --no-cpu --e2e-steps 1 --steps 2 --warmup 1 --workload teapot_scene > $O/r2j_teapot_scene.json 2> $O/r2j_teapot_scene.err; python -c "
import json; d=json.load(open('$O/r2j_teapot_scene.json')); print('teapot_scene', round(d['value']), d['kernel_ms'])"
