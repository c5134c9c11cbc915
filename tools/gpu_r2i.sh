#!/bin/bash
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2i_pytest.log 2>&1; tail -5 $O/r2i_pytest.log
B="python bench.py --no-cpu --e2e-steps 1 --steps 2 --warmup 1"
$B --workload teapot_scene > $O/r2i_teapot_scene.json 2> $O/r2i_teapot_scene.err; python -c "
import json; d=json.load(open('$O/r2i_teapot_scene.json')); print('teapot_scene', round(d['value']), d['kernel_ms'])"
python tools/parity_bigmesh.py 2290 1048576 2>&1 | tee $O/r2i_parity_bigmesh.log
