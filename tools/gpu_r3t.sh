#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
$B --workload cornell_smoke --spp 256 > $O/r3t_c4.json 2>/dev/null
$B --workload teapot_scene > $O/r3t_c1.json 2>/dev/null
$B --workload raybatch > $O/r3t_rb.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3t_c*.json")+glob.glob("gpurun_out/r3t_rb.json")):
    try: d=json.load(open(f)); print(f, round(d["value"]), round(d["ms_per_step"],2))
    except Exception as e: print(f, "FAIL")
PY
python -m pytest tests/test_gpu_parity.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -2
