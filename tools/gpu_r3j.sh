#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
$B --workload raybatch > $O/r3j_rb.json 2>/dev/null
$B --workload big_mesh --bvh lbvh_gpu --indexed > $O/r3j_c5.json 2>/dev/null
$B --workload teapot_scene > $O/r3j_c1.json 2>/dev/null
$B --workload cornell_smoke --spp 256 > $O/r3j_c4.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3j_*.json")):
    try: print(f, round(json.load(open(f))["value"]))
    except Exception as e: print(f, "FAIL")
PY
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
