#!/bin/bash
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r3a_pytest_gpu.log 2>&1; tail -3 $O/r3a_pytest_gpu.log
B="python bench.py --no-cpu --steps 3 --warmup 2"
$B --workload teapot_scene > $O/r3a_teapot_scene.json 2>/dev/null
$B --workload cornell_smoke --spp 512 --steps 2 > $O/r3a_cornell_smoke512.json 2>/dev/null
$B --workload big_mesh --bvh lbvh_gpu --indexed > $O/r3a_big_mesh_lbvh.json 2>/dev/null
$B --workload raybatch > $O/r3a_raybatch.json 2>/dev/null
$B > $O/r3a_shiny_teapot.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3a_*.json")):
    try:
        d = json.load(open(f)); print(f, round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms", round(d["ms_per_step"], 2))
    except Exception as e: print(f, "FAILED", e)
PY
