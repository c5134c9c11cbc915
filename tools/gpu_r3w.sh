#!/bin/bash
# r3w: wide kernels in two builds (with / without the fifth phase), picked per scene: gpu suite + the mesh-only configs
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r3w_pytest_gpu.log 2>&1; tail -3 $O/r3w_pytest_gpu.log
B="python bench.py --no-cpu --e2e-steps 1 --steps 2 --warmup 1"
$B --workload big_mesh --bvh lbvh_gpu --indexed > $O/r3w_big_mesh_lbvh.json 2>/dev/null
$B --workload raybatch --bvh lbvh_gpu > $O/r3w_raybatch.json 2>/dev/null
$B --workload teapot_scene > $O/r3w_teapot_scene.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3w_*.json")):
    try: d=json.load(open(f)); print(f, round(d["value"]), round(d["ms_per_step"],2))
    except Exception as e: print(f, "FAIL")
PY
