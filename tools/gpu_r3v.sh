#!/bin/bash
# r3v: every config once more at the final kernels (after r3t), short runs
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 2 --warmup 2"
$B --workload teapot_scene > $O/r3v_teapot_scene.json 2>/dev/null
$B --workload cornell_smoke --warmup 1 > $O/r3v_cornell_smoke.json 2>/dev/null
$B --workload big_mesh --bvh lbvh_gpu --indexed > $O/r3v_big_mesh_lbvh.json 2>/dev/null
$B --workload raybatch > $O/r3v_raybatch.json 2>/dev/null
$B --workload scattered_balls > $O/r3v_scattered_balls.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3v_*.json")):
    try: d=json.load(open(f)); print(f, round(d["value"]), "e2e", round(d["e2e"]["value"]), round(d["ms_per_step"],2), round(d["roofline"]["frac"],3))
    except Exception as e: print(f, "FAIL")
PY
