#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1"
$B --spp 128 --steps 5 --warmup 3 > $O/r2u_c2_spp128.json 2>/dev/null
$B --workload teapot_scene --spp 13 --steps 5 --warmup 3 > $O/r2u_c1_spp13.json 2>/dev/null
$B --workload teapot_scene --steps 3 --warmup 2 > $O/r2u_teapot_scene.json 2>/dev/null
$B --workload raybatch --steps 3 --warmup 2 > $O/r2u_raybatch.json 2>/dev/null
$B --workload big_mesh --bvh lbvh_gpu --indexed --steps 3 --warmup 2 > $O/r2u_big_mesh_lbvh.json 2>/dev/null
$B --workload cornell_smoke --spp 512 --steps 2 --warmup 2 > $O/r2u_cornell_smoke512.json 2>/dev/null
HRT_DEBUG_ITER=1 $B --workload teapot_scene --spp 13 --steps 1 --warmup 1 2>&1 >/dev/null | awk '/pass   0/{n++} n==2' | sed -n '36,62p'
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2u_*.json")):
    d = json.load(open(f)); print(f, round(d["value"]), "ms", round(d["ms_per_step"], 3), "launches", d["gpu_launches"], {k: round(v, 1) for k, v in d.get("kernel_ms", {}).items() if k != "note"})
PY
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
