#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 2 --warmup 1"
show() { python - "$1" <<'PY'
import json, sys
f = sys.argv[1]
try:
    d = json.load(open(f)); print(f, round(d["value"]), "Mrays/s", round(d["ms_per_step"], 1), "ms trace", round(d["kernel_ms"]["trace"], 1), "shade", round(d["kernel_ms"]["shade"], 1), "rays/sample", round(d["rays_per_sample"], 2))
except Exception as e:
    print(f, "FAILED", e); print(open(f.replace(".json", ".err")).read()[-1500:])
PY
}
$B --workload big_mesh --bvh sah --spp 16 > $O/r2g_big_mesh.json 2> $O/r2g_big_mesh.err; show $O/r2g_big_mesh.json
HRT_WAVE3_MERGED=1 $B --workload big_mesh --bvh sah --spp 16 > $O/r2g_big_mesh_merged.json 2> $O/r2g_big_mesh_merged.err; show $O/r2g_big_mesh_merged.json
HRT_NO_WIDE=1 $B --workload big_mesh --bvh sah --spp 16 > $O/r2g_big_mesh_old.json 2> $O/r2g_big_mesh_old.err; show $O/r2g_big_mesh_old.json
$B --workload teapot_scene > $O/r2g_teapot_scene.json 2> $O/r2g_teapot_scene.err; show $O/r2g_teapot_scene.json
$B --workload shiny_teapot --spp 256 > $O/r2g_shiny_teapot.json 2> $O/r2g_shiny_teapot.err; show $O/r2g_shiny_teapot.json
$B --workload scattered_balls --spp 128 > $O/r2g_scattered_balls.json 2> $O/r2g_scattered_balls.err; show $O/r2g_scattered_balls.json
NCU="ncu --set full --clock-control none --import-source on"
C1="python bench.py --no-cpu --e2e-steps 1 --workload teapot_scene --spp 16 --steps 1 --warmup 1"
$C1 > $O/plain_c1.log 2>&1 && $NCU -k regex:'k_wide_secondary' -s 8 -c 2 -o $O/r2g_c1 $C1 > $O/ncu_c1.log 2>&1; tail -1 $O/ncu_c1.log
