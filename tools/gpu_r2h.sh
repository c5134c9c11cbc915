#!/bin/bash
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2h_pytest.log 2>&1; tail -3 $O/r2h_pytest.log
B="python bench.py --no-cpu --e2e-steps 1 --steps 2 --warmup 1"
show() { python - "$1" <<'PY'
import json, sys
f = sys.argv[1]
try:
    d = json.load(open(f)); print(f, round(d["value"]), "Mrays/s e2e", round(d["e2e"]["value"]), " ", round(d["ms_per_step"], 1), "ms trace", round(d["kernel_ms"]["trace"], 1), "shade", round(d["kernel_ms"]["shade"], 1))
except Exception as e:
    print(f, "FAILED", e); print(open(f.replace(".json", ".err")).read()[-1500:])
PY
}
V=$PWD/hobbyraytracer_b200/csrc/variants
$B --workload big_mesh --bvh sah --spp 16 > $O/r2h_big_mesh.json 2> $O/r2h_big_mesh.err; show $O/r2h_big_mesh.json
HRT_PRODUCT_SO=$V/libhrt_b200_wt3.so $B --workload big_mesh --bvh sah --spp 16 > $O/r2h_big_mesh_wt3.json 2> $O/r2h_big_mesh_wt3.err; show $O/r2h_big_mesh_wt3.json
$B --workload big_mesh --bvh lbvh_gpu --indexed > $O/r2h_big_mesh_lbvh.json 2> $O/r2h_big_mesh_lbvh.err; show $O/r2h_big_mesh_lbvh.json
$B --workload teapot_scene > $O/r2h_teapot_scene.json 2> $O/r2h_teapot_scene.err; show $O/r2h_teapot_scene.json
$B --workload raybatch --bvh lbvh_gpu > $O/r2h_raybatch_lbvh.json 2> $O/r2h_raybatch_lbvh.err; python -c "
import json; d=json.load(open('$O/r2h_raybatch_lbvh.json')); print('raybatch lbvh_gpu', round(d['value']), d['parity'], 'commit', d['commit_s'])"
NCU="ncu --set full --clock-control none --import-source on"
C1="python bench.py --no-cpu --e2e-steps 1 --workload teapot_scene --spp 100 --steps 1 --warmup 1"
$C1 > $O/plain_c1.log 2>&1 && $NCU -k regex:'k_wide_secondary' -s 20 -c 1 -o $O/r2h_c1 $C1 > $O/ncu_c1.log 2>&1; tail -1 $O/ncu_c1.log
