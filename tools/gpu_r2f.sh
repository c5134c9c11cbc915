#!/bin/bash
# r2f: hybrid wide pipeline (phase kernel for secondary rays + lock-step camera paths): tests, configs, register variants
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2f_pytest.log 2>&1; tail -3 $O/r2f_pytest.log
V=hobbyraytracer_b200/csrc/variants
python tools/ab_raybatch.py hobbyraytracer_b200/csrc/libhrt_b200.so $V/libhrt_b200_b5.so $V/libhrt_b200_b6.so $V/libhrt_b200_b8.so 2>&1 | tee $O/r2f_ab_raybatch.log
B="python bench.py --no-cpu --e2e-steps 1 --steps 2 --warmup 1"
show() { python - "$1" <<'PY'
import json, sys
f = sys.argv[1]
try:
    d = json.load(open(f)); print(f, round(d["value"]), "Mrays/s", round(d["ms_per_step"], 1), "ms trace", round(d["kernel_ms"]["trace"], 1), "shade", round(d["kernel_ms"]["shade"], 1), "replayed", d["replayed_fraction"])
except Exception as e:
    print(f, "FAILED", e); print(open(f.replace(".json", ".err")).read()[-1500:])
PY
}
for w in "teapot_scene:0" "shiny_teapot:256" "scattered_balls:128" "cornell_smoke:256"; do
  n=${w%%:*}; spp=${w#*:}
  for lib in product b5 b6 b8; do
    if [ $lib = product ]; then unset HRT_PRODUCT_SO; else export HRT_PRODUCT_SO=$PWD/$V/libhrt_b200_$lib.so; fi
    $B --workload $n --spp $spp > $O/r2f_${n}_$lib.json 2> $O/r2f_${n}_$lib.err; show $O/r2f_${n}_$lib.json
  done
  unset HRT_PRODUCT_SO
  HRT_NO_WIDE=1 $B --workload $n --spp $spp > $O/r2f_${n}_old.json 2> $O/r2f_${n}_old.err; show $O/r2f_${n}_old.json
done
$B --workload big_mesh --bvh sah --spp 16 --steps 1 > $O/r2f_big_mesh.json 2> $O/r2f_big_mesh.err; show $O/r2f_big_mesh.json
$B --workload big_mesh --bvh lbvh_gpu --indexed --spp 16 --steps 1 > $O/r2f_big_mesh_lbvh.json 2> $O/r2f_big_mesh_lbvh.err; show $O/r2f_big_mesh_lbvh.json
