#!/bin/bash
O=gpurun_out; mkdir -p $O
HRT_DEBUG_BUILD=1 python tools/build_probe.py 1000 lbvh_gpu soup 3 2>&1 | tee $O/r2q_build_probe.log
HRT_DEBUG_BUILD=1 python tools/build_probe.py 2290 lbvh_gpu indexed 3 2>&1 | tee -a $O/r2q_build_probe.log
python -m pytest tests/test_gpu_lbvh.py -m gpu -x -q 2>&1 | tail -5
python tools/ab_raybatch.py hobbyraytracer_b200/csrc/variants/libhrt_b200_r2m.so hobbyraytracer_b200/csrc/libhrt_b200.so 2>&1 | tee $O/r2q_ab_raybatch.log
B="python bench.py --no-cpu"
$B --workload teapot_scene --steps 3 --warmup 2 > $O/r2q_teapot_scene.json 2>/dev/null
$B --workload raybatch --steps 3 --warmup 2 > $O/r2q_raybatch.json 2>$O/r2q_raybatch.err
$B --workload big_mesh --bvh lbvh_gpu --indexed --steps 3 --warmup 2 > $O/r2q_big_mesh_lbvh.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2q_*.json")):
    try:
        d = json.load(open(f)); print(f, round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"], 2), "launches", d["gpu_launches"], {k: round(v, 1) for k, v in d.get("kernel_ms", {}).items() if k != "note"})
    except Exception as e:
        print(f, "FAILED", e)
PY
tail -3 $O/r2q_raybatch.err
