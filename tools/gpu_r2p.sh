#!/bin/bash
# r2p (1 GPU): unified pop + mesh-entry phase + 32 Mi queues + pooled builder blocks: gpu suite, every config, capacity A/B on C2 / C3
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2p_pytest.log 2>&1; tail -5 $O/r2p_pytest.log
B="python bench.py --no-cpu"
$B --workload teapot_scene --steps 3 --warmup 2 > $O/r2p_teapot_scene.json 2>/dev/null
$B --workload raybatch --steps 3 --warmup 2 > $O/r2p_raybatch.json 2>/dev/null
$B --workload big_mesh --bvh lbvh_gpu --indexed --steps 3 --warmup 2 > $O/r2p_big_mesh_lbvh.json 2>/dev/null
$B --workload big_mesh --bvh sah --spp 16 --steps 2 --warmup 1 --e2e-steps 1 > $O/r2p_big_mesh_sah16.json 2>/dev/null
$B --workload big_mesh --bvh lbvh_gpu --indexed --spp 16 --steps 2 --warmup 1 --e2e-steps 1 > $O/r2p_big_mesh_lbvh16.json 2>/dev/null
$B --workload scattered_balls --steps 3 --warmup 2 > $O/r2p_scattered_balls.json 2>/dev/null
$B --workload cornell_smoke --spp 1024 --steps 2 --warmup 2 > $O/r2p_cornell_smoke1024.json 2>/dev/null
$B --steps 3 --warmup 3 > $O/r2p_shiny_teapot.json 2>/dev/null
for cap in 4194304 67108864; do
  HRT_WAVE_CAPACITY=$cap $B --steps 3 --warmup 3 --e2e-steps 1 > $O/r2p_c2_cap$cap.json 2>/dev/null
  HRT_WAVE_CAPACITY=$cap $B --workload scattered_balls --steps 3 --warmup 2 --e2e-steps 1 > $O/r2p_c3_cap$cap.json 2>/dev/null
  HRT_WAVE_CAPACITY=$cap $B --workload teapot_scene --steps 3 --warmup 2 --e2e-steps 1 > $O/r2p_c1_cap$cap.json 2>/dev/null
done
HRT_WAVE_CAPACITY=67108864 $B --workload big_mesh --bvh lbvh_gpu --indexed --steps 2 --warmup 2 --e2e-steps 1 > $O/r2p_c5_cap67108864.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2p_*.json")):
    try:
        d = json.load(open(f)); print(f, round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"], 2), "launches", d["gpu_launches"], {k: round(v, 1) for k, v in d.get("kernel_ms", {}).items() if k != "note"})
    except Exception as e:
        print(f, "FAILED", e)
PY
