#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1"
$B --spp 128 --steps 5 --warmup 3 > $O/r2t_c2_spp128.json 2>/dev/null
HRT_DEBUG_ITER=1 $B --spp 128 --steps 1 --warmup 1 > /dev/null 2> $O/r2t_c2_spp128_iter.log
HRT_DEBUG_ITER=1 $B --workload teapot_scene --spp 13 --steps 1 --warmup 1 > /dev/null 2> $O/r2t_c1_spp13_iter.log
$B --workload teapot_scene --spp 13 --steps 5 --warmup 3 > $O/r2t_c1_spp13.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2t_*.json")):
    d = json.load(open(f)); print(f, round(d["value"]), "ms", round(d["ms_per_step"], 3), "launches", d["gpu_launches"], {k: round(v, 1) for k, v in d.get("kernel_ms", {}).items() if k != "note"})
PY
grep -c pass $O/r2t_c2_spp128_iter.log
