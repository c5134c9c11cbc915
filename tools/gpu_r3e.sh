#!/bin/bash
# r3e: per-scene phase / refill thresholds (run-time): sweep on the scenes without meshes, check on the others
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
for pr in "8 24" "16 24" "16 16" "20 20" "24 24" "12 16" "20 28"; do
  set -- $pr
  HRT_WIDE_PHASE_MIN=$1 HRT_WIDE_REFILL=$2 $B --workload cornell_smoke --spp 256 > $O/r3e_c4_p$1_r$2.json 2>/dev/null
  HRT_WIDE_PHASE_MIN=$1 HRT_WIDE_REFILL=$2 HRT_FORCE_WIDE_SECONDARY=1 $B --workload scattered_balls > $O/r3e_c3w_p$1_r$2.json 2>/dev/null
done
$B --workload scattered_balls > $O/r3e_c3_default.json 2>/dev/null
$B --workload teapot_scene > $O/r3e_c1_default.json 2>/dev/null
$B --workload raybatch > $O/r3e_rb_default.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3e_*.json")):
    try: print(f, round(json.load(open(f))["value"]))
    except Exception as e: print(f, "FAIL")
PY
