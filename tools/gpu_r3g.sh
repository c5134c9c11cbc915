#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
for pr in "8 8" "8 12" "8 16" "8 24" "12 16" "6 12"; do
  set -- $pr
  HRT_WIDE_PHASE_MIN=$1 HRT_WIDE_PHASE_MIN_OBJ=$2 $B --workload teapot_scene > $O/r3g_c1_p$1_o$2.json 2>/dev/null
done
for pr in "16 16" "8 16" "12 20" "16 24"; do
  set -- $pr
  HRT_WIDE_PHASE_MIN=$1 HRT_WIDE_PHASE_MIN_OBJ=$2 $B --workload cornell_smoke --spp 256 > $O/r3g_c4_p$1_o$2.json 2>/dev/null
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3g_*.json")):
    try: print(f, round(json.load(open(f))["value"]))
    except Exception as e: print(f, "FAIL")
PY
