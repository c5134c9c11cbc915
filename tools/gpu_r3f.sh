#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
for pr in "16 16" "16 12" "16 8" "12 12" "24 16" "32 16"; do
  set -- $pr
  HRT_WIDE_PHASE_MIN=$1 HRT_WIDE_REFILL=$2 $B --workload cornell_smoke --spp 256 > $O/r3f_c4_p$1_r$2.json 2>/dev/null
done
$B --workload cornell_smoke --spp 256 > $O/r3f_c4_default.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3f_*.json")):
    try: print(f, round(json.load(open(f))["value"]))
    except Exception as e: print(f, "FAIL")
PY
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
