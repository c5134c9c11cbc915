#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
for o in 8 12; do
  HRT_WIDE_PHASE_MIN_OBJ=$o $B --workload raybatch > $O/r3i_rb_o$o.json 2>/dev/null
  HRT_WIDE_PHASE_MIN_OBJ=$o $B --workload big_mesh --bvh lbvh_gpu --indexed > $O/r3i_c5_o$o.json 2>/dev/null
  HRT_WIDE_PHASE_MIN_OBJ=$o $B --workload teapot_scene > $O/r3i_c1_o$o.json 2>/dev/null
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3i_*.json")):
    try: print(f, round(json.load(open(f))["value"]))
    except Exception as e: print(f, "FAIL")
PY
