#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
for b in 8 4; do
  HRT_TRACE_BUILD=$b $B --workload scattered_balls > $O/r3q_c3_b$b.json 2>/dev/null
  HRT_TRACE_BUILD=$b $B --workload cornell_smoke --spp 256 > $O/r3q_c4_b$b.json 2>/dev/null
  HRT_TRACE_BUILD=$b $B --workload teapot_scene > $O/r3q_c1_b$b.json 2>/dev/null
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3q_*.json")):
    try: d=json.load(open(f)); print(f, round(d["value"]), round(d["ms_per_step"],2), {k: round(v, 1) for k, v in d.get("kernel_ms", {}).items() if k != "note"})
    except Exception as e: print(f, "FAIL")
PY
