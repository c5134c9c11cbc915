#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
for m in 4 3 5; do
  HRT_TRACE_MINB=$m $B > $O/r3n_c2_minb$m.json 2>/dev/null
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3n_*.json")):
    try: d=json.load(open(f)); print(f, round(d["value"]), round(d["ms_per_step"],2))
    except Exception as e: print(f, "FAIL")
PY
