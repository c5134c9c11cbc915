#!/bin/bash
# r2w: forced wide-tree kernels on the scenes that use the lock-step kernels by default (A/B at HEAD)
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
for w in shiny_teapot scattered_balls cornell_smoke teapot_scene; do
  S=""; [ $w = cornell_smoke ] && S="--spp 512"
  $B --workload $w $S > $O/r2w_${w}_default.json 2>/dev/null
  HRT_FORCE_WIDE_PRIMARY=1 $B --workload $w $S > $O/r2w_${w}_wprim.json 2>/dev/null
  HRT_FORCE_WIDE_SECONDARY=1 $B --workload $w $S > $O/r2w_${w}_wsec.json 2>/dev/null
  HRT_FORCE_WIDE_PRIMARY=1 HRT_FORCE_WIDE_SECONDARY=1 $B --workload $w $S > $O/r2w_${w}_wboth.json 2>/dev/null
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2w_*.json")):
    try:
        d = json.load(open(f)); print(f, round(d["value"]), "ms", round(d["ms_per_step"], 3), {k: round(v, 1) for k, v in d.get("kernel_ms", {}).items() if k != "note"})
    except Exception as e: print(f, "FAILED", e)
PY
