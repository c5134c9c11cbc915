#!/bin/bash
# r2d: first B200 run of the wide wavefront (k_wave3_trace): gpu tests, all five configs at reduced spp, old pipeline beside it
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2d_pytest.log 2>&1; tail -5 $O/r2d_pytest.log
B="python bench.py --no-cpu --e2e-steps 1 --steps 2 --warmup 1"
for w in "teapot_scene:0" "shiny_teapot:256" "scattered_balls:128" "cornell_smoke:256"; do
  n=${w%%:*}; spp=${w#*:}
  $B --workload $n --spp $spp > $O/r2d_$n.json 2> $O/r2d_$n.err
  HRT_NO_WIDE=1 $B --workload $n --spp $spp > $O/r2d_${n}_old.json 2> $O/r2d_${n}_old.err
  python - <<PY
import json
for f in ("$O/r2d_$n.json", "$O/r2d_${n}_old.json"):
    try:
        d = json.load(open(f)); print(f, round(d["value"]), "Mrays/s", round(d["ms_per_step"], 1), "ms", d["kernel_ms"]["trace"], d["kernel_ms"]["shade"], "replayed", d["replayed_fraction"], "refwalk", d["ref_walk_rays"])
    except Exception as e:
        print(f, "FAILED", e); print(open(f.replace(".json", ".err")).read()[-2000:])
PY
done
