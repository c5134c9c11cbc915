#!/bin/bash
# r3h: state at HEAD: whole gpu suite, smoke, default bench with the CPU arm, reference arm, every config
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r3h_pytest_gpu.log 2>&1; tail -3 $O/r3h_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r3h_smoke.log 2>&1; tail -1 $O/r3h_smoke.log
python bench.py --steps 3 --warmup 3 > $O/r3h_bench.json 2> $O/r3h_bench.err; tail -c 200 $O/r3h_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/r3h_bench_reference.json 2> $O/r3h_bench_reference.err; tail -c 200 $O/r3h_bench_reference.err
B="python bench.py --no-cpu --steps 3 --warmup 3"
$B --workload teapot_scene > $O/r3h_teapot_scene.json 2>/dev/null
$B --workload scattered_balls > $O/r3h_scattered_balls.json 2>/dev/null
$B --workload cornell_smoke --steps 2 > $O/r3h_cornell_smoke.json 2>/dev/null
$B --workload big_mesh --bvh lbvh_gpu --indexed > $O/r3h_big_mesh_lbvh.json 2>/dev/null
$B --workload raybatch > $O/r3h_raybatch.json 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3h_*.json")):
    try:
        d = json.load(open(f)); print(f, round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms", round(d["ms_per_step"], 2), "frac", round(d.get("roofline", {}).get("frac") or 0, 3))
    except Exception as e: print(f, "FAILED", e)
PY
