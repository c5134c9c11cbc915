#!/bin/bash
# r2m (1 GPU): the allocation-free entry alone in a fresh process, the whole gpu suite, smoke, the default bench + every config
O=gpurun_out; mkdir -p $O
nvidia-smi -L
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/r2m_pytest_noalloc.log 2>&1; tail -4 $O/r2m_pytest_noalloc.log
python -m pytest tests -m gpu -x -q --durations=15 > $O/r2m_pytest.log 2>&1; tail -25 $O/r2m_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2m_smoke.log 2>&1; tail -2 $O/r2m_smoke.log
python bench.py --steps 3 --warmup 3 > $O/r2m_bench.json 2> $O/r2m_bench.err; tail -c 300 $O/r2m_bench.err
for w in teapot_scene scattered_balls cornell_smoke raybatch; do
  python bench.py --workload $w --no-cpu --steps 3 --warmup 3 > $O/r2m_$w.json 2> $O/r2m_$w.err; tail -c 200 $O/r2m_$w.err
done
python bench.py --workload big_mesh --bvh lbvh_gpu --indexed --no-cpu --steps 3 --warmup 3 > $O/r2m_big_mesh_lbvh.json 2> $O/r2m_big_mesh_lbvh.err; tail -c 200 $O/r2m_big_mesh_lbvh.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2m_*.json")):
    try:
        d = json.load(open(f)); print(f, d["config"]["workload"][:50], "value", round(d["value"]), "e2e", round(d["e2e"]["value"]) if d.get("e2e") else None, "ms", round(d["ms_per_step"], 2), "frac", d["roofline"]["frac"])
    except Exception as e:
        print(f, "FAILED", e)
PY
