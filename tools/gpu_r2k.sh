#!/bin/bash
# r2k (2 GPUs): multi-GPU tests, bench at N = 2 (peer group + reduce parity + pipelined e2e), bench at N = 1
O=gpurun_out; mkdir -p $O
nvidia-smi -L
python -m pytest tests -m gpu -x -q > $O/r2k_pytest.log 2>&1; tail -8 $O/r2k_pytest.log
python bench.py --steps 3 --warmup 3 > $O/r2k_bench_1gpu.json 2> $O/r2k_bench_1gpu.err; tail -c 300 $O/r2k_bench_1gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 3 --warmup 3 > $O/r2k_bench_2gpu.json 2> $O/r2k_bench_2gpu.err; tail -c 600 $O/r2k_bench_2gpu.err
python - <<'PY'
import json
for f in ("gpurun_out/r2k_bench_1gpu.json", "gpurun_out/r2k_bench_2gpu.json"):
    try:
        d = json.load(open(f)); print(f, "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"], 2), "parity", d.get("reduce_parity"), d["config"]["film_reduce"][:60])
    except Exception as e:
        print(f, "FAILED", e)
PY
