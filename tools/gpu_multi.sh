#!/bin/bash
# multi-GPU set: bash tools/gpu_multi.sh N TAG  (under gpurun --gpus N).  Every BASELINE config + the ray batch at N GPUs, one JSON line each;
# at N = 2 also the 2-GPU tests (peer reduce == NCCL reduce, the C++ host with --gpus 2).
N=$1; TAG=$2; O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8
if [ "$N" = "2" ]; then
  python -m pytest tests/test_multigpu_peer.py tests/test_cpp_host.py -m gpu -x -q > $O/${TAG}_pytest_2gpu.log 2>&1; tail -4 $O/${TAG}_pytest_2gpu.log
fi
P=29540
run() { name=$1; shift; P=$((P+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N "$@" > $O/${TAG}_${name}_${N}gpu.json 2> $O/${TAG}_${name}_${N}gpu.err || tail -c 400 $O/${TAG}_${name}_${N}gpu.err; }
run shiny_teapot --steps 3 --warmup 3
run teapot_scene --workload teapot_scene --no-cpu --steps 3 --warmup 3
[ -z "$SKIP_C3" ] && run scattered_balls --workload scattered_balls --no-cpu --steps 3 --warmup 3
run cornell_smoke --workload cornell_smoke --no-cpu --steps 2 --warmup 3 --e2e-steps 1
run big_mesh_lbvh --workload big_mesh --bvh lbvh_gpu --indexed --no-cpu --steps 3 --warmup 3
run raybatch --workload raybatch --bvh lbvh_gpu --no-cpu --steps 3 --warmup 3
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/${TAG}_*_${N}gpu.json")):
    try:
        d = json.load(open(f)); print(f, round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"], 2), "parity", d.get("reduce_parity"), d["config"].get("film_reduce", "")[:50])
    except Exception as e:
        print(f, "FAILED", e)
PY
