"""The N > 1 path without GPUs: hobbyraytracer_b200.multigpu.render_split on CPU tensors over gloo (world_size 2 and 3).

The CUDA library is replaced by tests/hostsim (the product's device headers compiled for the CPU, exporting the same
film entry points on host memory), so what runs here is the real driver logic: the contiguous sample split, the integer
SUM reduce of the films to rank 0, the resolve on rank 0.  The film adds integers, so the reduced film must equal the
single-process film of all samples bit for bit — the property the 1/2/4/8-GPU runs rely on.
"""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

WORKER = r'''
import os, sys
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
import numpy as np, torch, torch.distributed as dist
from hobbyraytracer_b200 import abi, scenes, multigpu
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
lib = abi.Lib(os.path.join({root!r}, "tests", "hostsim", "_build", "libhrt_hostsim.so"), "hsim_")
lib._sig("film_bytes", abi.C.c_int, [abi.C.c_int, abi.C.c_int])
lib._sig("render_accumulate_device", abi.C.c_int, [abi.C.c_void_p, abi.C.POINTER(abi.Camera), abi.C.POINTER(abi.RenderParams), abi.C.c_void_p, abi.C.c_void_p])
lib._sig("film_resolve_device", abi.C.c_int, [abi.C.c_void_p, abi.C.c_void_p, abi.C.c_int, abi.C.c_int, abi.C.c_int, abi.C.c_void_p, abi.C.c_void_p, abi.C.c_void_p, abi.C.c_void_p])
sd, cam, _ = scenes.teapot_scene(os.path.join({root!r}, "tests", "_cache"), n=12, env_w=64, env_h=32)
w, h, spp = 40, 28, 7                      # 7 samples over 2 or 3 ranks: uneven shares
c = scenes.camera_for(cam, w, h)
sc = sd.build(lib, abi.BVH_SAH, 3)
film = multigpu.new_film(lib, w, h, device="cpu")
rgb = torch.zeros(w * h * 3, dtype=torch.uint8); tot = torch.zeros(w * h * 3, dtype=torch.float32)
multigpu.render_split(sc, c, w, h, spp, film, rank, world, dist, seed=11, rgb8=rgb, total=tot)
shares = [multigpu.sample_range(spp, r, world) for r in range(world)]
assert sum(n for _, n in shares) == spp and all(shares[i][0] + shares[i][1] == shares[i + 1][0] for i in range(world - 1))
if rank == 0:
    whole = multigpu.new_film(lib, w, h, device="cpu")
    rgb1 = torch.zeros_like(rgb); tot1 = torch.zeros_like(tot)
    multigpu.render_split(sc, c, w, h, spp, whole, 0, 1, None, seed=11, rgb8=rgb1, total=tot1)
    assert torch.equal(film, whole), "reduced film differs from the single-process film"
    assert torch.equal(rgb, rgb1) and torch.equal(tot.view(torch.int32), tot1.view(torch.int32))
    assert float(tot.sum()) > 0
    print("OK", world)
dist.barrier(); dist.destroy_process_group()
'''


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


@pytest.mark.parametrize("world", [2, 3])
def test_render_split_over_gloo(world, tmp_path):
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "tests", "hostsim")], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    script = tmp_path / "worker.py"; script.write_text(WORKER.format(root=ROOT))
    port = _free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=600)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert f"OK {world}" in outs[0]


def test_sample_range_partitions():
    from hobbyraytracer_b200 import multigpu
    for spp in (1, 7, 100, 1024, 4096):
        for world in (1, 2, 3, 4, 8):
            r = [multigpu.sample_range(spp, k, world) for k in range(world)]
            assert r[0][0] == 0 and sum(n for _, n in r) == spp
            assert all(r[i][0] + r[i][1] == r[i + 1][0] for i in range(world - 1))
            assert max(n for _, n in r) - min(n for _, n in r) <= 1
