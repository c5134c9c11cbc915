"""2-GPU test (skipped on a 1-GPU box): the in-kernel reduces over NVLink peer memory - rank 0 pulling everything
(hrt_film_resolve_peers), every rank resolving its pixel slice into the root's buffers (hrt_film_resolve_slice, multigpu.PeerGroup) and
the single-process hrt_render_multi - give the same bits as the NCCL integer reduce and as a single-GPU render of all samples."""
import os
import socket
import subprocess
import sys

import pytest

from conftest import ROOT

WORKER = r'''
import os, sys
sys.path.insert(0, {root!r})
import numpy as np, torch, torch.distributed as dist
import hobbyraytracer_b200 as pkg
from hobbyraytracer_b200 import abi, scenes, multigpu
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
lib = pkg.load(); lib.check(lib.fn("set_device")(rank), "set_device")
sd, cam, _ = scenes.teapot_scene(os.path.join({root!r}, "tests", "_cache"), n=12, env_w=64, env_h=32)
w, h, spp = 96, 64, 13
c = scenes.camera_for(cam, w, h)
sc = sd.build(lib, abi.BVH_SAH, 3)
film = multigpu.new_film(lib, w, h)
rgb_n = torch.zeros(w * h * 3, dtype=torch.uint8, device="cuda"); tot_n = torch.zeros(w * h * 3, dtype=torch.float32, device="cuda")
multigpu.render_split(sc, c, w, h, spp, film, rank, world, dist, seed=11, rgb8=rgb_n, total=tot_n)
torch.cuda.synchronize()
pf = multigpu.PeerFilm(sc, w, h, rank, world, dist)
assert pf.ok, "CUDA IPC peer mapping is not available between these two GPUs"
rgb_p = torch.zeros_like(rgb_n); tot_p = torch.zeros_like(tot_n)
for _ in range(2):                                   # twice: the film is re-zeroed and the opened handles are reused
    multigpu.render_split_peer(pf, c, spp, seed=11, rgb8=rgb_p, total=tot_p)
torch.cuda.synchronize()
# the reduction spread over the ranks (PeerGroup / hrt_film_resolve_slice), everything queued on one stream per rank
work = torch.cuda.Stream()
with torch.cuda.stream(work):
    pg = multigpu.PeerGroup(sc, w, h, rank, world, dist)
    assert pg.ok
    for _ in range(3):
        pg.render(c, spp, seed=11)
    work.synchronize()
    if rank == 0:
        gs, gb = pg.root_views()
        assert torch.equal(gb.reshape(-1), rgb_n) and torch.equal(gs.reshape(-1).view(torch.int32), tot_n.view(torch.int32)), "slice reduce differs from the NCCL reduce"
    sc2 = sd.build(lib, abi.BVH_SAH, 3)              # a host that rebuilds its scene per frame keeps the group
    pg.render(c, spp, seed=11, scene=sc2); work.synchronize()
    if rank == 0:
        gs, gb = pg.root_views()
        assert torch.equal(gb.reshape(-1), rgb_n)
dist.barrier(); pg.close()
if rank == 0:
    assert torch.equal(rgb_n, rgb_p) and torch.equal(tot_n.view(torch.int32), tot_p.view(torch.int32)), "peer reduce differs from the NCCL reduce"
    S, B = sc.render(c, w, h, spp, seed=11)
    assert np.array_equal(S.reshape(-1).view(np.uint32), tot_p.cpu().numpy().view(np.uint32)) and np.array_equal(B.reshape(-1), rgb_p.cpu().numpy())
    assert float(tot_p.sum()) > 0
    # one process, two GPUs (hrt_render_multi): the same image again
    S2, B2 = sc.render_multi([0, 1], c, w, h, spp, seed=11)
    assert np.array_equal(S2.view(np.uint32), S.view(np.uint32)) and np.array_equal(B2, B)
    g = sd.build(lib, abi.BVH_LBVH_GPU, 3)           # device-built meshes are copied to the replica over NVLink
    S3, B3 = g.render(c, w, h, spp, seed=11); S4, B4 = g.render_multi([0, 1], c, w, h, spp, seed=11)
    assert np.array_equal(S3.view(np.uint32), S4.view(np.uint32)) and np.array_equal(B3, B4)
    print("OK")
dist.barrier(); pf.close(); dist.destroy_process_group()
'''


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


@pytest.mark.gpu
def test_peer_reduce_equals_nccl_reduce(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = tmp_path / "worker.py"; script.write_text(WORKER.format(root=ROOT))
    port = _free_port(); procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=300)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert "OK" in outs[0]
