"""Samples-per-pixel split across the GPUs of one box: one process per GPU, the scene replicated,
rank r renders a contiguous range of global sample indices of every pixel into its own device film (int64 fixed-point
sample sums, include/hrt.h), one NCCL reduce (integer sum, W*H*4 int64) brings the film to rank 0, which resolves it (mean -> ACES -> sqrt -> 8 bit).

Samples are keyed by (pixel, global sample index) in Philox, and the film adds integers, so the image does not depend on
the number of GPUs at all (bit-identical).  The reference has no counterpart: its render() is one process (main.cpp:81-140).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

from . import abi


def sample_range(spp: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced split of sample indices [0, spp) over `world` ranks: (first, count)."""
    base, extra = divmod(spp, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def new_film(lib: "abi.Lib", width: int, height: int, device="cuda"):
    """A zeroed film (int64 fixed-point sample sums, include/hrt.h) as a torch tensor on `device`."""
    import torch
    nbytes = lib.check(lib.fn("film_bytes")(width, height), "film_bytes")
    return torch.zeros(nbytes // 8, dtype=torch.int64, device=device)


def render_split(scene: "abi.Scene", cam: "abi.Camera", width: int, height: int, spp: int, film, rank: int = 0, world: int = 1,
                 dist=None, seed: int = 1, max_depth: int = 50, t_min: float = 0.001, rng_mode: int = 0, flags: int = 0,
                 stream_ptr: Optional[int] = None, rgb8=None, mean=None, total=None) -> None:
    """Render this rank's share of the samples into `film` (a zeroed int64 device tensor from new_film), reduce the
    films to rank 0 (integer SUM: exact and order independent, so the image is bit-identical for any number of GPUs)
    and, there, resolve into `rgb8` (uint8) / `mean` / `total` (float32 sums) when given.  `dist` is torch.distributed
    (initialised) or None for a single process.  All CUDA work is queued on the stream `stream_ptr` (cudaStream_t)."""
    lib = scene.lib
    first, count = sample_range(spp, rank, world)
    if count > 0:
        p = abi.RenderParams(width, height, spp, first, count, max_depth, t_min, seed, rng_mode, flags)
        lib.check(lib.fn("render_accumulate_device")(scene.h, C.byref(cam), C.byref(p), film.data_ptr(), stream_ptr), "render_accumulate_device")
    if dist is not None and world > 1:
        dist.reduce(film, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0 and (rgb8 is not None or mean is not None or total is not None):
        lib.check(lib.fn("film_resolve_device")(scene.h, film.data_ptr(), width, height, spp,
                                                None if total is None else total.data_ptr(),
                                                None if mean is None else mean.data_ptr(),
                                                None if rgb8 is None else rgb8.data_ptr(), stream_ptr), "film_resolve_device")


class PeerFilm:
    """This rank's film as a library allocation that the other ranks can map (CUDA IPC), plus — on rank 0 — the handles of
    all peers.  With it the reduce is not a collective at all: rank 0's resolve kernel pulls the peers' lanes over NVLink
    and sums them on the fly (include/hrt.h, hrt_film_resolve_peers).  `ok` is agreed on by all ranks (a probe resolve on
    rank 0 + one all_reduce); when it is False the caller keeps using render_split (NCCL reduce)."""

    def __init__(self, scene: "abi.Scene", width: int, height: int, rank: int, world: int, dist):
        import torch
        self.lib, self.scene, self.w, self.h, self.rank, self.world, self.dist = scene.lib, scene, width, height, rank, world, dist
        self.ptr = None; self.handles = None; self.ok = False
        lib = self.lib
        good = 1
        if not (lib.has("film_create") and lib.has("film_resolve_peers")):
            good = 0
        else:
            self.ptr = lib.fn("film_create")(width, height)
            if not self.ptr:
                good = 0
        mine = torch.zeros(64, dtype=torch.uint8)
        if good:
            buf = (C.c_ubyte * 64)()
            if lib.fn("film_ipc_handle")(self.ptr, buf) < 0:
                good = 0
            else:
                mine = torch.tensor(list(buf), dtype=torch.uint8)
        gathered = [torch.zeros(64, dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(gathered, mine.cuda())
        if rank == 0 and good:
            others = b"".join(bytes(g.cpu().tolist()) for r, g in enumerate(gathered) if r != 0)
            self.handles = C.create_string_buffer(others, len(others))
        flag = torch.tensor([good], dtype=torch.int32, device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 1:
            # probe: can rank 0 really map and read every peer?  (all films are zero here)
            if rank == 0:
                rc = lib.fn("film_resolve_peers")(scene.h, self.ptr, self.handles, world - 1, width, height, 1, 0, None, None, None, None)
                torch.cuda.synchronize()
                flag[0] = 1 if rc >= 0 else 0
            dist.broadcast(flag, src=0)
        self.ok = int(flag.item()) == 1

    def close(self):
        if self.ptr:
            self.lib.fn("film_destroy")(self.ptr); self.ptr = None


def render_split_peer(pf: PeerFilm, cam: "abi.Camera", spp: int, seed: int = 1, max_depth: int = 50, t_min: float = 0.001,
                      rng_mode: int = 0, flags: int = 0, stream_ptr: Optional[int] = None, rgb8=None, mean=None, total=None) -> None:
    """render_split without a collective on the data path: zero + render this rank's samples into its IPC-visible film,
    barrier, rank 0 reduces and resolves in ONE kernel over NVLink peer memory, barrier (the films are free again)."""
    import torch
    lib, scene, dist = pf.lib, pf.scene, pf.dist
    lib.check(lib.fn("film_zero")(pf.ptr, pf.w, pf.h, stream_ptr), "film_zero")
    first, count = sample_range(spp, pf.rank, pf.world)
    if count > 0:
        p = abi.RenderParams(pf.w, pf.h, spp, first, count, max_depth, t_min, seed, rng_mode, flags)
        lib.check(lib.fn("render_accumulate_device")(scene.h, C.byref(cam), C.byref(p), pf.ptr, stream_ptr), "render_accumulate_device")
    torch.cuda.synchronize()
    dist.barrier()
    if pf.rank == 0:
        lib.check(lib.fn("film_resolve_peers")(scene.h, pf.ptr, pf.handles, pf.world - 1, pf.w, pf.h, spp, 0,
                                               None if total is None else total.data_ptr(), None if mean is None else mean.data_ptr(),
                                               None if rgb8 is None else rgb8.data_ptr(), stream_ptr), "film_resolve_peers")
        torch.cuda.synchronize()
    dist.barrier()
