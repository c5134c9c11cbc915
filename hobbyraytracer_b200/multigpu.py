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
