"""Samples-per-pixel split across the GPUs of one box: one process per GPU, the scene replicated,
rank r renders a contiguous range of global sample indices of every pixel into its own device accumulation buffer,
one NCCL reduce (sum, fp32, W*H*3) brings the film to rank 0, which resolves it (mean -> ACES -> sqrt -> 8 bit).

Samples are keyed by (pixel, global sample index) in Philox, so the image does not depend on the number of GPUs up
to fp32 summation order.  The reference has no counterpart: its render() is one process (main.cpp:81-140).
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


def render_split(scene: "abi.Scene", cam: "abi.Camera", width: int, height: int, spp: int, accum, rank: int = 0, world: int = 1,
                 dist=None, seed: int = 1, max_depth: int = 50, t_min: float = 0.001, rng_mode: int = 0, flags: int = 0,
                 stream_ptr: Optional[int] = None, rgb8=None, mean=None) -> None:
    """Render this rank's share into `accum` (a zeroed float32 device tensor of width*height*3 elements), reduce to
    rank 0 and, there, resolve into `rgb8` (uint8 device tensor) / `mean` when given.  `dist` is torch.distributed
    (initialised) or None for a single process.  All work is queued on the stream `stream_ptr` (cudaStream_t)."""
    lib = scene.lib
    first, count = sample_range(spp, rank, world)
    if count > 0:
        p = abi.RenderParams(width, height, spp, first, count, max_depth, t_min, seed, rng_mode, flags)
        lib.check(lib.fn("render_accumulate_device")(scene.h, C.byref(cam), C.byref(p), accum.data_ptr(), stream_ptr), "render_accumulate_device")
    if dist is not None and world > 1:
        dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0 and (rgb8 is not None or mean is not None):
        lib.check(lib.fn("film_resolve_device")(scene.h, accum.data_ptr(), width, height, spp,
                                                None if mean is None else mean.data_ptr(),
                                                None if rgb8 is None else rgb8.data_ptr(), stream_ptr), "film_resolve_device")
