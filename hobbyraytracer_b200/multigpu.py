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
        if not stream_ptr and film.is_cuda:
            # the reduce is ordered on torch's current stream, the resolve below goes to the scene's own stream (NULL): wait here
            import torch
            torch.cuda.current_stream().synchronize()
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


class PeerGroup:
    """The reduction spread over the ranks (include/hrt.h, hrt_film_resolve_slice).  Every rank owns a film and — the root's are the
    ones that get written — output buffers, all of them library allocations that the other processes map through CUDA IPC handles
    exchanged ONCE here (one all_gather).  A step is then queued entirely on one CUDA stream per rank, without host-side waits on
    other ranks:
        zero own film -> render own samples -> barrier -> sum + resolve OWN pixel slice from all films into the root's buffers -> barrier
    where a barrier is a one-word NCCL all_reduce on the same stream (stream ordered: the host never blocks on it).  `ok` is agreed
    on by all ranks; when False (no IPC / peer access) the caller keeps to render_split."""

    def __init__(self, scene: "abi.Scene", width: int, height: int, rank: int, world: int, dist):
        import torch
        self.lib, self.scene, self.w, self.h, self.rank, self.world, self.dist = scene.lib, scene, width, height, rank, world, dist
        self.npix = width * height
        lib = self.lib
        self.film = self.out = None; self.films = None; self.root_out = None; self.ok = False
        self.token = torch.zeros(1, dtype=torch.int32, device="cuda")
        good = 1 if all(lib.has(n) for n in ("peer_alloc", "peer_handle", "peer_open", "film_resolve_slice")) else 0
        mine = torch.zeros(2 * 64, dtype=torch.uint8)
        if good:
            self.film = lib.fn("peer_alloc")(lib.check(lib.fn("film_bytes")(width, height), "film_bytes"))
            self.out = lib.fn("peer_alloc")(self.npix * 15)                       # float sums (12 B / pixel) then the 8-bit image (3 B / pixel)
            if not self.film or not self.out:
                good = 0
            else:
                buf = (C.c_ubyte * 128)()
                if lib.fn("peer_handle")(self.film, buf) < 0 or lib.fn("peer_handle")(self.out, C.byref(buf, 64)) < 0:
                    good = 0
                else:
                    mine = torch.tensor(list(buf), dtype=torch.uint8)
        gathered = [torch.zeros(128, dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(gathered, mine.cuda())
        if good:
            ptrs = []
            for r, g in enumerate(gathered):
                raw = bytes(g.cpu().tolist())
                if r == rank:
                    ptrs.append(self.film)
                    if r == 0:
                        self.root_out = self.out
                    continue
                hb = C.create_string_buffer(raw[:64], 64)
                p = lib.fn("peer_open")(scene.h, hb)
                if not p:
                    good = 0; break
                ptrs.append(p)
                if r == 0:
                    ob = C.create_string_buffer(raw[64:], 64)
                    self.root_out = lib.fn("peer_open")(scene.h, ob)
                    if not self.root_out:
                        good = 0; break
            if good:
                self.films = (C.c_void_p * world)(*ptrs)
        flag = torch.tensor([good], dtype=torch.int32, device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        self.ok = int(flag.item()) == 1

    def slice(self):
        base, extra = divmod(self.npix, self.world)
        first = self.rank * base + min(self.rank, extra)
        return first, base + (1 if self.rank < extra else 0)

    def barrier(self):
        self.dist.all_reduce(self.token)          # on torch's current stream: a stream-ordered barrier, the host does not wait

    def render(self, cam: "abi.Camera", spp: int, seed: int = 1, max_depth: int = 50, t_min: float = 0.001, rng_mode: int = 0, flags: int = 0,
               want_sum: bool = True, want_rgb8: bool = True, scene: Optional["abi.Scene"] = None) -> None:
        """One step; everything is queued on torch's CURRENT stream (use a non-default one: `with torch.cuda.stream(s):`).  Afterwards —
        in stream order — the root's buffers hold the image: see root_views().  `scene`: render another scene of this process (same
        device) than the one the mappings were opened with — a host that rebuilds its scene every frame keeps one PeerGroup."""
        import torch
        lib = self.lib
        stream_ptr = torch.cuda.current_stream().cuda_stream
        if not stream_ptr:
            raise abi.HrtError("PeerGroup.render needs a non-default CUDA stream (the NCCL barriers are ordered on it)")
        lib.check(lib.fn("film_zero")(self.film, self.w, self.h, stream_ptr), "film_zero")
        first, count = sample_range(spp, self.rank, self.world)
        if count > 0:
            p = abi.RenderParams(self.w, self.h, spp, first, count, max_depth, t_min, seed, rng_mode, flags)
            lib.check(lib.fn("render_accumulate_device")((scene or self.scene).h, C.byref(cam), C.byref(p), self.film, stream_ptr), "render_accumulate_device")
        self.barrier()                                                            # every film is complete
        pf, pc = self.slice()
        d_sum = self.root_out if want_sum else None
        d_rgb = (self.root_out + self.npix * 12) if want_rgb8 else None
        lib.check(lib.fn("film_resolve_slice")(self.scene.h, self.films, self.world, self.w, self.h, spp, pf, pc, d_sum, None, d_rgb, stream_ptr), "film_resolve_slice")
        self.barrier()                                                            # every slice has landed; the films may be zeroed again

    def root_views(self):
        """(float sums [h, w, 3], uint8 image [h, w, 3]) as torch tensors over the root's own buffers (rank 0 only)."""
        import torch
        assert self.rank == 0
        class _Arr:      # __cuda_array_interface__ views of library memory
            def __init__(self, ptr, shape, typestr):
                self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 3}
        s = torch.as_tensor(_Arr(self.out, (self.h, self.w, 3), "<f4"), device="cuda")
        b = torch.as_tensor(_Arr(self.out + self.npix * 12, (self.h, self.w, 3), "|u1"), device="cuda")
        return s, b

    def close(self):
        if self.film:
            self.lib.fn("peer_free")(self.film); self.film = None
        if self.out:
            self.lib.fn("peer_free")(self.out); self.out = None
