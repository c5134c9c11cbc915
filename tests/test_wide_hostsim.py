"""The 8-wide quantised trees (hrt_bvh_rules.h (4), hrt_wide.cuh) on the CPU: the same headers the CUDA kernels are built from,
compiled with g++ (tests/hostsim), against the oracle — hit primitive ids and t bit for bit on every scene family of
test_hostsim_vs_oracle.py — plus structural checks of the collapse (containment of the quantised boxes, numbering)."""
import ctypes as C

import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes
from test_hostsim_vs_oracle import OPS, bits, hsim, mesh_scene, nan_rays, rects_scene, spheres_scene  # noqa: F401

F = np.float32


def wide_trace(lib, sc, rays, seed=0):
    lib._sig("trace_batch_wide", C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p])
    rays = np.ascontiguousarray(rays, F).reshape(-1, 8)
    n = rays.shape[0]
    prim = np.empty(n, np.uint32); t = np.empty(n, F); steps = C.c_uint64(0)
    lib.check(lib.fn("trace_batch_wide")(sc.h, rays.ctypes.data, n, prim.ctypes.data, t.ctypes.data, seed, C.byref(steps)), "trace_batch_wide")
    return dict(prim=prim, t=t, steps=int(steps.value))


def same(o, w):
    assert np.array_equal(o["prim"], w["prim"]), f"prim ids differ at {np.nonzero(o['prim'] != w['prim'])[0][:10]}"
    assert np.array_equal(bits(o["t"]), bits(w["t"]))


def test_spheres_rects_wide(orc, hsim):
    for sd, center, rad, seed in ((spheres_scene(), (0, 0, -1), 3.0, 1), (rects_scene(), (0, 2.5, 0), 2.4, 2)):
        o = sd.build(orc, abi.BVH_REFERENCE, 5); h = sd.build(hsim, abi.BVH_SAH, 5)
        rays = scenes.random_rays(30000, center, rad, seed=seed)
        rays[::7, 3] = 0.5; rays[::5, 7] = 2.0; rays[::11, 3] = -1.0
        rays[:100, 4:7] = [0, 0, -1]; rays[100:200, 4:7] = [1, 0, 0]
        same(o.trace(rays, extras=False), wide_trace(hsim, h, rays))


@pytest.mark.parametrize("mode", [abi.BVH_SAH, abi.BVH_LBVH_HOST])
@pytest.mark.parametrize("n", [8, 24, 56])
def test_mesh_wide(orc, hsim, mode, n):
    sd = mesh_scene(n)
    o = sd.build(orc, abi.BVH_REFERENCE, 9); h = sd.build(hsim, mode, 9)
    rays = scenes.random_rays(60000, (0, 1, 0), 2.5, seed=4)
    rays[:6000, 4:7] *= 4.1; rays[6000:9000, 4:7] *= 0.03
    # HRT_BVH_LBVH_* order equal keys of the reference topology stably where the oracle's std::sort permutes them (include/hrt.h):
    # bottom-node boxes, hence sub-t_min self hits, can differ from the oracle there; the yardstick is that mode's two-child walk
    truth = (lambda r: o.trace(r, extras=False)) if mode == abi.BVH_SAH else (lambda r: h.trace(r, extras=False))
    ro = truth(rays); rw = wide_trace(hsim, h, rays)
    assert (ro["prim"] != abi.MISS).mean() > 0.3
    same(ro, rw)
    hit = ro["prim"] != abi.MISS                       # secondary rays from the surface (sub-t_min self hits)
    p = rays[hit, 0:3] + ro["t"][hit, None] * rays[hit, 4:7]
    rng = np.random.default_rng(0)
    sec = np.empty((p.shape[0], 8), F); sec[:, 0:3] = p; sec[:, 3] = 1e-3; sec[:, 7] = np.inf
    d = rng.normal(size=(p.shape[0], 3)); sec[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    same(truth(sec), wide_trace(hsim, h, sec))


def test_mesh_degenerate_wide(orc, hsim):
    for n_tri in (0, 1, 2, 3, 5):
        sd = abi.SceneDesc(); m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
        pos, nrm, uv = scenes.blob_mesh(2, 2, 1.0, 0.0, seed=1)
        sd.add(sd.mesh(pos[:n_tri], nrm[:n_tri], uv[:n_tri], m))
        sd.add(sd.sphere((0, 0, 0), 0.3, m))
        o = sd.build(orc, 0, 5); h = sd.build(hsim, abi.BVH_SAH, 5)
        rays = scenes.random_rays(3000, (0, 0, 0), 2.0, seed=n_tri)
        same(o.trace(rays, extras=False), wide_trace(hsim, h, rays))
    # a mesh alone in the world (the top-level root is a leaf), axis-parallel rays (zero direction components: trace_ref decides)
    sd = mesh_scene(8, with_floor=False)
    o = sd.build(orc, 0, 5); h = sd.build(hsim, abi.BVH_SAH, 5)
    rays = np.array([[0, 1, -3, 1e-3, 0, 0, 1, np.inf], [3, 1, 0, 1e-3, -1, 0, 0, np.inf],
                     [0, 4, 0, 1e-3, 0, -1, 0, np.inf], [0.3, 1.2, -3, 1e-3, 0, 0, 1, np.inf]], F)
    same(o.trace(rays, extras=False), wide_trace(hsim, h, rays))
    rays = scenes.random_rays(20000, (0, 1, 0), 2.5, seed=12)
    same(o.trace(rays, extras=False), wide_trace(hsim, h, rays))


@pytest.mark.parametrize("ops", OPS)
def test_instances_wide(orc, hsim, ops):
    for kind in ("mesh", "sphere", "box"):
        sd = abi.SceneDesc()
        m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
        sd.add(sd.rect(1, -4, 4, -4, 4, -0.5, m))
        if kind == "mesh":
            pos, nrm, uv = scenes.blob_mesh(12, 12, 1.0, 0.2, seed=3, center=(0, 0.5, 0)); sh = sd.mesh(pos, nrm, uv, m)
        elif kind == "sphere":
            sh = sd.sphere((0.2, 0.5, -0.1), 0.8, m)
        else:
            sh = sd.box((0, 0.5, 0), (1.0, 1.5, 0.8), m)
        sd.add(sh, ops)
        sd.add(sd.sphere((2.5, 0.5, 0.5), 0.7, m))
        o = sd.build(orc, 0, 5); h = sd.build(hsim, abi.BVH_SAH, 5)
        rays = scenes.random_rays(20000, (0, 0.5, 0), 3.5, seed=5)
        rays[:5000, 4:7] *= 3.7; rays[5000:8000, 4:7] *= 0.05
        same(o.trace(rays, extras=False), wide_trace(hsim, h, rays))


def test_many_spheres_wide(orc, hsim):
    sd, cam, _ = scenes.scattered_balls_scene(seed=42, grid=11)
    o = sd.build(orc, 0, 5); h = sd.build(hsim, abi.BVH_SAH, 5)
    rays = np.concatenate([scenes.random_rays(20000, (0, 1, 0), 12.0, seed=11), scenes.camera_rays(cam, 1200, 800, 20000, seed=2)])
    rw = wide_trace(hsim, h, rays)
    same(o.trace(rays, extras=False), rw)
    assert rw["steps"] / len(rays) < 40, "the wide top-level tree should cut the ~485 sphere tests per ray of the flat list"


def test_media_and_nan_wide(orc, hsim):
    # the product's two traversal orders draw the same Philox number per (ray, medium): wide == two-child walk exactly
    sd, cam, _ = scenes.cornell_smoke_scene(glass_sphere=True)
    h = sd.build(hsim, abi.BVH_SAH, 5)
    rays = scenes.random_rays(40000, (278, 278, 278), 270, seed=6); rays[:10000, 4:7] *= 2.5
    same(h.trace(rays, rng_seed=77, extras=False), wide_trace(hsim, h, rays, seed=77))
    cases = [spheres_scene(), rects_scene(), mesh_scene(8), mesh_scene(8, ops=OPS[4]), mesh_scene(8, ops=OPS[3], with_floor=False)]
    for sd in cases:
        o = sd.build(orc, 0, 5); h = sd.build(hsim, abi.BVH_SAH, 5)
        rays = nan_rays()
        same(o.trace(rays, rng_seed=3, extras=False), wide_trace(hsim, h, rays, seed=3))


def test_teapot_scene_wide(orc, hsim, cache_dir):
    """C1's geometry: the mesh under rotate / scale / translate inside the Cornell walls, camera rays and diffuse bounces."""
    sd, cam, _ = scenes.teapot_scene(cache_dir, n=56, env_w=64, env_h=32)
    o = sd.build(orc, 0, 3); h = sd.build(hsim, abi.BVH_SAH, 3)
    rays = np.concatenate([scenes.camera_rays(cam, 640, 640, 30000, seed=3), scenes.random_rays(30000, (0, 2.5, 0), 2.4, seed=9)])
    ro = o.trace(rays, extras=False); rw = wide_trace(hsim, h, rays)
    same(ro, rw)
    two = h.trace(rays, extras=False)
    same(ro, two)


@pytest.mark.parametrize("mode", [abi.BVH_SAH, abi.BVH_LBVH_HOST])
def test_wide_collapse_structure(hsim, mode):
    """hrt_bvh_rules.h (4): every leaf record is reachable exactly once, wide nodes are numbered breadth first (the children
    of node i come after i, in slot order), and the quantised box of a child contains everything below it."""
    sd = mesh_scene(40, with_floor=False)
    h = sd.build(hsim, mode, 3)
    leaves = h.mesh_array(0, 2).reshape(-1, 32)
    raw = h.mesh_array(0, 5).view(np.uint8).reshape(-1, 96)
    n = raw.shape[0]
    p = raw[:, 0:12].copy().view(F).reshape(n, 3).astype(np.float64)
    meta = raw[:, 12:16].copy().view(np.uint32).reshape(n)
    q = raw[:, 16:64].reshape(n, 6, 8).astype(np.float64)
    child = raw[:, 64:96].copy().view(np.int32).reshape(n, 8)
    step = np.stack([np.ldexp(1.0, ((meta >> (8 * a)) & 0xff).astype(np.int64) - 127) for a in range(3)], 1)      # n x 3
    valid = (meta >> 24) & 0xff
    lo = p[:, :, None] + q[:, 0::2, :] * step[:, :, None]          # n x 3 x 8
    hi = p[:, :, None] + q[:, 1::2, :] * step[:, :, None]
    seen = np.zeros(leaves.shape[0], np.int64)
    nxt = 1
    sub_lo = np.full((n, 3), np.inf); sub_hi = np.full((n, 3), -np.inf)       # true bounds of everything below node i
    order = []
    for i in range(n):
        for s in range(8):
            if not (valid[i] >> s) & 1:
                continue
            c = child[i, s]
            if c >= 0:
                assert c == nxt, "breadth-first numbering in slot order"
                nxt += 1
            else:
                seen[~c] += 1
            order.append((i, s, c))
    assert nxt == n and (seen == 1).all()
    for i, s, c in reversed(order):                                            # children have larger indices: one backwards sweep
        if c >= 0:
            blo, bhi = sub_lo[c], sub_hi[c]
        else:
            blo, bhi = leaves[~c, 0:3].astype(np.float64), leaves[~c, 4:7].astype(np.float64)
        assert (lo[i, :, s] <= blo).all() and (hi[i, :, s] >= bhi).all(), (i, s, c)
        sub_lo[i] = np.minimum(sub_lo[i], blo); sub_hi[i] = np.maximum(sub_hi[i], bhi)
        assert (hi[i, :, s] - lo[i, :, s] <= (bhi - blo) + 2.0 * step[i] + 1e-30).all(), "quantisation adds at most one grid step per side"


def test_visit_counts_by_kind_add_up(hsim):
    """hsim_wide_visit_counts (the figures behind tools/b_logical.py): node steps + triangle leaves + mesh entries + other entries = the steps
    trace_batch_wide reports; a scene with one mesh enters it at most once per ray."""
    sc = mesh_scene(24).build(hsim, abi.BVH_SAH, 9)
    rays = np.ascontiguousarray(scenes.random_rays(20000, (0, 1, 0), 2.5, seed=4), F)
    f = hsim.dll.hsim_wide_visit_counts
    f.restype = C.c_int; f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_uint64, C.c_void_p]
    out = (C.c_uint64 * 5)()
    assert f(sc.h, rays.ctypes.data, len(rays), 0, out) == 0
    node, leaf, mesh, entry, pend = list(out)
    assert node + leaf + mesh + entry == wide_trace(hsim, sc, rays)["steps"]
    assert 0 < mesh <= len(rays) and node > leaf > 0 and pend < len(rays) // 10
