"""The product's device headers (hrt_hit.cuh / hrt_shade.cuh) compiled for the CPU (tests/hostsim) against the oracle.

This is the no-GPU rehearsal of tests/test_gpu_parity.py: the same source the CUDA kernels are built from runs here
with g++, so traversal logic (reference order, SAH order + tie re-trace), the host flattener and BVH builders, and the
shading switch are checked bit-for-bit (hits) / statistically (renders) in a container without a GPU.
"""
import os
import subprocess

import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes
from conftest import ROOT

F = np.float32
HSIM_SO = os.path.join(ROOT, "tests", "hostsim", "_build", "libhrt_hostsim.so")


@pytest.fixture(scope="session")
def hsim():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "tests", "hostsim")], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    lib = abi.Lib(HSIM_SO, "hsim_")
    lib._sig("counters", abi.C.c_int, [abi.C.c_void_p, abi.C.c_void_p])
    return lib


def bits(a):
    """Bit pattern of fp32 values, with every NaN mapped to one canonical pattern: the sign/payload of a generated NaN is a
    hardware convention (x86 SSE produces 0xFFC00000, NVIDIA GPUs 0x7FFFFFFF), not arithmetic — NaN *positions* must agree."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = a.view(np.uint32).copy()
    b[np.isnan(a)] = 0x7FC00000
    return b


def check_hits(o, h, uv_exact=True, front=True, t_exact=True, skip_uv=None):
    assert np.array_equal(o["prim"], h["prim"]), f"prim ids differ at {np.nonzero(o['prim'] != h['prim'])[0][:10]}"
    if t_exact:
        assert np.array_equal(bits(o["t"]), bits(h["t"]))
    assert np.array_equal(bits(o["normal"]), bits(h["normal"]))
    if skip_uv is not None:
        o["uv"][skip_uv] = 0; h["uv"][skip_uv] = 0
    if uv_exact:
        assert np.array_equal(bits(o["uv"]), bits(h["uv"]))
    if front:
        assert np.array_equal(o["front"], h["front"])


def fallbacks(lib, sc):
    c = np.zeros(3, np.uint64); lib.fn("counters")(sc.h, c.ctypes.data); return int(c[2]), int(c[0]), int(c[1])


def spheres_scene():
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.sphere((0, 0, -1), 0.5, m)); sd.add(sd.sphere((0, 0, -1), 0.5, m))
    sd.add(sd.sphere((1.5, 0.2, -2), 0.75, m)); sd.add(sd.sphere((0, -100.5, -1), 100, m))
    return sd


def rects_scene():
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.rect(0, 0, 5, -2.5, 2.5, 2.5, m)); sd.add(sd.rect(0, 0, 5, -2.5, 2.5, -2.5, m))
    sd.add(sd.rect(1, -2.5, 2.5, -2.5, 2.5, 0, m)); sd.add(sd.rect(1, -2.5, 2.5, -2.5, 2.5, 5, m))
    sd.add(sd.rect(2, -2.5, 2.5, 0, 5, -2.5, m))
    sd.add(sd.box((0, 1, 0), (1.5, 2, 1), m))
    sd.add(sd.box_minmax((-2, 0, -2), (-1, 0.7, -1.2), m))
    sd.add(sd.rect(2, -0.75, 0.75, 0, 2, 0.5, m))
    return sd


def mesh_scene(n=24, ops=(), with_floor=True, center=(0, 1, 0)):
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    pos, nrm, uv = scenes.blob_mesh(n, n, 1.0, 0.2, seed=7, center=center)
    if with_floor:
        sd.add(sd.rect(1, -3, 3, -3, 3, 0, m))
    sd.add(sd.mesh(pos, nrm, uv, m), list(ops))
    return sd


@pytest.mark.parametrize("mode", [abi.BVH_REFERENCE, abi.BVH_SAH])
def test_spheres_rects(orc, hsim, mode):
    for sd, center, rad, seed in ((spheres_scene(), (0, 0, -1), 3.0, 1), (rects_scene(), (0, 2.5, 0), 2.4, 2)):
        o = sd.build(orc, abi.BVH_REFERENCE, 5); h = sd.build(hsim, mode, 5)
        rays = scenes.random_rays(30000, center, rad, seed=seed)
        rays[::7, 3] = 0.5; rays[::5, 7] = 2.0
        rays[:100, 4:7] = [0, 0, -1]; rays[100:200, 4:7] = [1, 0, 0]
        # sphere u,v go through acos/atan2 of the same libm here, so they are exact on the CPU too
        check_hits(o.trace(rays), h.trace(rays))


@pytest.mark.parametrize("mode", [abi.BVH_REFERENCE, abi.BVH_SAH])
def test_mesh(orc, hsim, mode):
    sd = mesh_scene(24)
    o = sd.build(orc, abi.BVH_REFERENCE, 9); h = sd.build(hsim, mode, 9)
    # the host builder reproduces the oracle's (= the reference's) topology node for node
    nb = orc.fn("mesh_bvh_nodes")(o.h, 1, None, None)
    assert nb == hsim.fn("mesh_bvh_nodes")(h.h, 1, None, None)
    bo = np.zeros((nb, 6), F); co = np.zeros((nb, 2), np.int32); bh = bo.copy(); ch = co.copy()
    orc.fn("mesh_bvh_nodes")(o.h, 1, bo.ctypes.data, co.ctypes.data); hsim.fn("mesh_bvh_nodes")(h.h, 1, bh.ctypes.data, ch.ctypes.data)
    # the oracle numbers nodes in creation order (= pre-order), so ids agree as well
    assert np.array_equal(co, ch) and np.array_equal(bits(bo), bits(bh))
    rays = scenes.random_rays(60000, (0, 1, 0), 2.5, seed=4)
    ro, rh = o.trace(rays), h.trace(rays)
    assert (ro["prim"] != abi.MISS).mean() > 0.3
    check_hits(ro, rh, front=False)
    if mode == abi.BVH_SAH:
        fb, nodes, prims = fallbacks(hsim, h)
        assert fb < 0.01 * len(rays), f"{fb} rays re-traced in reference order"
    # secondary rays from the surface: sub-t_min self hits (ITriangle ignores t_min) must agree
    hit = ro["prim"] != abi.MISS
    p = rays[hit, 0:3] + ro["t"][hit, None] * rays[hit, 4:7]
    rng = np.random.default_rng(0)
    sec = np.empty((p.shape[0], 8), F); sec[:, 0:3] = p; sec[:, 3] = 1e-3; sec[:, 7] = np.inf
    d = rng.normal(size=(p.shape[0], 3)); sec[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    so, sh = o.trace(sec), h.trace(sec)
    check_hits(so, sh, front=False)
    assert ((so["prim"] != abi.MISS) & (so["t"] < 1e-3)).sum() > 0


@pytest.mark.parametrize("mode", [abi.BVH_REFERENCE, abi.BVH_SAH])
def test_mesh_degenerate(orc, hsim, mode):
    sd = mesh_scene(8, with_floor=False)
    o = sd.build(orc, 0, 5); h = sd.build(hsim, mode, 5)
    rays = np.array([[0, 1, -3, 1e-3, 0, 0, 1, np.inf], [3, 1, 0, 1e-3, -1, 0, 0, np.inf],
                     [0, 4, 0, 1e-3, 0, -1, 0, np.inf], [0.3, 1.2, -3, 1e-3, 0, 0, 1, np.inf]], F)
    ro, rh = o.trace(rays), h.trace(rays)
    assert np.array_equal(ro["prim"], rh["prim"]) and np.array_equal(bits(ro["t"]), bits(rh["t"]))
    # empty mesh, single triangle, two triangles
    for n_tri in (0, 1, 2, 3):
        sd = abi.SceneDesc(); m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
        pos, nrm, uv = scenes.blob_mesh(2, 2, 1.0, 0.0, seed=1)
        sd.add(sd.mesh(pos[:n_tri], nrm[:n_tri], uv[:n_tri], m))
        sd.add(sd.sphere((0, 0, 0), 0.3, m))
        o = sd.build(orc, 0, 5); h = sd.build(hsim, mode, 5)
        rays = scenes.random_rays(3000, (0, 0, 0), 2.0, seed=n_tri)
        check_hits(o.trace(rays), h.trace(rays), front=False)


OPS = [
    [(abi.XF_TRANSLATE, (0.5, -0.25, 1.0))],
    [(abi.XF_SCALE, (1.5, 0.5, 2.0))],
    [(abi.XF_ROTATE_EULER_DEG, (10, 180, -35))],
    [(abi.XF_ROTATE_Y_DEG, (15,))],
    [(abi.XF_ROTATE_EULER_DEG, (0, 180, 0)), (abi.XF_SCALE, (1.4, 1.4, 1.4)), (abi.XF_TRANSLATE, (0, 1, 0))],
    [(abi.XF_ROTATE_Y_DEG, (-18,)), (abi.XF_TRANSLATE, (1.3, 0, 0.65))],
    [(abi.XF_SCALE, (0.5, 0.5, 0.5)), (abi.XF_ROTATE_EULER_DEG, (20, 30, 40)), (abi.XF_SCALE, (3, 2, 2.5)), (abi.XF_TRANSLATE, (0.1, 0.2, 0.3))],
]


@pytest.mark.parametrize("mode", [abi.BVH_REFERENCE, abi.BVH_SAH])
@pytest.mark.parametrize("ops", OPS)
def test_instances(orc, hsim, mode, ops):
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
        o = sd.build(orc, 0, 5); h = sd.build(hsim, mode, 5)
        rays = scenes.random_rays(20000, (0, 0.5, 0), 3.5, seed=5)
        rays[:5000, 4:7] *= 3.7; rays[5000:8000, 4:7] *= 0.05
        check_hits(o.trace(rays), h.trace(rays))


@pytest.mark.parametrize("mode", [abi.BVH_REFERENCE, abi.BVH_SAH])
def test_world_list_front_flag(orc, hsim, mode):
    """A bare mesh hit reports the frontFace left by the latest earlier successful hit (tempRec reuse)."""
    sd = abi.SceneDesc(); m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.sphere((0, 1, 0), 2.5, m))
    pos, nrm, uv = scenes.blob_mesh(10, 10, 1.0, 0.2, seed=3, center=(0, 1, 0)); sd.add(sd.mesh(pos, nrm, uv, m))
    o = sd.build(orc, 0, 5); h = sd.build(hsim, mode, 5)
    rays = scenes.random_rays(20000, (0, 1, 0), 4.0, seed=8)
    ro, rh = o.trace(rays), h.trace(rays)
    check_hits(ro, rh, front=(mode == abi.BVH_REFERENCE))


def test_many_spheres_tlas(orc, hsim):
    sd, cam, _ = scenes.scattered_balls_scene(seed=42, grid=11)
    o = sd.build(orc, 0, 5); h = sd.build(hsim, abi.BVH_SAH, 5)
    rays = np.concatenate([scenes.random_rays(20000, (0, 1, 0), 12.0, seed=11),
                           scenes.camera_rays(cam, 1200, 800, 20000, seed=2)])
    check_hits(o.trace(rays), h.trace(rays))
    fb, nodes, prims = fallbacks(hsim, h)
    assert fb < 0.01 * len(rays)
    assert prims / len(rays) < 20, "top-level tree should cut the ~485 sphere tests per ray of the flat list"


def test_media_boundaries(orc, hsim):
    """Media consume a random number inside hit(): the Philox stream differs from the oracle's rand() stream, so
    only rays whose result does not involve a medium acceptance test are compared bit-for-bit; medium hit rates
    are compared statistically."""
    sd, cam, _ = scenes.cornell_smoke_scene(glass_sphere=True)
    o = sd.build(orc, 0, 5)
    rays = scenes.random_rays(40000, (278, 278, 278), 270, seed=6); rays[:10000, 4:7] *= 2.5
    ro = o.trace(rays, rng_seed=77)
    media = [o.prim_base(8), o.prim_base(9)]
    res = {}
    for mode in (abi.BVH_REFERENCE, abi.BVH_SAH):
        h = sd.build(hsim, mode, 5); rh = h.trace(rays, rng_seed=77); res[mode] = rh
        for mp in media:
            fo, fh = (ro["prim"] == mp).mean(), (rh["prim"] == mp).mean()
            assert abs(fo - fh) < 0.01 + 0.1 * fo, (mp, fo, fh)
            to, th = ro["t"][ro["prim"] == mp].mean(), rh["t"][rh["prim"] == mp].mean()
            assert abs(to - th) < 0.05 * to
    # the two traversal orders of the product agree with each other exactly (same Philox draw per (ray, medium))
    a, b = res[abi.BVH_REFERENCE], res[abi.BVH_SAH]
    assert np.array_equal(a["prim"], b["prim"]) and np.array_equal(bits(a["t"]), bits(b["t"]))
    # without media the scene must agree with the oracle exactly
    sd2, _, _ = scenes.cornell_smoke_scene(glass_sphere=True)
    sd2.objects = [ob for ob in sd2.objects if ob[0] == "object"]
    o2 = sd2.build(orc, 0, 5)
    for mode in (abi.BVH_REFERENCE, abi.BVH_SAH):
        check_hits(o2.trace(rays), sd2.build(hsim, mode, 5).trace(rays))


def nan_rays(n=64, seed=0):
    """Rays a NaN-t hit leaves behind (origin NaN in all components; direction finite, NaN or zero), plus partially
    non-finite rays that must take the reference-order walk."""
    rng = np.random.default_rng(seed)
    r = np.empty((n, 8), F); r[:, 3] = 1e-3; r[:, 7] = np.inf
    r[:, 0:3] = np.nan
    r[:, 4:7] = rng.normal(size=(n, 3))
    r[n // 4: n // 2, 4:7] = np.nan
    r[n // 2: 5 * n // 8, 4:7] = 0
    k = 3 * n // 4
    r[k:, 0:3] = rng.normal(size=(n - k, 3)) + np.array([0, 1, 0])      # partially non-finite from here on
    r[k::3, 0] = np.nan; r[k + 1::3, 1] = np.inf; r[k + 2::3, 5] = np.nan
    return r


@pytest.mark.parametrize("mode", [abi.BVH_REFERENCE, abi.BVH_SAH])
def test_nan_rays(orc, hsim, mode):
    """An all-NaN origin makes every primitive accept (all comparisons false): the reference ends with the last one in
    its visiting order.  The ordered traversal answers that without walking the scene."""
    cases = [spheres_scene(), rects_scene(), mesh_scene(8), mesh_scene(8, ops=OPS[4]), mesh_scene(8, ops=OPS[3], with_floor=False)]
    sd, _, _ = scenes.cornell_smoke_scene(glass_sphere=False); cases.append(sd)          # last object is a medium
    sd = mesh_scene(8); m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    pos, nrm, uv = scenes.blob_mesh(2, 2, 1.0, 0.0, seed=1); sd.add(sd.mesh(pos[:0], nrm[:0], uv[:0], m)); cases.append(sd)   # last object: empty mesh
    for sd in cases:
        o = sd.build(orc, 0, 5); h = sd.build(hsim, mode, 5)
        rays = nan_rays()
        ro, rh = o.trace(rays, rng_seed=3, extras=False), h.trace(rays, rng_seed=3, extras=False)
        assert np.array_equal(ro["prim"], rh["prim"]) and np.array_equal(bits(ro["t"]), bits(rh["t"]))
        assert (ro["prim"][:32] != abi.MISS).all()


def test_env_map_rgbe_repack_is_lossless(orc, hsim):
    """Decoded Radiance images are stored as one RGBE word per texel on the device (hrt_host.cpp try_pack_rgbe): lookups
    must equal the float lookups of texture.cpp:76-97 bit for bit; anything not exactly representable keeps its floats."""
    hsim._sig("texture_kind", abi.C.c_int, [abi.C.c_void_p, abi.C.c_int])
    rng = np.random.default_rng(3)
    rgbe = rng.integers(0, 256, (64, 128, 4), dtype=np.uint8)
    rgbe[..., 3] = rng.integers(100, 160, (64, 128))            # exponents around 1.0
    rgbe[0, :8] = 0                                             # black texels (e = 0)
    rgbe[1, :8, :3] = [[255, 0, 1]] * 8                         # un-normalised mantissas decode and repack too
    exact = scenes.rgbe_to_float(rgbe)
    inexact = exact.copy(); inexact[5, 7, 1] = np.float32(0.1)  # 0.1 has no 8-bit mantissa
    negative = exact.copy(); negative[9, 3, 0] = -1.0
    uvp = rng.random((5000, 5)).astype(np.float32) * 1.2 - 0.1
    for data, packed in ((exact, True), (inexact, False), (negative, False)):
        sd = abi.SceneDesc()
        t = sd.env(data)
        sd.background = t
        sd.add(sd.sphere((0, 0, 0), 1, sd.material(abi.MAT_LAMBERTIAN)))
        o, h = sd.build(orc), sd.build(hsim)
        assert (hsim.fn("texture_kind")(h.h, t) == 4) == packed
        assert np.array_equal(bits(o.texture_eval(t, uvp)), bits(h.texture_eval(t, uvp)))
