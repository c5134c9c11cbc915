"""GPU parity tests proper: the CUDA library, through its C ABI, against the oracle on the same seeded inputs.

Gate 1: on fixed ray batches primitive ids are bit-exact and t is bit-exact (stricter than the
1e-5 relative the north star asks for) in BOTH traversal orders.  Gate 2: converged renders match the oracle's render
within its own noise floor.  Size-independent properties cover the full-size configurations.
"""
import os

import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes
from test_hostsim_vs_oracle import OPS, bits, mesh_scene, rects_scene, spheres_scene

pytestmark = pytest.mark.gpu
F = np.float32
MODES = [abi.BVH_REFERENCE, abi.BVH_SAH]


def check(o, g, uv_tol=None, front=True):
    assert np.array_equal(o["prim"], g["prim"]), f"prim ids differ at {np.nonzero(o['prim'] != g['prim'])[0][:10]}"
    assert np.array_equal(bits(o["t"]), bits(g["t"])), "t not bit-exact"
    assert np.array_equal(bits(o["normal"]), bits(g["normal"])), "normal not bit-exact"
    if uv_tol is None:
        assert np.array_equal(bits(o["uv"]), bits(g["uv"]))
    else:   # sphere u,v go through acos/atan2: glibc on the CPU, libdevice on the GPU
        assert np.allclose(o["uv"], g["uv"], atol=uv_tol, rtol=0, equal_nan=True)   # |outward.y| > 1 by an ulp -> acos = NaN on both
    if front:
        assert np.array_equal(o["front"], g["front"])


@pytest.mark.parametrize("mode", MODES)
def test_spheres_rects(orc, hrt, mode):
    for sd, center, rad, seed, uv_tol in ((spheres_scene(), (0, 0, -1), 3.0, 1, 2e-6), (rects_scene(), (0, 2.5, 0), 2.4, 2, None)):
        o = sd.build(orc, 0, 5); g = sd.build(hrt, mode, 5)
        rays = scenes.random_rays(200000, center, rad, seed=seed)
        rays[::7, 3] = 0.5; rays[::5, 7] = 2.0
        rays[:100, 4:7] = [0, 0, -1]; rays[100:200, 4:7] = [1, 0, 0]
        check(o.trace(rays), g.trace(rays), uv_tol=uv_tol)
    kat = np.array([[0, 0, 0, 1e-3, 0, 0, -1, np.inf], [0, 0, 0, 1e-3, 0, 0, -2, np.inf]], F)     # SURVEY §8c KATs
    r = spheres_scene().build(hrt, mode, 5).trace(kat)
    assert r["prim"].tolist() == [1, 1] and r["t"].tolist() == [0.5, 0.25] and r["front"].tolist() == [1, 1]


@pytest.mark.parametrize("mode", MODES)
def test_mesh(orc, hrt, mode):
    sd = mesh_scene(56)
    o = sd.build(orc, 0, 9); g = sd.build(hrt, mode, 9)
    rays = scenes.random_rays(500000, (0, 1, 0), 2.5, seed=4)
    ro, rg = o.trace(rays), g.trace(rays)
    assert (ro["prim"] != abi.MISS).mean() > 0.3
    check(ro, rg, front=False)
    hit = ro["prim"] != abi.MISS
    p = rays[hit, 0:3] + ro["t"][hit, None] * rays[hit, 4:7]
    rng = np.random.default_rng(0)
    sec = np.empty((p.shape[0], 8), F); sec[:, 0:3] = p; sec[:, 3] = 1e-3; sec[:, 7] = np.inf
    d = rng.normal(size=(p.shape[0], 3)); sec[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    so, sg = o.trace(sec), g.trace(sec)
    check(so, sg, front=False)
    assert ((so["prim"] != abi.MISS) & (so["t"] < 1e-3)).sum() > 0          # sub-t_min self hits reproduced


@pytest.mark.parametrize("mode", MODES)
def test_mesh_degenerate_and_tiny(orc, hrt, mode):
    sd = mesh_scene(8, with_floor=False)
    o = sd.build(orc, 0, 5); g = sd.build(hrt, mode, 5)
    rays = np.array([[0, 1, -3, 1e-3, 0, 0, 1, np.inf], [3, 1, 0, 1e-3, -1, 0, 0, np.inf],
                     [0, 4, 0, 1e-3, 0, -1, 0, np.inf], [0.3, 1.2, -3, 1e-3, 0, 0, 1, np.inf]], F)
    ro, rg = o.trace(rays), g.trace(rays)
    assert np.array_equal(ro["prim"], rg["prim"]) and np.array_equal(bits(ro["t"]), bits(rg["t"]))   # incl. NaN-t hits
    for n_tri in (0, 1, 2, 3):
        sd = abi.SceneDesc(); m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
        pos, nrm, uv = scenes.blob_mesh(2, 2, 1.0, 0.0, seed=1)
        sd.add(sd.mesh(pos[:n_tri], nrm[:n_tri], uv[:n_tri], m)); sd.add(sd.sphere((0, 0, 0), 0.3, m))
        o = sd.build(orc, 0, 5); g = sd.build(hrt, mode, 5)
        rays = scenes.random_rays(3000, (0, 0, 0), 2.0, seed=n_tri)
        check(o.trace(rays), g.trace(rays), uv_tol=2e-6, front=False)
    # empty batch / empty world
    sd = abi.SceneDesc(); sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    g = sd.build(hrt, mode, 5)
    r = g.trace(scenes.random_rays(100, (0, 0, 0), 1.0))
    assert (r["prim"] == abi.MISS).all()
    assert g.trace(np.zeros((0, 8), F))["prim"].shape == (0,)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("ops", OPS)
def test_instances(orc, hrt, mode, ops):
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
        o = sd.build(orc, 0, 5); g = sd.build(hrt, mode, 5)
        rays = scenes.random_rays(50000, (0, 0.5, 0), 3.5, seed=5)
        rays[:5000, 4:7] *= 3.7; rays[5000:8000, 4:7] *= 0.05
        check(o.trace(rays), g.trace(rays), uv_tol=2e-6)


def test_scattered_balls_tlas(orc, hrt):
    sd, cam, _ = scenes.scattered_balls_scene(seed=42, grid=11)
    o = sd.build(orc, 0, 5)
    rays = np.concatenate([scenes.random_rays(100000, (0, 1, 0), 12.0, seed=11), scenes.camera_rays(cam, 1200, 800, 100000, seed=2)])
    ro = o.trace(rays)
    for mode in MODES:
        check(ro, sd.build(hrt, mode, 5).trace(rays), uv_tol=2e-6)


def test_media(orc, hrt):
    sd, cam, _ = scenes.cornell_smoke_scene(glass_sphere=True)
    o = sd.build(orc, 0, 5)
    rays = scenes.random_rays(200000, (278, 278, 278), 270, seed=6); rays[:50000, 4:7] *= 2.5
    ro = o.trace(rays, rng_seed=77)
    media = [o.prim_base(8), o.prim_base(9)]
    res = {}
    for mode in MODES:
        rg = sd.build(hrt, mode, 5).trace(rays, rng_seed=77); res[mode] = rg
        for mp in media:     # media draw a random number inside hit(): statistical agreement only
            fo, fg = (ro["prim"] == mp).mean(), (rg["prim"] == mp).mean()
            assert abs(fo - fg) < 0.004 + 0.05 * fo, (mp, fo, fg)
            assert abs(ro["t"][ro["prim"] == mp].mean() - rg["t"][rg["prim"] == mp].mean()) < 0.03 * ro["t"][ro["prim"] == mp].mean()
    a, b = res[abi.BVH_REFERENCE], res[abi.BVH_SAH]
    assert np.array_equal(a["prim"], b["prim"]) and np.array_equal(bits(a["t"]), bits(b["t"]))
    sd.objects = [ob for ob in sd.objects if ob[0] == "object"]
    o2 = sd.build(orc, 0, 5)
    for mode in MODES:
        check(o2.trace(rays), sd.build(hrt, mode, 5).trace(rays), uv_tol=2e-6)


def test_teapot_scene_rays(orc, hrt, cache_dir):
    """C1 geometry (Cornell rects + mesh under rotate/scale/translate): camera rays and incoherent rays."""
    sd, cam, _ = scenes.teapot_scene(cache_dir, env_w=64, env_h=32)
    o = sd.build(orc, 0, 3)
    rays = np.concatenate([scenes.camera_rays(cam, 640, 640, 150000, seed=1), scenes.random_rays(150000, (0, 2.5, 0), 2.4, seed=2)])
    ro = o.trace(rays)
    for mode in MODES:
        check(ro, sd.build(hrt, mode, 3).trace(rays))


# ---- renders ---------------------------------------------------------------------------------------------------------------
def render_case(name, cache_dir):
    if name == "teapot_scene":
        sd, cam, _ = scenes.teapot_scene(cache_dir, n=24, env_w=128, env_h=64); return sd, cam, 128, 128, 128
    if name == "shiny_teapot":
        sd, cam, _ = scenes.shiny_teapot_scene(cache_dir, n=24, env_w=256, env_h=128); return sd, cam, 192, 112, 128
    if name == "shiny_teapot_glass":
        sd, cam, _ = scenes.shiny_teapot_scene(cache_dir, n=24, env_w=256, env_h=128, rough_dielectric=True); return sd, cam, 192, 112, 128
    if name == "scattered_balls":
        sd, cam, _ = scenes.scattered_balls_scene(seed=42, grid=4); return sd, cam, 192, 128, 64
    if name == "cornell_smoke":
        sd, cam, _ = scenes.cornell_smoke_scene(glass_sphere=True); return sd, cam, 96, 96, 256
    raise KeyError(name)


def block_sums(img, ok, bs=16):
    """Sums over bs x bs pixel blocks (all channels) of the pixels flagged ok -> (sums, element counts), one entry per block."""
    h, w, _ = img.shape
    hb, wb = h // bs, w // bs
    v = np.where(ok[..., None], img, 0.0)[:hb * bs, :wb * bs].astype(np.float64).reshape(hb, bs, wb, bs, 3).sum(axis=(1, 3, 4))
    n = ok[:hb * bs, :wb * bs].reshape(hb, bs, wb, bs).sum(axis=(1, 3)) * 3
    return v.reshape(-1), n.reshape(-1)


def block_gate(a, b, g, ok, bs=16):
    """16 x 16 block means of g against the mean of the two oracle renders a, b.  The noise of ONE render's block mean is measured in
    the block itself, from the per-pixel sample variance over the three renders (heavy tails: a firefly that only one render caught
    widens that block's limit, a bias of g over a whole block adds delta^2 / 3 per pixel to it and stays visible).  Returns the largest
    |g_b - (a_b + b_b) / 2| in units of its limit max(1 % of the block mean, 7 sigma of that difference: oracle against oracle reaches 4.5 - 5.8)."""
    sa, n = block_sums(a, ok, bs); sb, _ = block_sums(b, ok, bs); sg, _ = block_sums(g, ok, bs)
    m = (a.astype(np.float64) + b + g) / 3.0
    sd2, _ = block_sums(((a - m) ** 2 + (b - m) ** 2 + (g - m) ** 2) / 2.0, ok, bs)
    full = n == n.max()
    n = np.maximum(n, 1)
    mid = 0.5 * (sa + sb) / n
    sigma = np.sqrt(sd2) / n
    limit = np.maximum(0.01 * np.abs(mid), 7.0 * sigma * np.sqrt(1.5))
    return float((np.abs(sg / n - mid)[full] / np.maximum(limit[full], 1e-12)).max()), int(full.sum())


@pytest.mark.parametrize("name", ["teapot_scene", "shiny_teapot", "shiny_teapot_glass", "scattered_balls", "cornell_smoke"])
def test_render_matches_oracle_within_noise(orc, hrt, cache_dir, name):
    """Noise-floor-calibrated image gate, GPU against two oracle renders A, B of other seeds at equal spp:
      * RMSE(gpu, A) <= 1.25 RMSE(A, B);
      * 16 x 16 block means: |gpu - (A + B) / 2| <= 1 % of the block mean, or 7 sigma of that block's own noise, measured from the
        per-pixel differences A - B inside it (a biased lobe that keeps the global mean shows here);
      * global mean within 0.5 %, or 3 |mean A - mean B|;
      * rays per camera sample within 1 % of the oracle's (the path-length statistics: scatter verdicts, absorption, misses);
      * no sample was clamped by the film;
    the three image gates on radiance clipped at the 99.9th percentile of A (the reference is a brute-force path tracer: a handful of
    fireflies carries most of the variance of a Cornell-type frame)."""
    sd, cam, w, h, spp = render_case(name, cache_dir)
    c = scenes.camera_for(cam, w, h)
    o = sd.build(orc, 0, 3)
    A, rays_a = o.render_f32(c, w, h, spp, seed=101); B, rays_b = o.render_f32(c, w, h, spp, seed=202)
    A /= spp; B /= spp
    ok = np.isfinite(A).all(-1) & np.isfinite(B).all(-1)
    lim = float(np.percentile(A[ok], 99.9))
    for mode in MODES:
        g = sd.build(hrt, mode, 3)
        S, rgb = g.render(c, w, h, spp, seed=7)
        st = g.stats()
        G = S / spp
        okg = ok & np.isfinite(G).all(-1)
        assert okg.mean() > 0.99
        a, b, gg = np.minimum(A, lim), np.minimum(B, lim), np.minimum(G, lim)
        rmse = lambda x, y: float(np.sqrt(np.mean((x[okg] - y[okg]) ** 2)))
        floor = rmse(a, b)
        assert rmse(gg, a) <= 1.25 * floor + 1e-6, (name, mode, rmse(gg, a), floor)
        ma, mb, mg = a[okg].mean(), b[okg].mean(), gg[okg].mean()
        assert abs(mg - 0.5 * (ma + mb)) <= max(0.005 * ma, 3 * abs(ma - mb)), (name, mode, ma, mb, mg)
        worst, n_blocks = block_gate(a, b, gg, okg)
        assert worst <= 1.0 and n_blocks >= 30, (name, mode, worst, n_blocks)
        assert abs(st.rays / st.paths - 0.5 * (rays_a + rays_b) / (w * h * spp)) <= 0.01 * rays_a / (w * h * spp), (name, mode, st.rays / st.paths, rays_a / (w * h * spp))
        assert st.film_clamped == 0
        # 8-bit film: the GPU's tonemap of its own sums equals the oracle's film code on the same sums, bit for bit
        ref8 = np.empty((h, w, 3), np.uint8)
        orc.check(orc.fn("film_resolve")(np.ascontiguousarray(S).ctypes.data, w, h, spp, ref8.ctypes.data))
        assert np.array_equal(ref8, rgb)


def eight_bit_case(name, cache_dir):
    """-> scene, camera, oracle spp: enough for the oracle's own 8-bit noise to stay under the bound (measured on the CPU, oracle against
    oracle: RMSE(A, B) at 256 spp each is 3.2 on the shiny teapot, 2.0 on the balls, 3.7 on the big mesh; shiny teapot at 384 spp against
    4096 spp 2.18 - the GPU measured 2.17 against the same 384 - and at 768 spp 1.85; the big mesh, whose bright-window reflections are
    heavy tailed, 768 against 2048 spp 2.42 - the GPU 2.52 - and 1024 against 2048 still 2.28: it takes 2048 spp)"""
    if name == "shiny_teapot":
        sd, cam, _ = scenes.shiny_teapot_scene(cache_dir, n=24, env_w=256, env_h=128); return sd, cam, 1024
    if name == "scattered_balls":
        sd, cam, _ = scenes.scattered_balls_scene(seed=42, grid=4); return sd, cam, 256
    sd, cam, _ = scenes.big_mesh_scene(cache_dir, quads=120, env_w=256, env_h=128); return sd, cam, 2048


@pytest.mark.parametrize("name", ["shiny_teapot", "scattered_balls", "big_mesh"])
def test_eight_bit_image_against_oracle(orc, hrt, cache_dir, name):
    """SURVEY.md §8d, env / sky-lit configs (C2, C3, C5 style) at 480 x 270: the final 8-bit image (ACES + gamma) within 2 / 255 RMSE
    of the CPU oracle's.  Spp on each side: the oracle at 256-2048 (see eight_bit_case), the GPU at 8192 - its own noise then adds
    little to the oracle's, which is what the bound has to absorb."""
    sd, cam, spp_cpu = eight_bit_case(name, cache_dir)
    w, h, spp_gpu = 480, 270, 8192
    c = scenes.camera_for(cam, w, h)
    A, rays_a = sd.build(orc, 0, 3).render_f32(c, w, h, spp_cpu, seed=31)
    a8 = np.empty((h, w, 3), np.uint8)
    orc.check(orc.fn("film_resolve")(np.ascontiguousarray(A).ctypes.data, w, h, spp_cpu, a8.ctypes.data))
    g = sd.build(hrt, abi.BVH_SAH, 3)
    _, g8 = g.render(c, w, h, spp_gpu, seed=7)
    st = g.stats()
    rmse8 = float(np.sqrt(np.mean((a8.astype(np.float64) - g8.astype(np.float64)) ** 2)))
    assert rmse8 <= 2.0, (name, rmse8)
    assert abs(a8.astype(np.float64).mean() - g8.astype(np.float64).mean()) <= 0.5, "mean 8-bit level"
    assert abs(st.rays / st.paths - rays_a / (w * h * spp_cpu)) <= 0.01 * rays_a / (w * h * spp_cpu)
    assert st.film_clamped == 0


def test_wavefront_equals_megakernel_and_sample_split(hrt, cache_dir):
    sd, cam, _ = scenes.teapot_scene(cache_dir, n=24, env_w=128, env_h=64)
    w, h, spp = 64, 48, 24
    c = scenes.camera_for(cam, w, h)
    for mode in MODES:
        g = sd.build(hrt, mode, 3)
        wf, _ = g.render(c, w, h, spp, seed=9)
        mk, _ = g.render(c, w, h, spp, seed=9, flags=abi.RENDER_MEGAKERNEL)
        assert np.array_equal(bits(wf), bits(mk)), "wavefront pipeline differs from the one-thread-per-path loop"
        assert g.stats().rays % 2 == 0              # two renders of the same paths counted the same rays
        # opt-in Russian roulette: off by default (the two renders above), the same bits from both pipelines when on, fewer rays
        rays_plain = g.stats().rays // 2
        rr, _ = g.render(c, w, h, spp, seed=9, flags=abi.RENDER_RR)
        rays_rr = g.stats().rays - 2 * rays_plain
        rm, _ = g.render(c, w, h, spp, seed=9, flags=abi.RENDER_RR | abi.RENDER_MEGAKERNEL)
        assert np.array_equal(bits(rr), bits(rm)) and not np.array_equal(bits(rr), bits(wf))
        assert rays_rr < 0.8 * rays_plain, (rays_rr, rays_plain)
        # samples are keyed by (pixel, global sample index) and the film adds integers: the wavefront capacity (hence the
        # order in which paths start and finish) does not change a single bit
        for cap in (1024, 5000):
            os.environ["HRT_WAVE_CAPACITY"] = str(cap)
            try:
                wf2, _ = g.render(c, w, h, spp, seed=9)
            finally:
                del os.environ["HRT_WAVE_CAPACITY"]
            assert np.array_equal(bits(wf), bits(wf2))
        a, _ = g.render(c, w, h, spp, seed=9, sample_first=0, sample_count=10, want_rgb8=False)
        b, _ = g.render(c, w, h, spp, seed=9, sample_first=10, sample_count=14, want_rgb8=False)
        assert np.allclose(a + b, wf, rtol=1e-5, atol=1e-6)
    ref = sd.build(hrt, abi.BVH_REFERENCE, 3).render(c, w, h, spp, seed=9)[0]
    fast = sd.build(hrt, abi.BVH_SAH, 3).render(c, w, h, spp, seed=9)[0]
    assert np.array_equal(bits(ref), bits(fast)), "SAH-order render differs from reference-order render"


@pytest.mark.parametrize("name", ["teapot_scene", "cornell_smoke"])
def test_russian_roulette_is_unbiased(orc, hrt, cache_dir, name):
    """HRT_RENDER_RR against the oracle, which has no roulette (main.cpp:43-76): same image within the noise floor (the roulette adds
    variance: the GPU renders 4 x the oracle's spp), global mean within 0.5 % or the oracle's own spread, rays per sample down."""
    sd, cam, w, h, spp = render_case(name, cache_dir)
    c = scenes.camera_for(cam, w, h)
    o = sd.build(orc, 0, 3)
    A, rays_a = o.render_f32(c, w, h, spp, seed=101); B, _ = o.render_f32(c, w, h, spp, seed=202)
    A /= spp; B /= spp
    ok = np.isfinite(A).all(-1) & np.isfinite(B).all(-1)
    lim = float(np.percentile(A[ok], 99.9))
    g = sd.build(hrt, abi.BVH_SAH, 3)
    S, _ = g.render(c, w, h, 4 * spp, seed=7, flags=abi.RENDER_RR)
    st = g.stats()
    G = S / (4 * spp)
    a, b, gg = np.minimum(A, lim), np.minimum(B, lim), np.minimum(G, lim)
    rmse = lambda x, y: float(np.sqrt(np.mean((x[ok] - y[ok]) ** 2)))
    assert rmse(gg, a) <= 1.1 * rmse(a, b), (rmse(gg, a), rmse(a, b))
    ma, mb, mg = a[ok].mean(), b[ok].mean(), gg[ok].mean()
    assert abs(mg - 0.5 * (ma + mb)) <= max(0.005 * ma, 3 * abs(ma - mb)), (ma, mb, mg)
    worst, _ = block_gate(a, b, gg, ok)
    assert worst <= 1.0, worst
    assert st.rays / st.paths < 0.75 * rays_a / (w * h * spp), (st.rays / st.paths, rays_a / (w * h * spp))


@pytest.mark.parametrize("case", ["teapot_scene", "cornell_smoke", "scattered_balls", "instances"])
def test_wide_kernels_equal_megakernel(hrt, cache_dir, case):
    """The kernels that walk the 8-wide trees — k_wide_secondary (phases + per-lane refill), k_wave2_trace<.., WIDE> (phased walk of a
    claim) and the pending list + k_wave3_resolve behind them — are picked by the host from scene statistics (large meshes, many
    secondary rays).  Here they are forced onto small scenes of every kind: the film must equal the one-thread-per-path loop over the
    two-child trees bit for bit."""
    if case == "instances":
        sd = mesh_scene(16, ops=OPS[6]); m = sd.material(abi.MAT_METAL, (0.8, 0.8, 0.8), p0=0.1)
        sd.add(sd.sphere((2, 1, 0), 0.7, m), OPS[1]); sd.add(sd.box((-2, 1, 0), (1, 2, 1), m), OPS[3]); sd.background = sd.solid((0.7, 0.8, 1.0))
        cam = dict(position=(0, 2, 7), look_at=(0, 1, 0), up=(0, 1, 0), vfov_deg=50, aperture=0.0, focal_distance=7)
        w, h, spp = 64, 48, 16
    else:
        sd, cam, w, h, spp = render_case(case, cache_dir); w, h, spp = 64, 48, 16
    c = scenes.camera_for(cam, w, h)
    g = sd.build(hrt, abi.BVH_SAH, 3)
    mk, _ = g.render(c, w, h, spp, seed=9, flags=abi.RENDER_MEGAKERNEL)
    for env in (("HRT_FORCE_WIDE_SECONDARY",), ("HRT_FORCE_WIDE_PRIMARY",), ("HRT_FORCE_WIDE_SECONDARY", "HRT_FORCE_WIDE_PRIMARY")):
        for e in env:
            os.environ[e] = "1"
        try:
            wf, _ = g.render(c, w, h, spp, seed=9)
        finally:
            for e in env:
                del os.environ[e]
        assert np.array_equal(bits(wf), bits(mk)), (case, env)


def test_film_is_split_invariant(hrt, cache_dir):
    """The multi-GPU contract on one GPU: films of disjoint sample ranges added as integers (what the NCCL reduce does)
    equal the film of one call over all samples, bit for bit, and so do the resolved images."""
    import torch
    from hobbyraytracer_b200 import multigpu
    sd, cam, _ = scenes.teapot_scene(cache_dir, n=24, env_w=128, env_h=64)
    w, h, spp = 96, 64, 20                      # 96 x 64: whole 8x4 tiles; the odd size below takes the row-order mapping
    c = scenes.camera_for(cam, w, h)
    g = sd.build(hrt, abi.BVH_SAH, 3)
    whole = multigpu.new_film(hrt, w, h)
    multigpu.render_split(g, c, w, h, spp, whole, 0, 1, None, seed=5)
    for world in (2, 3, 8):
        parts = torch.zeros_like(whole)
        for rank in range(world):
            f = multigpu.new_film(hrt, w, h)
            first, count = multigpu.sample_range(spp, rank, world)
            p = abi.RenderParams(w, h, spp, first, count, 50, 0.001, 5, 0, 0)
            import ctypes as C
            hrt.check(hrt.fn("render_accumulate_device")(g.h, C.byref(c), C.byref(p), f.data_ptr(), None), "render_accumulate_device")
            torch.cuda.synchronize()
            parts += f
        assert torch.equal(parts, whole), f"film differs for a {world}-way sample split"
    rgb = torch.zeros(w * h * 3, dtype=torch.uint8, device="cuda"); tot = torch.zeros(w * h * 3, dtype=torch.float32, device="cuda")
    hrt.check(hrt.fn("film_resolve_device")(g.h, whole.data_ptr(), w, h, spp, tot.data_ptr(), None, rgb.data_ptr(), None), "film_resolve_device")
    torch.cuda.synchronize()
    S, B = g.render(c, w, h, spp, seed=5)
    assert np.array_equal(bits(S.reshape(-1)), bits(tot.cpu().numpy())) and np.array_equal(B.reshape(-1), rgb.cpu().numpy())
    # a film that is not a whole number of tiles
    w2, h2 = 61, 37
    c2 = scenes.camera_for(cam, w2, h2)
    a, _ = g.render(c2, w2, h2, 6, seed=3)
    b, _ = g.render(c2, w2, h2, 6, seed=3, flags=abi.RENDER_MEGAKERNEL)
    assert np.array_equal(bits(a), bits(b)) and np.isfinite(a).all() and a.mean() > 0


def test_large_batch_orders_agree(hrt, cache_dir):
    """Full-size property (16 M-ray micro-benchmark scale is run by bench.py; 4 M here): the ordered traversal and the
    reference-order traversal return identical (prim, t) for every ray."""
    sd, cam, _ = scenes.shiny_teapot_scene(cache_dir, env_w=64, env_h=32)
    rays = scenes.random_rays(4_000_000, (0, 0.8, 0), 1.5 * 1.2)
    a = sd.build(hrt, abi.BVH_REFERENCE, 3).trace(rays, extras=False)
    b = sd.build(hrt, abi.BVH_SAH, 3).trace(rays, extras=False)
    assert np.array_equal(a["prim"], b["prim"]) and np.array_equal(bits(a["t"]), bits(b["t"]))
    assert (a["prim"] != abi.MISS).mean() > 0.15


def test_full_resolution_wavefront_equals_megakernel(hrt, cache_dir):
    """BASELINE config 2 at its own resolution (1920 x 1080; 40 spp = 83 M camera paths, more than the 32 Mi queue entries, so the
    camera-path kernel fills the queues and stops several times): the persistent wavefront and the one-thread-per-path kernel produce the
    same film bit for bit, nothing is clamped, and the rays-per-sample figure is the one the bench reports (1.10)."""
    sd, cam, _ = scenes.shiny_teapot_scene(cache_dir, env_w=512, env_h=256)
    w, h, spp = 1920, 1080, 40
    c = scenes.camera_for(cam, w, h)
    g = sd.build(hrt, abi.BVH_SAH, 3)
    a, img = g.render(c, w, h, spp, seed=77)
    st = g.stats()
    b, _ = g.render(c, w, h, spp, seed=77, flags=abi.RENDER_MEGAKERNEL)
    assert np.array_equal(bits(a), bits(b))
    assert st.film_clamped == 0 and st.paths == w * h * spp
    assert 1.05 < st.rays / st.paths < 1.15
    assert img.shape == (h, w, 3) and img.mean() > 10


@pytest.mark.parametrize("name,w,h,spp", [("teapot_scene", 640, 640, 12), ("cornell_smoke", 1024, 1024, 6), ("scattered_balls", 1200, 800, 6)])
def test_other_configs_at_their_resolution(hrt, cache_dir, name, w, h, spp):
    """BASELINE configs 1, 4 and 3 at their own resolutions (a few samples per pixel): every pipeline the host picks for them - wide
    secondary pass with per-scene thresholds, lock-step pass, media inside hit() - gives the megakernel's film bit for bit."""
    if name == "teapot_scene":
        sd, cam, _ = scenes.teapot_scene(cache_dir, n=56, env_w=64, env_h=32)
    elif name == "cornell_smoke":
        sd, cam = scenes.cornell_smoke_scene()[:2]
    else:
        sd, cam = scenes.scattered_balls_scene()[:2]
    c = scenes.camera_for(cam, w, h)
    g = sd.build(hrt, abi.BVH_SAH, 3)
    for rep in range(2):                                   # the second render starts from the kernel choice the first one found
        a, _ = g.render(c, w, h, spp, seed=5 + rep)
        assert g.stats().film_clamped == 0
        b, _ = g.render(c, w, h, spp, seed=5 + rep, flags=abi.RENDER_MEGAKERNEL)
        assert np.array_equal(bits(a), bits(b)), (name, rep)


def test_chunked_host_batch_equals_one_launch(hrt, cache_dir):
    """hrt_trace_batch on a large host batch runs in chunks on three streams (copies overlap the walks): the answers are those of the
    one-launch path, bit for bit, including the ragged last chunk."""
    sd, cam, _ = scenes.shiny_teapot_scene(cache_dir, env_w=64, env_h=32)
    rays = scenes.random_rays(5 * (1 << 20) + 12345, (0, 0.8, 0), 1.5 * 1.2, seed=21)
    sc = sd.build(hrt, abi.BVH_SAH, 3)
    a = sc.trace(rays, extras=False)                                  # >= 4 Mi rays, no extras, no media: chunked
    os.environ["HRT_TRACE_BATCH_ONE_LAUNCH"] = "1"
    try:
        b = sc.trace(rays, extras=False)
    finally:
        del os.environ["HRT_TRACE_BATCH_ONE_LAUNCH"]
    assert np.array_equal(a["prim"], b["prim"]) and np.array_equal(bits(a["t"]), bits(b["t"]))
    assert sc.stats().rays == 2 * len(rays)


@pytest.mark.parametrize("mode", MODES)
def test_nan_rays(orc, hrt, mode):
    """Rays left behind by NaN-t hits (all-NaN origin: answered without a walk) and partially non-finite rays
    (reference-order walk) agree with the oracle, which is pinned to the reference on the same rays."""
    from test_hostsim_vs_oracle import nan_rays
    cases = [spheres_scene(), rects_scene(), mesh_scene(8), mesh_scene(8, ops=OPS[4])]
    sd, _, _ = scenes.cornell_smoke_scene(glass_sphere=False); cases.append(sd)
    for sd in cases:
        o = sd.build(orc, 0, 5); g = sd.build(hrt, mode, 5)
        rays = nan_rays()
        ro, rg = o.trace(rays, rng_seed=3, extras=False), g.trace(rays, rng_seed=3, extras=False)
        assert np.array_equal(ro["prim"], rg["prim"]) and np.array_equal(bits(ro["t"]), bits(rg["t"]))
