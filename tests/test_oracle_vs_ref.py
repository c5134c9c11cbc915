"""Pins the CPU restatement (oracle/oracle.cpp) against the UNMODIFIED reference (oracle/_ref, built from
/root/reference by oracle/Makefile): every comparison here is bit-exact, on the same rand() stream.

Skipped automatically where neither /root/reference nor a prebuilt oracle/_ref exists (then tests/test_golden.py
pins the restatement against committed vectors generated from the reference by tests/golden/make_golden.py).
"""
import ctypes

import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes

F = np.float32


def same_bits(a, b):
    a = np.ascontiguousarray(a); b = np.ascontiguousarray(b)
    if a.dtype.kind == "f":
        return np.array_equal(a.view(np.uint32), b.view(np.uint32)) or np.array_equal(a, b, equal_nan=True) and np.array_equal(np.signbit(a), np.signbit(b))
    return np.array_equal(a, b)


def assert_trace_equal(o, r, check_front=True, check_uv=True):
    assert np.array_equal(o["prim"], r["prim"]), f"prim ids differ at {np.nonzero(o['prim'] != r['prim'])[0][:10]}"
    assert same_bits(o["t"], r["t"])
    assert same_bits(o["normal"], r["normal"])
    if check_uv:
        assert same_bits(o["uv"], r["uv"])
    if check_front:
        assert np.array_equal(o["front"], r["front"])


def both(sd, orc, ref, seed=5):
    return sd.build(orc, abi.BVH_REFERENCE, seed), sd.build(ref, abi.BVH_REFERENCE, seed, tagged=True)


# ---------------------------------------------------------------------------------------------------------------
def test_rng_streams(orc, ref):
    for kind, n, width in ((0, 4096, 1), (1, 4096, 1), (2, 2048, 3), (3, 2048, 3), (4, 2048, 1)):
        a = np.zeros(n * width, np.float64); b = np.zeros(n * width, np.float64)
        orc.check(orc.fn("rng_draw")(kind, 11, 3, n, a.ctypes.data))
        ref.check(ref.fn("rng_draw")(kind, 11, 3, n, b.ctypes.data))
        assert np.array_equal(a, b), f"rng kind {kind}"
    # the byte quirk of random.inl:25: rand() % 255 -> linearRand(0f,1f) never exceeds 0xFEFEFEFE / 2^32
    assert a.max() <= 0.99608


def test_spheres_and_ties(orc, ref):
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.sphere((0, 0, -1), 0.5, m))
    sd.add(sd.sphere((0, 0, -1), 0.5, m))          # identical sphere: the later list entry must win (SURVEY §8c KAT)
    sd.add(sd.sphere((1.5, 0.2, -2), 0.75, m))
    sd.add(sd.sphere((0, -100.5, -1), 100, m))
    o, r = both(sd, orc, ref)
    kat = np.array([[0, 0, 0, 1e-3, 0, 0, -1, np.inf], [0, 0, 0, 1e-3, 0, 0, -2, np.inf]], F)
    ro = o.trace(kat)
    assert ro["prim"].tolist() == [1, 1] and ro["t"].tolist() == [0.5, 0.25]
    assert ro["uv"][0].tolist() == [0.25, 0.5] and ro["front"].tolist() == [1, 1]
    rays = scenes.random_rays(20000, (0, 0, -1), 3.0, seed=1)
    rays[::7, 3] = 0.5; rays[::5, 7] = 2.0          # varied t_min / t_max
    assert_trace_equal(o.trace(rays), r.trace(rays))


def test_rects_boxes(orc, ref):
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.rect(0, 0, 5, -2.5, 2.5, 2.5, m)); sd.add(sd.rect(0, 0, 5, -2.5, 2.5, -2.5, m))
    sd.add(sd.rect(1, -2.5, 2.5, -2.5, 2.5, 0, m)); sd.add(sd.rect(1, -2.5, 2.5, -2.5, 2.5, 5, m))
    sd.add(sd.rect(2, -2.5, 2.5, 0, 5, -2.5, m))
    sd.add(sd.box((0, 1, 0), (1.5, 2, 1), m))
    sd.add(sd.box_minmax((-2, 0, -2), (-1, 0.7, -1.2), m))
    sd.add(sd.rect(2, -0.75, 0.75, 0, 2, 0.5, m))    # coplanar with the first box's +z face: tie -> later wins
    o, r = both(sd, orc, ref)
    rays = scenes.random_rays(30000, (0, 2.5, 0), 2.4, seed=2)
    rays[:100, 4:7] = [0, 0, -1]                     # axis-aligned directions: division by zero paths
    rays[100:200, 4:7] = [1, 0, 0]
    assert_trace_equal(o.trace(rays), r.trace(rays))


def test_box_class_equals_six_rects(ref):
    """Tagged scenes replace Box by its six rects; check the untagged (real Box) build reports identical hits."""
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.box((0, 1, 0), (1.5, 2, 1), m), [(abi.XF_ROTATE_Y_DEG, (15,)), (abi.XF_TRANSLATE, (0.2, 0, 0.1))])
    a = sd.build(ref, 0, 1, tagged=True); b = sd.build(ref, 0, 1, tagged=False)
    rays = scenes.random_rays(5000, (0, 1, 0), 3.0, seed=3)
    ra, rb = a.trace(rays), b.trace(rays)
    assert np.array_equal(ra["prim"] == abi.MISS, rb["prim"] == abi.MISS)
    assert same_bits(ra["t"], rb["t"]) and same_bits(ra["normal"], rb["normal"])


def _mesh_scene(n=24, ops=(), with_floor=True):
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    pos, nrm, uv = scenes.blob_mesh(n, n, 1.0, 0.2, seed=7, center=(0, 1, 0))
    if with_floor:
        sd.add(sd.rect(1, -3, 3, -3, 3, 0, m))
    sd.add(sd.mesh(pos, nrm, uv, m), list(ops))
    return sd


def test_mesh_bvh(orc, ref):
    sd = _mesh_scene(24)
    o, r = both(sd, orc, ref, seed=9)
    rays = scenes.random_rays(40000, (0, 1, 0), 2.5, seed=4)
    ro, rr = o.trace(rays), r.trace(rays)
    assert (ro["prim"] != abi.MISS).mean() > 0.3
    # frontFace of a bare ITriangle hit is never written by the reference (triangle.cpp:57-131): not compared
    assert_trace_equal(ro, rr, check_front=False)
    # rays that start on the surface: ITriangle ignores t_min, so self-hits at tiny t must agree too
    hit = ro["prim"] != abi.MISS
    p = rays[hit, 0:3] + ro["t"][hit, None] * rays[hit, 4:7]
    rng = np.random.default_rng(0)
    sec = np.empty((p.shape[0], 8), F); sec[:, 0:3] = p; sec[:, 3] = 1e-3; sec[:, 7] = np.inf
    d = rng.normal(size=(p.shape[0], 3)); sec[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    so, sr = o.trace(sec), r.trace(sec)
    assert_trace_equal(so, sr, check_front=False)
    tiny = (so["prim"] != abi.MISS) & (so["t"] < 1e-3)
    assert tiny.sum() > 0, "expected some sub-t_min self hits (reference quirk)"


def test_mesh_degenerate_rays(orc, ref):
    """d[kZ] == 0 makes ITriangle return a NaN-t 'hit' (SURVEY §8a row 11)."""
    sd = _mesh_scene(8, with_floor=False)
    o, r = both(sd, orc, ref)
    rays = np.array([[0, 1, -3, 1e-3, 0, 0, 1, np.inf],      # o.z smallest -> kZ from origin; d[kZ] may be 0
                     [3, 1, 0, 1e-3, -1, 0, 0, np.inf],
                     [0, 4, 0, 1e-3, 0, -1, 0, np.inf],
                     [0.3, 1.2, -3, 1e-3, 0, 0, 1, np.inf]], F)
    ro, rr = o.trace(rays), r.trace(rays)
    assert np.array_equal(ro["prim"], rr["prim"])
    assert np.array_equal(np.isnan(ro["t"]), np.isnan(rr["t"]))
    assert same_bits(ro["t"], rr["t"])


def test_nan_origin_rays(orc, ref):
    """What a NaN-t hit leaves behind: rays whose origin is NaN.  Every comparison in hit() is false, so the reference
    accepts every primitive and ends with the last one in its visiting order; the oracle must agree (the CUDA path
    answers these without walking the scene, hrt_trace.cuh)."""
    from test_hostsim_vs_oracle import nan_rays
    sd = _mesh_scene(8)
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.sphere((0.5, 0.5, 0.5), 0.25, m)); sd.add(sd.box((0, 1, 0), (1.5, 2, 1), m))
    for extra in (None, "mesh_last"):
        if extra:
            pos, nrm, uv = scenes.blob_mesh(6, 6, 1.0, 0.2, seed=5, center=(1, 1, 0)); sd.add(sd.mesh(pos, nrm, uv, m), [(abi.XF_TRANSLATE, (0.5, 0, 0))])
        o, r = both(sd, orc, ref)
        rays = nan_rays()
        ro, rr = o.trace(rays), r.trace(rays)
        assert np.array_equal(ro["prim"], rr["prim"]) and np.array_equal(np.isnan(ro["t"]), np.isnan(rr["t"]))
        assert same_bits(np.nan_to_num(ro["t"], nan=0.0), np.nan_to_num(rr["t"], nan=0.0))
        assert (ro["prim"][:32] != abi.MISS).all() and len(set(ro["prim"][:32].tolist())) == 1


@pytest.mark.parametrize("ops", [
    [(abi.XF_TRANSLATE, (0.5, -0.25, 1.0))],
    [(abi.XF_SCALE, (1.5, 0.5, 2.0))],
    [(abi.XF_ROTATE_EULER_DEG, (10, 180, -35))],
    [(abi.XF_ROTATE_Y_DEG, (15,))],
    [(abi.XF_ROTATE_EULER_DEG, (0, 180, 0)), (abi.XF_SCALE, (1.4, 1.4, 1.4)), (abi.XF_TRANSLATE, (0, 1, 0))],   # teapot_scene.yaml
    [(abi.XF_ROTATE_Y_DEG, (-18,)), (abi.XF_TRANSLATE, (1.3, 0, 0.65))],
])
def test_instances(orc, ref, ops):
    for shape_kind in ("mesh", "sphere", "box"):
        sd = abi.SceneDesc()
        m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
        sd.add(sd.rect(1, -4, 4, -4, 4, -0.5, m))
        if shape_kind == "mesh":
            pos, nrm, uv = scenes.blob_mesh(12, 12, 1.0, 0.2, seed=3, center=(0, 0.5, 0))
            sh = sd.mesh(pos, nrm, uv, m)
        elif shape_kind == "sphere":
            sh = sd.sphere((0.2, 0.5, -0.1), 0.8, m)
        else:
            sh = sd.box((0, 0.5, 0), (1.0, 1.5, 0.8), m)
        sd.add(sh, ops)
        o, r = both(sd, orc, ref)
        rays = scenes.random_rays(20000, (0, 0.5, 0), 3.5, seed=5)
        rays[:5000, 4:7] *= 3.7                      # unnormalised directions: RotateQuat changes the units of t
        assert_trace_equal(o.trace(rays), r.trace(rays))


def test_media(orc, ref):
    sd, cam, _ = scenes.cornell_smoke_scene(glass_sphere=True)
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add_medium(sd.sphere((300, 300, 300), 120, m), [], 0.02, (0.2, 0.4, 0.9))
    o, r = both(sd, orc, ref)
    rays = scenes.random_rays(30000, (278, 278, 278), 270, seed=6)
    rays[:10000, 4:7] *= 2.5
    ro, rr = o.trace(rays, rng_seed=77), r.trace(rays, rng_seed=77)
    nmedia = np.isin(ro["prim"], [o.prim_base(8), o.prim_base(9), o.prim_base(12)]).sum()
    assert nmedia > 500
    # u, v of a medium hit are never written by the reference (constantMedium.cpp:31-35): compare them only elsewhere
    med = np.isin(ro["prim"], [o.prim_base(8), o.prim_base(9), o.prim_base(12)])
    ro["uv"][med] = 0; rr["uv"][med] = 0
    assert_trace_equal(ro, rr)


def test_camera(orc, ref):
    cams = [dict(position=(0, 1, 3), look_at=(0, 1, 0), up=(0, 1, 0), vfov_deg=60, aperture=0.001, focal_distance=8.5),
            dict(position=(13, 2, 3), look_at=(0, 0, 0), up=(0, 1, 0), vfov_deg=20, aperture=0.1, focal_distance=10.0),
            dict(position=(278, 278, -800), look_at=(278, 278, 0), up=(0, 1, 0), vfov_deg=40, aperture=0.0, focal_distance=10.0)]
    st = np.random.default_rng(1).random((1000, 2)).astype(F) * 1.2 - 0.1
    for c in cams:
        for (w, h) in ((1920, 1080), (640, 640), (1200, 800)):
            cc = scenes.camera_for(c, w, h)
            a = np.empty((1000, 6), F); b = np.empty((1000, 6), F)
            orc.check(orc.fn("camera_rays")(ctypes.byref(cc), st.ctypes.data, 1000, a.ctypes.data))
            ref.check(ref.fn("camera_rays")(ctypes.byref(cc), st.ctypes.data, 1000, b.ctypes.data))
            assert same_bits(a, b)


def _hit_records(n, seed):
    rng = np.random.default_rng(seed)
    x = np.empty((n, 14), F)
    x[:, 0:3] = rng.normal(size=(n, 3)); d = rng.normal(size=(n, 3)); x[:, 3:6] = d * rng.uniform(0.5, 9, (n, 1))
    x[:, 6:9] = rng.normal(size=(n, 3)) * 2
    nn = rng.normal(size=(n, 3)); nn /= np.linalg.norm(nn, axis=1, keepdims=True)
    nn[: n // 3] *= rng.uniform(0.3, 2.0, (n // 3, 1))      # unnormalised mesh normals
    x[:, 9:12] = nn; x[:, 12:14] = rng.random((n, 2))
    front = (rng.random(n) < 0.5).astype(np.uint8)
    return x, front


def test_materials_and_textures(orc, ref, cache_dir):
    sd = abi.SceneDesc()
    t_even, t_odd = sd.solid((0.9, 0.9, 0.9)), sd.solid((0.2, 0.3, 0.1))
    t_chk = sd.checker(t_even, t_odd)
    img = np.random.default_rng(3).integers(0, 256, (33, 47, 3), dtype=np.uint8)
    t_img = sd.image(img)
    t_env = scenes.env_texture(sd, cache_dir, 64, 32)
    t_rough = sd.solid((0.1, 0.2, 0.05))
    mats = [
        sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.6, 0.7)),
        sd.material(abi.MAT_LAMBERTIAN, albedo_tex=t_chk),
        sd.material(abi.MAT_LAMBERTIAN, albedo_tex=t_img),
        sd.material(abi.MAT_METAL, (0.73, 0.73, 0.73), p0=0.2),
        sd.material(abi.MAT_METAL, (0.73, 0.73, 0.73), p0=0.0),
        sd.material(abi.MAT_METAL, albedo_tex=t_img, p0_tex=t_rough),
        sd.material(abi.MAT_METAL, (0.5, 0.5, 0.5), p0=-3.0),
        sd.material(abi.MAT_DIELECTRIC, p0=1.5, p1=0.0),
        sd.material(abi.MAT_DIELECTRIC, p0=1.5, p1=0.1),
        sd.material(abi.MAT_DIELECTRIC, p0=2.4, p1_tex=t_rough),
        sd.material(abi.MAT_ISOTROPIC, (0.3, 0.4, 0.5)),
        sd.material(abi.MAT_ISOTROPIC, albedo_tex=t_chk),
        sd.material(abi.MAT_DIFFUSE_LIGHT, (0.93, 0.75, 0.43), p0=4.5),
        sd.material(abi.MAT_DIFFUSE_LIGHT, albedo_tex=t_img, p0_tex=t_chk),
        sd.material(abi.MAT_PBR, (0.8, 0.3, 0.3), p0=0.3, p1=0.9),
        sd.material(abi.MAT_PBR, (0.8, 0.3, 0.3), p0=0.3, p1=0.1),
        sd.material(abi.MAT_UVTEST),
    ]
    sd.add(sd.sphere((0, 0, 0), 1, mats[0]))
    o, r = both(sd, orc, ref)
    uvp = np.random.default_rng(5).random((4000, 5)).astype(F) * 1.4 - 0.2
    uvp[:, 2:] *= 3
    for t in (t_even, t_chk, t_img, t_env):
        assert same_bits(o.texture_eval(t, uvp), r.texture_eval(t, uvp)), f"texture {t}"
    x, front = _hit_records(6000, 8)
    for m in mats:
        a, b = o.scatter(m, x, front, seed=21), r.scatter(m, x, front, seed=21)
        assert same_bits(a, b), f"material {m}: first diff row {np.nonzero((a != b).any(1))[0][:5]}"


def test_empty_image_texture_is_cyan(orc):
    sd = abi.SceneDesc()
    t = sd.image(np.zeros((0, 0, 3), np.uint8))
    sd.add(sd.sphere((0, 0, 0), 1, sd.material(abi.MAT_LAMBERTIAN, albedo_tex=t)))
    o = sd.build(orc)
    assert o.texture_eval(t, np.zeros((1, 5), F)).tolist() == [[0, 1, 1]]


RENDER_CASES = ["teapot", "shiny", "balls", "smoke", "shiny_glass"]


def _render_case(name, cache_dir):
    if name == "teapot":
        sd, cam, _ = scenes.teapot_scene(cache_dir, n=10, env_w=64, env_h=32)
    elif name == "shiny":
        sd, cam, _ = scenes.shiny_teapot_scene(cache_dir, n=10, env_w=64, env_h=32)
    elif name == "shiny_glass":
        sd, cam, _ = scenes.shiny_teapot_scene(cache_dir, n=10, env_w=64, env_h=32, rough_dielectric=True)
        # a bare dielectric mesh reads the never-written frontFace (UB in the reference): wrap it so it is defined
        kind, shape, ops = sd.objects[-1]
        sd.objects[-1] = (kind, shape, [(abi.XF_TRANSLATE, (0, 0, 0))])
    elif name == "balls":
        sd, cam, _ = scenes.scattered_balls_scene(grid=3)
    else:
        sd, cam, _ = scenes.cornell_smoke_scene(glass_sphere=True)
    return sd, cam


@pytest.mark.parametrize("name", RENDER_CASES)
def test_render_bit_exact(orc, ref, cache_dir, name):
    """Whole renders through the unmodified render()/rayColour() vs the restatement, same rand() stream."""
    sd, cam = _render_case(name, cache_dir)
    w, h, spp = 40, 30, 3
    o = sd.build(orc, 0, 4); r = sd.build(ref, 0, 4, tagged=False)
    c = scenes.camera_for(cam, w, h)
    a8, b8 = o.render_u8(c, w, h, spp, seed=3), r.render_u8(c, w, h, spp, seed=3)
    assert np.array_equal(a8, b8), f"{(a8 != b8).sum()} bytes differ"
    (af, ar), (bf, br) = o.render_f32(c, w, h, spp, seed=3, threads=2), r.render_f32(c, w, h, spp, seed=3, threads=3)
    assert ar == br and ar >= w * h * spp
    assert same_bits(af, bf)
    assert af.sum() > 0
