"""The shading half of the path on the GPU against the oracle, record by record (SURVEY.md §4 test plan (iii), §8a rows 20-27):
Texture::colourValue through hrt_texture_eval — solid, checkered, image, environment (float and RGBE-repacked) — bit for bit, and
Material::scatter / emitted through hrt_scatter_batch for every material kind incl. texture-valued MatVec3 / MatScalar, PBR and UVTest:
what does not depend on random numbers (attenuation, emitted, the scatter verdict of non-metals) bit for bit, the rest by moments
over 4 x 10^5 draws (the device draws from Philox, the oracle from the reference's rand() construction: same distributions)."""
import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes

pytestmark = pytest.mark.gpu
F = np.float32
N = 400_000


def bits(a):
    a = np.ascontiguousarray(a, F); b = a.view(np.uint32).copy(); b[np.isnan(a)] = 0x7FC00000; return b


def textured_scene():
    rng = np.random.default_rng(4)
    sd = abi.SceneDesc()
    t = {}
    t["red"] = sd.solid((0.65, 0.05, 0.05)); t["white"] = sd.solid((0.9, 0.9, 0.9))
    t["checker"] = sd.checker(t["red"], t["white"])
    t["nested"] = sd.checker(t["checker"], sd.solid((0.1, 0.2, 0.3)))
    t["photo"] = sd.image(rng.integers(0, 256, (37, 53, 3), dtype=np.uint8))
    t["empty"] = sd.image(np.zeros((0, 0, 3), np.uint8))                                 # a file that failed to load: cyan (texture.cpp:57)
    t["grey"] = sd.image(np.full((8, 8, 3), 60, np.uint8))                               # length(rgb) ~ 0.41: a roughness / strength map
    rgbe = rng.integers(0, 256, (32, 64, 4), dtype=np.uint8); rgbe[..., 3] = rng.integers(110, 150, (32, 64))
    t["env_rgbe"] = sd.env(scenes.rgbe_to_float(rgbe))
    t["env_f32"] = sd.env((rng.random((16, 32, 3)) * 3).astype(F))
    m = {}
    m["lambert"] = sd.material(abi.MAT_LAMBERTIAN, (0.7, 0.6, 0.5))
    m["lambert_checker"] = sd.material(abi.MAT_LAMBERTIAN, albedo_tex=t["checker"])
    m["lambert_photo"] = sd.material(abi.MAT_LAMBERTIAN, albedo_tex=t["photo"])
    m["metal0"] = sd.material(abi.MAT_METAL, (0.8, 0.8, 0.9), p0=0.0)
    m["metal_rough"] = sd.material(abi.MAT_METAL, (0.8, 0.8, 0.9), p0=0.35)
    m["metal_over"] = sd.material(abi.MAT_METAL, (0.8, 0.8, 0.9), p0=-3.0)               # min(abs(r), 1) (material.h:169-170)
    m["metal_tex"] = sd.material(abi.MAT_METAL, albedo_tex=t["photo"], p0_tex=t["grey"])
    m["glass"] = sd.material(abi.MAT_DIELECTRIC, p0=1.5, p1=0.0)
    m["glass_rough"] = sd.material(abi.MAT_DIELECTRIC, p0=1.33, p1=0.1)
    m["glass_tex"] = sd.material(abi.MAT_DIELECTRIC, p0=1.5, p1_tex=t["grey"])
    m["iso"] = sd.material(abi.MAT_ISOTROPIC, albedo_tex=t["checker"])
    m["light"] = sd.material(abi.MAT_DIFFUSE_LIGHT, (0.9, 0.8, 0.7), p0=4.5)
    m["light_tex"] = sd.material(abi.MAT_DIFFUSE_LIGHT, albedo_tex=t["photo"], p0_tex=t["grey"])
    m["pbr_metal"] = sd.material(abi.MAT_PBR, (0.7, 0.7, 0.2), p0=0.3, p1=0.9)
    m["pbr_diffuse"] = sd.material(abi.MAT_PBR, (0.7, 0.7, 0.2), p0=0.3, p1=0.1)
    m["pbr_tex"] = sd.material(abi.MAT_PBR, albedo_tex=t["checker"], p0=0.3, p1_tex=t["photo"])
    m["uvtest"] = sd.material(abi.MAT_UVTEST)
    sd.add(sd.sphere((0, 0, 0), 1.0, m["lambert"]))
    sd.background = t["env_rgbe"]
    return sd, t, m


@pytest.fixture(scope="module")
def pair(hrt, orc):
    sd, t, m = textured_scene()
    return sd.build(hrt, abi.BVH_SAH, 1), sd.build(orc, abi.BVH_REFERENCE, 1), t, m


def test_textures_bit_exact(pair):
    g, o, t, _ = pair
    rng = np.random.default_rng(9)
    uvp = np.empty((200_000, 5), F)
    uvp[:, 0:2] = rng.random((len(uvp), 2)) * 1.3 - 0.15                                 # beyond [0, 1]: the clamps
    uvp[:, 2:5] = rng.normal(size=(len(uvp), 3)) * 2.0
    uvp[:64, 0:2] = [[0, 0], [1, 1], [0, 1], [1, 0]] * 16                                # corners: int(u * W) == W
    for name, tex in t.items():
        a, b = g.texture_eval(tex, uvp), o.texture_eval(tex, uvp)
        if name in ("checker", "nested"):                                                # sign of sinf products: libdevice vs libm within 1 ulp of a zero
            assert (bits(a) != bits(b)).any(axis=1).mean() < 1e-4, name
        else:
            assert np.array_equal(bits(a), bits(b)), name


def hit_records(n, seed):
    """Random incoming rays, hit points, unnormalised normals (meshes leave them so) facing the ray or not, uv."""
    rng = np.random.default_rng(seed)
    x = np.empty((n, 14), F)
    x[:, 0:3] = rng.normal(size=(n, 3))
    d = rng.normal(size=(n, 3)); x[:, 3:6] = d * rng.uniform(0.2, 3.0, (n, 1))
    x[:, 6:9] = rng.normal(size=(n, 3)) * 2
    nn = rng.normal(size=(n, 3)); nn /= np.linalg.norm(nn, axis=1, keepdims=True)
    flip = (np.einsum("ij,ij->i", nn, d) > 0)
    nn[flip] *= -1                                                                       # setFaceNormal: the record's normal opposes the ray
    x[:, 9:12] = nn * rng.uniform(0.5, 2.0, (n, 1))
    x[:, 12:14] = rng.random((n, 2))
    front = rng.integers(0, 2, n).astype(np.uint8)
    return x, front


KINDS = ["lambert", "lambert_checker", "lambert_photo", "metal0", "metal_rough", "metal_over", "metal_tex", "glass", "glass_rough", "glass_tex",
         "iso", "light", "light_tex", "pbr_metal", "pbr_diffuse", "pbr_tex", "uvtest"]


@pytest.mark.parametrize("kind", KINDS)
def test_scatter_moments(pair, kind):
    g, o, _, m = pair
    x, front = hit_records(N, 17)
    a = g.scatter(m[kind], x, front, seed=5)
    b = o.scatter(m[kind], x, front, seed=5)
    ok_a, ok_b = a[:, 0] > 0, b[:, 0] > 0
    assert np.array_equal(bits(a[:, 7:10]), bits(b[:, 7:10])), "emitted"                 # no random numbers in emitted()
    if kind.startswith("light"):
        assert not ok_a.any() and not ok_b.any()
        return
    # the verdict: always true except for metals (material.h:176), where it follows the random lobe
    sig = np.sqrt(max(ok_b.mean() * (1 - ok_b.mean()), 1e-6) / N)
    assert abs(ok_a.mean() - ok_b.mean()) < 5 * sig + 1e-4, (ok_a.mean(), ok_b.mean())
    if not (kind.startswith("metal") or kind.startswith("pbr")):
        assert ok_a.all() and ok_b.all()
    # attenuation never depends on the draw, except through PBR's metal / diffuse choice (a texture lookup, not a draw): bit for bit
    both = ok_a & ok_b if kind.startswith("metal") or kind.startswith("pbr") else slice(None)
    if kind in ("lambert_checker", "iso", "pbr_tex"):
        assert (bits(a[both, 1:4]) != bits(b[both, 1:4])).any(axis=1).mean() < 1e-4
    else:
        assert np.array_equal(bits(a[both, 1:4]), bits(b[both, 1:4])), "attenuation"
    # scattered direction: first and second moments in the frame of the record (mean, mean of d.n, mean |d|, spread)
    da, db = a[ok_a, 4:7].astype(np.float64), b[ok_b, 4:7].astype(np.float64)
    na = x[ok_a, 9:12].astype(np.float64); nb = x[ok_b, 9:12].astype(np.float64)
    tol = 6.0 / np.sqrt(min(len(da), len(db))) * max(1.0, float(np.abs(db).mean()))
    assert np.abs(da.mean(0) - db.mean(0)).max() < tol, (da.mean(0), db.mean(0))
    ca = np.einsum("ij,ij->i", da, na) / np.linalg.norm(na, axis=1); cb = np.einsum("ij,ij->i", db, nb) / np.linalg.norm(nb, axis=1)
    assert abs(ca.mean() - cb.mean()) < tol and abs(ca.std() - cb.std()) < tol
    la, lb = np.linalg.norm(da, axis=1), np.linalg.norm(db, axis=1)
    assert abs(la.mean() - lb.mean()) < tol and abs(la.std() - lb.std()) < tol
    assert abs((da ** 2).mean() - (db ** 2).mean()) < 2 * tol * max(1.0, float((db ** 2).mean()))


def test_scatter_is_the_render_stream(pair):
    """hrt_scatter_batch draws what a path draws at bounce 0: same seed -> same records; another seed -> other directions."""
    g, _, _, m = pair
    x, front = hit_records(10_000, 3)
    a, b, c = g.scatter(m["lambert"], x, front, seed=5), g.scatter(m["lambert"], x, front, seed=5), g.scatter(m["lambert"], x, front, seed=6)
    assert np.array_equal(bits(a), bits(b)) and (bits(a[:, 4:7]) != bits(c[:, 4:7])).any(axis=1).mean() > 0.99
    u = g.scatter(m["lambert"], x, front, seed=5, rng_mode=1)                            # plain uniform instead of the base-255 digits
    assert (bits(a[:, 4:7]) != bits(u[:, 4:7])).any(axis=1).mean() > 0.9
