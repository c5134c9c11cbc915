"""HRT_BVH_LBVH_GPU: the per-mesh build of hrt_scene_commit done on the device (csrc/hrt_bvh_gpu.cu).

(1) Every array the device builder produces — reference-topology nodes, Morton-tree nodes, leaf records, packed triangles
and attributes — is BYTE-identical to the host mirror's (HRT_BVH_LBVH_HOST, pinned on the CPU by tests/test_lbvh_host.py);
(2) hits through the device-built trees are bit-exact with the oracle; (3) films rendered over the device-built and the
host-built trees are identical; (4) inputs the device builder hands back to the host (non-finite positions) still commit and
still agree with the oracle; (5) at 2 M triangles (beyond what the oracle traces in seconds) the same byte-equality holds and
the ray batch agrees with the SAH build.
"""
import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes
from test_hostsim_vs_oracle import bits, mesh_scene
from test_lbvh_host import scene_of, tie_mesh
from test_gpu_parity import check

pytestmark = pytest.mark.gpu
F = np.float32
NAMES = ["ref_nodes", "fast_nodes", "leaves", "tris", "attrs", "wide_nodes"]


def same_arrays(a, b):
    for which, name in enumerate(NAMES):
        x, y = a.mesh_array(0, which), b.mesh_array(0, which)
        assert x.shape == y.shape, f"{name}: {x.shape} vs {y.shape}"
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), f"{name} differs at words {np.nonzero(x.view(np.uint32) != y.view(np.uint32))[0][:8]}"


def random_soup(n, seed):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-1, 1, size=(n, 1, 3)); pos = (c + rng.normal(scale=0.05, size=(n, 3, 3))).astype(F)
    nrm = rng.normal(size=(n, 3, 3)).astype(F); uv = rng.uniform(size=(n, 3, 2)).astype(F)
    return pos, nrm, uv


@pytest.mark.parametrize("n_tri", [1, 2, 3, 4, 5, 7, 33, 1000, 4097])
def test_arrays_equal_host_mirror_small_counts(hrt, n_tri):
    sd = scene_of(*random_soup(n_tri, n_tri))
    same_arrays(sd.build(hrt, abi.BVH_LBVH_GPU, 7), sd.build(hrt, abi.BVH_LBVH_HOST, 7))


@pytest.mark.parametrize("which", ["blob", "ties", "duplicates"])
def test_arrays_equal_host_mirror(hrt, which):
    if which == "blob":
        sd = mesh_scene(56, with_floor=False)
    elif which == "ties":
        sd = scene_of(*tie_mesh(64))
    else:                                                  # every triangle eight times: equal keys AND equal Morton codes
        pos, nrm, uv = random_soup(500, 3)
        sd = scene_of(np.tile(pos, (8, 1, 1)), np.tile(nrm, (8, 1, 1)), np.tile(uv, (8, 1, 1)))
    for seed in (1, 9):
        same_arrays(sd.build(hrt, abi.BVH_LBVH_GPU, seed), sd.build(hrt, abi.BVH_LBVH_HOST, seed))


@pytest.mark.parametrize("which", ["blob", "ties", "transformed"])
def test_hits_bit_exact_with_oracle(orc, hrt, which):
    if which == "blob":
        sd = mesh_scene(56); center, rad = (0, 1, 0), 2.5
    elif which == "ties":
        sd = scene_of(*tie_mesh(40)); center, rad = (0, 1, 0), 2.0
    else:
        sd = mesh_scene(40, ops=[(abi.XF_ROTATE_EULER_DEG, (10, 30, -20)), (abi.XF_SCALE, (1.5, 0.8, 1.2)), (abi.XF_TRANSLATE, (0.2, 0.1, -0.3))]); center, rad = (0, 1, 0), 3.0
    o = sd.build(orc, abi.BVH_LBVH_GPU, 9); g = sd.build(hrt, abi.BVH_LBVH_GPU, 9)
    rays = scenes.random_rays(300000, center, rad, seed=11)
    ro, rg = o.trace(rays), g.trace(rays)
    assert (ro["prim"] != abi.MISS).mean() > 0.1
    check(ro, rg, front=False)
    hit = ro["prim"] != abi.MISS
    p = rays[hit, 0:3] + ro["t"][hit, None] * rays[hit, 4:7]
    rng = np.random.default_rng(0)
    sec = np.empty((p.shape[0], 8), F); sec[:, 0:3] = p; sec[:, 3] = 1e-3; sec[:, 7] = np.inf
    d = rng.normal(size=(p.shape[0], 3)); sec[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    check(o.trace(sec), g.trace(sec), front=False)


def test_render_identical_over_device_and_host_built_trees(hrt, cache_dir):
    sd, cam, _ = scenes.teapot_scene(cache_dir, n=32, env_w=128, env_h=64)
    c = scenes.camera_for(cam, 160, 160)
    a = sd.build(hrt, abi.BVH_LBVH_GPU, 3).render(c, 160, 160, 16, seed=5)
    b = sd.build(hrt, abi.BVH_LBVH_HOST, 3).render(c, 160, 160, 16, seed=5)
    s = sd.build(hrt, abi.BVH_SAH, 3).render(c, 160, 160, 16, seed=5)
    assert np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32)) and np.array_equal(a[1], b[1])
    # against the SAH build the image may differ only where std::sort and the stable rule order equal keys differently
    assert np.mean(a[1] != s[1]) < 1e-3


def test_non_finite_positions_fall_back_to_the_host_builder(orc, hrt):
    pos, nrm, uv = random_soup(2000, 4)
    pos[17, 1, 2] = np.nan; pos[900, 0, 0] = np.inf
    sd = scene_of(pos, nrm, uv)
    g = sd.build(hrt, abi.BVH_LBVH_GPU, 5); o = sd.build(orc, abi.BVH_REFERENCE, 5)   # the fallback is the SAH path (std::sort order)
    rays = scenes.random_rays(100000, (0, 0, 0), 1.5, seed=2)
    ro, rg = o.trace(rays, extras=False), g.trace(rays, extras=False)
    assert np.array_equal(ro["prim"], rg["prim"]) and np.array_equal(bits(ro["t"]), bits(rg["t"]))


def test_two_million_triangles(hrt):
    pos, nrm, uv = scenes.blob_mesh(1000, 1000, 1.0, 0.15, seed=7)
    sd = scene_of(pos, nrm, uv)
    g = sd.build(hrt, abi.BVH_LBVH_GPU, 3)
    st = g.stats()
    h = sd.build(hrt, abi.BVH_LBVH_HOST, 3)
    same_arrays(g, h)
    # the build proper (the upload part also holds cudaMalloc of the scratch and a pageable-memory copy: 25 ms alone, 150-900 ms late in a long process)
    dev_ms = st.ms_build - st.ms_build_upload
    assert dev_ms < h.stats().ms_build, f"device build {dev_ms:.1f} ms vs host mirror {h.stats().ms_build:.1f} ms"
    rays = scenes.random_rays(1 << 20, (0, 0, 0), 1.5, seed=8)
    rg = g.trace(rays, extras=False); rs = sd.build(hrt, abi.BVH_SAH, 3).trace(rays, extras=False)
    assert (rg["prim"] != abi.MISS).mean() > 0.3
    same = (rg["prim"] == rs["prim"]) & (bits(rg["t"]) == bits(rs["t"]))
    assert same.mean() > 0.9999, f"{(~same).sum()} rays differ between the stable-order and the std::sort topology"


# ---- hrt_scene_add_mesh_indexed: the gather of mesh.cpp:24-33 done by the device builder (k_mesh_pack<true>) ----------
def indexed_scene(v, n, t, idx, sub=None, flatten=False, borrow=False):
    sd = abi.SceneDesc(); sd.flatten_for_all = flatten; sd.borrow = borrow
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.mesh_indexed(v, n, t, idx, m, sub))
    return sd


@pytest.mark.parametrize("mode", [abi.BVH_LBVH_GPU, abi.BVH_SAH, abi.BVH_REFERENCE])
def test_indexed_mesh_arrays_equal_flattened_mesh(hrt, mode):
    v, n, t, idx = scenes.blob_mesh(40, 40, 1.0, 0.2, seed=7, center=(0, 1, 0), indexed=True)
    a = indexed_scene(v, n, t, idx).build(hrt, mode, 5)
    b = indexed_scene(v, n, t, idx, flatten=True).build(hrt, mode, 5)
    if mode == abi.BVH_REFERENCE:
        for which in (0, 3, 4):
            assert np.array_equal(a.mesh_array(0, which).view(np.uint32), b.mesh_array(0, which).view(np.uint32))
    else:
        same_arrays(a, b)
    if mode == abi.BVH_LBVH_GPU:
        same_arrays(a, indexed_scene(v, n, t, idx).build(hrt, abi.BVH_LBVH_HOST, 5))
        assert a.stats().h2d_bytes < 0.5 * b.stats().h2d_bytes          # only the compact buffers cross PCIe


def test_indexed_mesh_hits_bit_exact_with_oracle(orc, hrt):
    va, na, ta, ia = scenes.blob_mesh(30, 30, 0.6, 0.2, seed=1, center=(-0.8, 1, 0), indexed=True)
    vb, nb, tb, ib = scenes.blob_mesh(20, 34, 0.5, 0.2, seed=2, center=(0.9, 1, 0), indexed=True)
    v = np.concatenate([va, vb]); n = np.concatenate([na, nb]); t = np.concatenate([ta, tb]); idx = np.concatenate([ia, ib])
    sub = [(0, ia.shape[0], 0), (ia.shape[0], ib.shape[0], va.shape[0])]
    rays = scenes.random_rays(200000, (0, 1, 0), 2.5, seed=8)
    for table in (sub, None):                       # corrected base-vertex offset / the reference's as-is indexing (mesh.cpp:111-114)
        sd = indexed_scene(v, n, t, idx, table)
        o = sd.build(orc, abi.BVH_LBVH_GPU, 5); g = sd.build(hrt, abi.BVH_LBVH_GPU, 5)
        ro = o.trace(rays)
        assert (ro["prim"] != abi.MISS).mean() > 0.02
        check(ro, g.trace(rays), front=False)
        check(sd.build(orc, abi.BVH_REFERENCE, 5).trace(rays), sd.build(hrt, abi.BVH_SAH, 5).trace(rays), front=False)


def test_indexed_two_million_triangles(hrt):
    v, n, t, idx = scenes.blob_mesh(1000, 1000, 1.0, 0.15, seed=7, indexed=True)
    a = indexed_scene(v, n, t, idx).build(hrt, abi.BVH_LBVH_GPU, 3)
    b = indexed_scene(v, n, t, idx, flatten=True).build(hrt, abi.BVH_LBVH_GPU, 3)
    same_arrays(a, b)
    vv, nn, tt, ii = v.copy(), n.copy(), t.copy(), idx.copy()
    c = indexed_scene(vv, nn, tt, ii, borrow=True).build(hrt, abi.BVH_LBVH_GPU, 3)      # HRT_MESH_BORROW: read in place until the commit returns
    vv[:] = np.nan; ii[:] = 0xFFFFFFFF
    same_arrays(a, c)
    sa, sb = a.stats(), b.stats()
    assert sa.h2d_bytes < 0.35 * sb.h2d_bytes, f"{sa.h2d_bytes} vs {sb.h2d_bytes} bytes host -> device"
    rays = scenes.random_rays(1 << 18, (0, 0, 0), 1.5, seed=8)
    ra, rb = a.trace(rays, extras=False), b.trace(rays, extras=False)
    assert np.array_equal(ra["prim"], rb["prim"]) and np.array_equal(bits(ra["t"]), bits(rb["t"]))
