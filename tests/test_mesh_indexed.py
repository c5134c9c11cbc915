"""hrt_scene_add_mesh_indexed — the mesh ingest fast path (mesh.cpp:13-41, 53-120).

The entry point takes the vertex / index buffers as Mesh::assimpLoadFile concatenates them and does Mesh::Mesh's gather
(mesh.cpp:24-33) inside the library.  Checked here on the CPU through the host half of the product (tests/hostsim links
csrc/hrt_builder.cpp + hrt_host.cpp): (1) every per-mesh array equals the one hrt_scene_add_mesh produces from the
flattened triangles, in all build modes; (2) hits agree bit for bit with the oracle, which is fed the flattened triangles;
(3) without a sub-mesh table the indices are used as they are - the reference's behaviour for files with several aiMeshes
(mesh.cpp:111-114) - and with a table every sub-mesh is offset by its base vertex; (4) the inputs on which the reference
reads out of bounds return HRT_ERR_ARG with a message.  tests/test_gpu_lbvh.py repeats (1)-(2) for the device gather.
"""
import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes
from test_hostsim_vs_oracle import hsim, check_hits  # noqa: F401

F = np.float32
NAMES = ["ref_nodes", "fast_nodes", "leaves", "tris", "attrs"]


def indexed_scene(v, n, t, idx, sub=None, flatten=False, borrow=False):
    sd = abi.SceneDesc(); sd.flatten_for_all = flatten; sd.borrow = borrow
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.mesh_indexed(v, n, t, idx, m, sub))
    return sd


def same_arrays(a, b, names=NAMES):
    for which, name in enumerate(NAMES):
        if name not in names:
            continue
        x, y = a.mesh_array(0, which), b.mesh_array(0, which)
        assert x.shape == y.shape, f"{name}: {x.shape} vs {y.shape}"
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), name


def test_generator_flattens_to_the_soup():
    v, n, t, idx = scenes.blob_mesh(9, 7, 1.0, 0.2, seed=3, indexed=True)
    pos, nrm, uv = scenes.blob_mesh(9, 7, 1.0, 0.2, seed=3)
    fp, fn, ft = abi.SceneDesc.flatten_indexed(v, n, t, idx, None)
    assert np.array_equal(fp, pos) and np.array_equal(fn, nrm) and np.array_equal(ft, uv)


@pytest.mark.parametrize("mode", [abi.BVH_REFERENCE, abi.BVH_SAH, abi.BVH_LBVH_HOST])
def test_arrays_equal_flattened_mesh(hsim, mode):
    v, n, t, idx = scenes.blob_mesh(20, 20, 1.0, 0.2, seed=7, center=(0, 1, 0), indexed=True)
    a = indexed_scene(v, n, t, idx).build(hsim, mode, 5)
    b = indexed_scene(v, n, t, idx, flatten=True).build(hsim, mode, 5)
    names = NAMES if mode != abi.BVH_REFERENCE else ["ref_nodes", "tris", "attrs"]
    same_arrays(a, b, names)
    assert a.prim_count() == b.prim_count() == idx.shape[0] // 3


@pytest.mark.parametrize("mode", [abi.BVH_REFERENCE, abi.BVH_SAH])
def test_hits_bit_exact_with_oracle(orc, hsim, mode):
    v, n, t, idx = scenes.blob_mesh(24, 24, 1.0, 0.2, seed=7, center=(0, 1, 0), indexed=True)
    sd = indexed_scene(v, n, t, idx)
    o = sd.build(orc, abi.BVH_REFERENCE, 5); h = sd.build(hsim, mode, 5)     # the oracle has no indexed entry: numpy flattens
    rays = scenes.random_rays(40000, (0, 1, 0), 2.5, seed=4)
    ro, rh = o.trace(rays), h.trace(rays)
    assert (ro["prim"] != abi.MISS).mean() > 0.1
    check_hits(ro, rh, front=False)


def test_missing_normals_and_uvs_are_zero(orc, hsim):
    v, _, _, idx = scenes.blob_mesh(10, 10, 1.0, 0.2, seed=2, center=(0, 1, 0), indexed=True)
    sd = indexed_scene(v, None, None, idx)
    o = sd.build(orc, abi.BVH_REFERENCE, 5); h = sd.build(hsim, abi.BVH_SAH, 5)
    rays = scenes.random_rays(20000, (0, 1, 0), 2.5, seed=6)
    ro, rh = o.trace(rays), h.trace(rays)
    check_hits(ro, rh, front=False)
    hit = rh["prim"] != abi.MISS
    assert hit.any() and not rh["uv"][hit].any()                            # mesh.cpp:96-99: uv (0, 0) everywhere


def two_submeshes():
    """Two aiMesh-like pieces concatenated the way Mesh::assimpLoadFile does: vertices appended, indices local to each piece."""
    va, na, ta, ia = scenes.blob_mesh(8, 8, 0.6, 0.2, seed=1, center=(-0.8, 1, 0), indexed=True)
    vb, nb, tb, ib = scenes.blob_mesh(6, 10, 0.5, 0.2, seed=2, center=(0.9, 1, 0), indexed=True)
    v = np.concatenate([va, vb]); n = np.concatenate([na, nb]); t = np.concatenate([ta, tb]); idx = np.concatenate([ia, ib])
    sub = [(0, ia.shape[0], 0), (ia.shape[0], ib.shape[0], va.shape[0])]
    return v, n, t, idx, sub, (va, vb, ia, ib)


def test_submesh_table_applies_the_base_vertex(orc, hsim):
    v, n, t, idx, sub, (va, vb, ia, ib) = two_submeshes()
    fixed = indexed_scene(v, n, t, idx, sub).build(hsim, abi.BVH_SAH, 5)
    tris = fixed.mesh_array(0, 3).reshape(-1, 3, 4)[:, :, :3]
    want = np.concatenate([va[ia.astype(np.int64)], vb[ib.astype(np.int64)]]).reshape(-1, 3, 3)
    assert np.array_equal(tris, want)
    o = indexed_scene(v, n, t, idx, sub).build(orc, abi.BVH_REFERENCE, 5)
    rays = scenes.random_rays(30000, (0, 1, 0), 2.5, seed=8)
    check_hits(o.trace(rays), fixed.trace(rays), front=False)


def test_without_a_table_indices_are_used_as_they_are(orc, hsim):
    """The reference's behaviour on multi-aiMesh files (mesh.cpp:111-114): the second piece's faces index the first piece's vertices."""
    v, n, t, idx, _, (va, vb, ia, ib) = two_submeshes()
    asis = indexed_scene(v, n, t, idx).build(hsim, abi.BVH_REFERENCE, 5)
    tris = asis.mesh_array(0, 3).reshape(-1, 3, 4)[:, :, :3]
    assert np.array_equal(tris, v[idx.astype(np.int64)].reshape(-1, 3, 3))
    o = indexed_scene(v, n, t, idx).build(orc, abi.BVH_REFERENCE, 5)
    rays = scenes.random_rays(30000, (0, 1, 0), 2.5, seed=9)
    check_hits(o.trace(rays), asis.trace(rays), front=False)


@pytest.mark.parametrize("mode", [abi.BVH_SAH, abi.BVH_LBVH_HOST])
def test_borrowed_arrays_give_the_same_scene(hsim, mode):
    """HRT_MESH_BORROW: arrays read in place until the commit returns, then never touched again."""
    v, n, t, idx, sub, _ = two_submeshes()
    rays = scenes.random_rays(20000, (0, 1, 0), 2.5, seed=3)
    for table in (None, sub):
        a = indexed_scene(v, n, t, idx, table).build(hsim, mode, 5)
        vv, nn, tt, ii = v.copy(), n.copy(), t.copy(), idx.copy()
        b = indexed_scene(vv, nn, tt, ii, table, borrow=True).build(hsim, mode, 5)
        vv[:] = np.nan; nn[:] = np.nan; tt[:] = np.nan; ii[:] = 0xFFFFFFFF           # after the commit the caller owns them again
        same_arrays(a, b)
        ra, rb = a.trace(rays), b.trace(rays)
        check_hits(ra, rb, front=False)


def test_unknown_flags_are_an_error(hsim):
    v, n, t, idx = scenes.blob_mesh(4, 4, 1.0, 0.2, seed=1, indexed=True)
    h = hsim.fn("scene_create")()
    m = abi.Material(); m.kind = 0; m.albedo_tex = m.p0_tex = m.p1_tex = -1
    assert hsim.fn("scene_add_material")(h, abi.C.byref(m)) == 0
    rc = hsim.fn("scene_add_mesh_indexed")(h, v.ctypes.data, n.ctypes.data, t.ctypes.data, v.shape[0], idx.ctypes.data, idx.shape[0], None, 0, 0, 2)
    assert rc == -1 and b"flags" in hsim.fn("last_error")()
    hsim.fn("scene_destroy")(h)


def test_empty_mesh(hsim):
    sd = indexed_scene(np.zeros((0, 3), F), None, None, np.zeros(0, np.uint32))
    sd.add(sd.sphere((0, 0, -3), 1.0, 0))
    h = sd.build(hsim, abi.BVH_SAH, 5)
    rays = scenes.random_rays(1000, (0, 0, -3), 2.0, seed=1)
    r = h.trace(rays)
    assert set(np.unique(r["prim"])) <= {0, abi.MISS}


@pytest.mark.parametrize("case", ["not_multiple_of_3", "index_out_of_range", "overlap", "uncovered", "base_out_of_range"])
def test_bad_input_is_an_error(hsim, case):
    v, n, t, idx = scenes.blob_mesh(4, 4, 1.0, 0.2, seed=1, indexed=True)
    sub = None
    if case == "not_multiple_of_3":
        idx = idx[:-1]
    elif case == "index_out_of_range":
        idx = idx.copy(); idx[5] = v.shape[0]
    elif case == "overlap":
        sub = [(0, 48, 0), (45, idx.shape[0] - 45, 0)]
    elif case == "uncovered":
        sub = [(0, 48, 0)]
    else:
        sub = [(0, idx.shape[0], 1)]                       # the largest index + 1 leaves the vertex array
    with pytest.raises(abi.HrtError):
        indexed_scene(v, n, t, idx, sub).build(hsim, abi.BVH_SAH, 5)
