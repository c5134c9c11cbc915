"""HRT_BVH_LBVH_HOST — the host mirror of the device builder (hrt_bvh_gpu.cu) — checked on the CPU.

The device builder and this mirror share their rules as source (csrc/hrt_bvh_rules.h); tests/test_gpu_lbvh.py proves the
two produce identical arrays on the GPU box.  Here, without a GPU, the mirror itself is pinned: (1) its reference-topology
nodes equal the oracle's stable-order build node for node, and equal the reference's own std::sort topology whenever no two
keys of a range are equal; (2) hits through the Morton tree are bit-exact with the oracle; (3) the Morton tree is a proper
BVH over the leaf records (every leaf once, child boxes = exact unions); (4) a plain-Python restatement of the rules
(tree shape from the triangle count, Morton codes, Karras splits) reproduces the arrays.
"""
import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes
from test_hostsim_vs_oracle import hsim, bits, check_hits, mesh_scene, fallbacks  # noqa: F401

F = np.float32


def arrays(sc):
    ref = sc.mesh_array(0, 0).reshape(-1, 8)
    fast = sc.mesh_array(0, 1).reshape(-1, 16)
    leaves = sc.mesh_array(0, 2).reshape(-1, 32)
    return ref, fast, leaves


def ref_children(ref):
    return ref[:, 3].view(np.int32), ref[:, 7].view(np.int32)


def tie_mesh(n=20):
    """A flat grid: whole rows of triangles share box.min on every axis — the worst case for equal keys."""
    pos, nrm, uv = scenes.grid_mesh(n, n, lambda u, v: (np.stack([u * 2 - 1, 0 * u + 1.0, v * 2 - 1], -1), np.stack([0 * u, 0 * u + 1, 0 * u], -1)))
    return pos, nrm, uv


def scene_of(pos, nrm, uv):
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.mesh(pos, nrm, uv, m))
    return sd


def test_topology_equals_oracle_stable_build(orc, hsim):
    sd = mesh_scene(24)
    o = sd.build(orc, abi.BVH_LBVH_HOST, 9); h = sd.build(hsim, abi.BVH_LBVH_HOST, 9)
    nb = orc.fn("mesh_bvh_nodes")(o.h, 1, None, None)
    bo = np.zeros((nb, 6), F); co = np.zeros((nb, 2), np.int32)
    orc.fn("mesh_bvh_nodes")(o.h, 1, bo.ctypes.data, co.ctypes.data)
    ref, fast, leaves = arrays(h)
    l, r = ref_children(ref)
    assert ref.shape[0] == nb
    assert np.array_equal(co[:, 0], l) and np.array_equal(co[:, 1], r)
    assert np.array_equal(bits(bo[:, :3]), bits(ref[:, 0:3])) and np.array_equal(bits(bo[:, 3:]), bits(ref[:, 4:7]))


def test_stable_rule_equals_reference_rule_without_equal_keys(orc):
    """Jittered vertices: no two triangles of a range share box.min[axis], so std::sort (bvh.cpp:39) and the stable rule agree."""
    rng = np.random.default_rng(5)
    pos = rng.uniform(-1, 1, size=(3000, 3, 3)).astype(F)
    nrm = np.zeros_like(pos); nrm[..., 1] = 1; uv = np.zeros((3000, 3, 2), F)
    sd = scene_of(pos, nrm, uv)
    out = []
    for mode in (abi.BVH_REFERENCE, abi.BVH_LBVH_HOST):
        o = sd.build(orc, mode, 3)
        nb = orc.fn("mesh_bvh_nodes")(o.h, 0, None, None)
        b = np.zeros((nb, 6), F); c = np.zeros((nb, 2), np.int32)
        orc.fn("mesh_bvh_nodes")(o.h, 0, b.ctypes.data, c.ctypes.data)
        out.append((b, c))
    assert np.array_equal(out[0][1], out[1][1]) and np.array_equal(bits(out[0][0]), bits(out[1][0]))


@pytest.mark.parametrize("which", ["blob", "ties", "transformed"])
def test_hits_bit_exact_with_oracle(orc, hsim, which):
    if which == "blob":
        sd = mesh_scene(24); center, rad = (0, 1, 0), 2.5
    elif which == "ties":
        sd = scene_of(*tie_mesh(20)); center, rad = (0, 1, 0), 2.0
    else:
        sd = mesh_scene(20, ops=[(abi.XF_ROTATE_EULER_DEG, (10, 30, -20)), (abi.XF_SCALE, (1.5, 0.8, 1.2)), (abi.XF_TRANSLATE, (0.2, 0.1, -0.3))]); center, rad = (0, 1, 0), 3.0
    o = sd.build(orc, abi.BVH_LBVH_HOST, 9); h = sd.build(hsim, abi.BVH_LBVH_HOST, 9)
    rays = scenes.random_rays(40000, center, rad, seed=11)
    ro, rh = o.trace(rays), h.trace(rays)
    assert (ro["prim"] != abi.MISS).mean() > 0.1
    check_hits(ro, rh, front=False)
    hit = ro["prim"] != abi.MISS                                   # secondary rays: sub-t_min self hits (triangle.cpp:57-131)
    p = rays[hit, 0:3] + ro["t"][hit, None] * rays[hit, 4:7]
    rng = np.random.default_rng(0)
    sec = np.empty((p.shape[0], 8), F); sec[:, 0:3] = p; sec[:, 3] = 1e-3; sec[:, 7] = np.inf
    d = rng.normal(size=(p.shape[0], 3)); sec[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    check_hits(o.trace(sec), h.trace(sec), front=False)


def test_morton_tree_is_a_bvh_over_the_leaf_records(hsim):
    for sd in (mesh_scene(24, with_floor=False), scene_of(*tie_mesh(16))):
        h = sd.build(hsim, abi.BVH_LBVH_HOST, 9)
        ref, fast, leaves = arrays(h)
        n = leaves.shape[0]
        assert fast.shape[0] == n - 1
        l, r = ref_children(ref)
        assert (l < 0).sum() == n                                   # one leaf record per bottom node, in pre-order
        bottoms = np.nonzero(l < 0)[0]
        assert np.array_equal(bits(leaves[:, 0:3]), bits(ref[bottoms, 0:3])) and np.array_equal(bits(leaves[:, 4:7]), bits(ref[bottoms, 4:7]))
        assert np.array_equal(leaves[:, 3].view(np.int32), ~l[bottoms]) and np.array_equal(leaves[:, 7].view(np.int32), ~r[bottoms])
        seen = np.zeros(n, np.int32)

        def box_of(c):                                              # returns (lo[3], hi[3]) of child reference c, checking the stored boxes
            if c < 0:
                seen[~c] += 1
                return leaves[~c, 0:3], leaves[~c, 4:7]
            nd = fast[c]; c0, c1 = nd[12:14].view(np.int32)
            lo0, hi0 = box_of(int(c0)); lo1, hi1 = box_of(int(c1))
            assert np.array_equal(bits(nd[[0, 2, 8]]), bits(lo0)) and np.array_equal(bits(nd[[1, 3, 9]]), bits(hi0)), f"node {c} child 0 box"
            assert np.array_equal(bits(nd[[4, 6, 10]]), bits(lo1)) and np.array_equal(bits(nd[[5, 7, 11]]), bits(hi1)), f"node {c} child 1 box"
            return np.minimum(lo0, lo1), np.maximum(hi0, hi1)

        import sys
        sys.setrecursionlimit(10000)
        box_of(0)
        assert (seen == 1).all()


# ---- plain-Python restatement of csrc/hrt_bvh_rules.h -------------------------------------------------------------

def py_tree_shape(n_tri):
    """(start, n, node id) of every range of the median-split tree over n_tri items, in pre-order."""
    out = []

    def nodes(m):
        return 1 if m <= 2 else 1 + nodes(m // 2) + nodes(m - m // 2)

    def walk(start, m, self_):
        out.append((start, m, self_))
        if m > 2:
            walk(start, m // 2, self_ + 1)
            walk(start + m // 2, m - m // 2, self_ + 1 + nodes(m // 2))
    walk(0, n_tri, 0)
    return out


def py_float_key(x):
    b = np.asarray(x, F).view(np.uint32).astype(np.uint64)
    b = np.where((b << np.uint64(33)) == 0, np.uint64(0), b)       # -0 -> +0 (32-bit shift of the sign out)
    return np.where(b & np.uint64(0x80000000), (~b) & np.uint64(0xffffffff), b | np.uint64(0x80000000)).astype(np.uint32)


def py_morton(leaves):
    c = (F(0.5) * leaves[:, 0:3] + F(0.5) * leaves[:, 4:7]).astype(F)
    lo, hi = c.min(0), c.max(0)
    codes = np.zeros(len(c), np.uint64)
    for a in range(3):
        scale = F(2097152.0) / F(hi[a] - lo[a]) if hi[a] > lo[a] else F(0)
        f = ((c[:, a] - lo[a]).astype(F) * scale).astype(F)
        cell = np.where(f > 0, np.minimum(f, F(2097151.0)), 0).astype(np.uint64)
        spread = np.zeros(len(c), np.uint64)
        for bit in range(21):
            spread |= ((cell >> np.uint64(bit)) & np.uint64(1)) << np.uint64(3 * bit)
        codes |= spread << np.uint64(2 - a)
    return codes


def py_delta(codes, i, j):
    n = len(codes)
    if j < 0 or j >= n:
        return -1
    a, b = int(codes[i]), int(codes[j])
    if a != b:
        return 64 - (a ^ b).bit_length()
    return 64 + 64 - (i ^ j).bit_length()


def py_karras(codes, i):
    d = 1 if py_delta(codes, i, i + 1) - py_delta(codes, i, i - 1) >= 0 else -1
    dmin = py_delta(codes, i, i - d)
    lmax = 2
    while py_delta(codes, i, i + lmax * d) > dmin:
        lmax *= 2
    l, t = 0, lmax // 2
    while t >= 1:
        if py_delta(codes, i, i + (l + t) * d) > dmin:
            l += t
        t //= 2
    j = i + l * d
    dn = py_delta(codes, i, j)
    s, div = 0, 2
    while True:
        t = (l + div - 1) // div
        if py_delta(codes, i, i + (s + t) * d) > dn:
            s += t
        if t <= 1:
            break
        div *= 2
    split = i + s * d + min(d, 0)
    return min(i, j), max(i, j), split


def test_python_restatement_of_the_rules(hsim):
    sd = mesh_scene(14, with_floor=False)
    h = sd.build(hsim, abi.BVH_LBVH_HOST, 9)
    ref, fast, leaves = arrays(h)
    n_tri = sd.shapes[0][1].shape[0]
    l, r = ref_children(ref)
    shape = py_tree_shape(n_tri)
    assert len(shape) == ref.shape[0]
    for start, m, self_ in shape:                                  # internal nodes: children ids follow from the count alone
        if m > 2:
            assert l[self_] == self_ + 1 and r[self_] > self_ + 1
        else:
            assert l[self_] < 0 and r[self_] < 0
    codes = py_morton(leaves)
    order = np.argsort(codes, kind="stable")
    sorted_codes = codes[order]
    n = len(codes)
    for i in range(n - 1):
        first, last, split = py_karras(sorted_codes, i)
        c0, c1 = fast[i, 12:14].view(np.int32)
        assert c0 == (~int(order[split]) if split == first else split), f"node {i} left"
        assert c1 == (~int(order[split + 1]) if split + 1 == last else split + 1), f"node {i} right"
    # the ordering key: float order == key order, -0 == +0
    x = np.array([-np.inf, -3.5, -1e-40, -0.0, 0.0, 1e-40, 2.0, np.inf], F)
    k = py_float_key(x)
    assert k[3] == k[4] and (np.diff(k.astype(np.int64))[[0, 1, 2, 4, 5, 6]] > 0).all()
