"""hrt_trace_batch_static: the ray batch whose scratch is a file-scope device array (no cudaMalloc inside the call).

It has no counterpart in the reference (include/hrt.h); the check is that it returns exactly what hrt_trace_batch_device
returns - bit-exact primitive ids and t, the tolerance of every ray-batch test here - on two calls in a row (the array keeps
the previous call's contents), on a scene that takes the wide-tree kernels (mesh: hit records + pending list + counter all in
the array) and on one that takes the plain kernel, and that sizes beyond the stated maximum are rejected with a clear error.
Collected last on purpose (file name).
"""
import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes
from test_hostsim_vs_oracle import mesh_scene, spheres_scene

pytestmark = pytest.mark.gpu
STATIC_MAX = 1 << 20          # HRT_TRACE_BATCH_STATIC_MAX (include/hrt.h)


def _both(hrt, sc, rays_h, torch):
    n = rays_h.shape[0]
    rays = torch.from_numpy(rays_h).cuda()
    out = {}
    for name in ("trace_batch_device", "trace_batch_static"):
        prim = torch.full((n,), 0x7fffffff, dtype=torch.int32, device="cuda")
        t = torch.full((n,), -7.0, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        hrt.check(hrt.fn(name)(sc.h, rays.data_ptr(), n, prim.data_ptr(), t.data_ptr(), 11, None), name)
        torch.cuda.synchronize()
        out[name] = (prim.cpu().numpy().view(np.uint32), t.cpu().numpy().view(np.uint32))
    return out["trace_batch_device"], out["trace_batch_static"]


def test_static_scratch_equals_allocating_entry(hrt):
    import torch
    sc_mesh = mesh_scene(56).build(hrt, abi.BVH_SAH, 9)            # wide trees: k_trace_batch + k_resolve_batch + k_split_hits
    sc_sph = spheres_scene().build(hrt, abi.BVH_REFERENCE, 5)      # no trees: k_trace_batch_plain + k_split_hits
    # two calls in a row on the same array, different sizes and scenes: the second must not see the first one's leftovers
    cases = ((sc_mesh, 300000, (0, 1, 0), 2.5, 4), (sc_mesh, 4097, (0, 1, 0), 2.5, 5), (sc_sph, 100000, (0, 0, -1), 3.0, 6),
             (sc_mesh, STATIC_MAX, (0, 1, 0), 2.5, 7), (sc_mesh, 1, (0, 1, 0), 2.5, 8))
    hits = 0
    for sc, n, center, rad, seed in cases:
        rays = scenes.random_rays(n, center, rad, seed=seed)
        (p0, t0), (p1, t1) = _both(hrt, sc, rays, torch)
        assert np.array_equal(p0, p1), f"n={n}: primitive ids differ at {np.nonzero(p0 != p1)[0][:8]}"
        assert np.array_equal(t0, t1), f"n={n}: t not bit-exact"
        assert (p1 != 0x7fffffff).all()                             # every output written
        hits += int((p1 != abi.MISS).sum())
    assert hits > 100000
    launches = sc_mesh.stats().launches
    assert launches > 0


def test_static_scratch_rejects_sizes_beyond_the_maximum(hrt):
    import torch
    sc = spheres_scene().build(hrt, abi.BVH_REFERENCE, 5)
    dummy = torch.zeros(64, dtype=torch.float32, device="cuda")
    rc = hrt.fn("trace_batch_static")(sc.h, dummy.data_ptr(), STATIC_MAX + 1, dummy.data_ptr(), dummy.data_ptr(), 0, None)
    assert rc == -1                                                  # HRT_ERR_ARG
    msg = hrt.fn("last_error")().decode()
    assert "HRT_TRACE_BATCH_STATIC_MAX" in msg and str(STATIC_MAX) in msg
    assert hrt.fn("trace_batch_static")(sc.h, None, 0, None, None, 0, None) == 0     # empty batch: nothing to do
