#!/usr/bin/env python
"""Generates tests/golden/trace_bigmesh.npz: the first 2^18 rays of the 0xC0FFEE incoherent batch against the
10,488,200-triangle mesh of bench.py's big_mesh workload, answered by the UNMODIFIED reference (oracle/_ref/libhrt_ref.so:
ITriangle::hit triangle.cpp:57-131 under BVHNode::hit bvh.cpp:69-78, the tree built by bvh.cpp:6-61 on the seeded rand() stream).
Only the answers are stored (prim, t: 2 MiB): mesh and rays are regenerated from their seeds by the test.
Run in the build container (about 10 minutes, ~6 GB):  python tests/golden/make_golden_bigmesh.py"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from hobbyraytracer_b200 import abi, scenes  # noqa: E402

QUADS, N_RAYS, SEED = 2290, 1 << 18, 3


def inputs():
    sd, cam, info = scenes.big_mesh_scene(os.path.join(ROOT, "tests", "_cache"), quads=QUADS, env_w=64, env_h=32)
    rays = scenes.random_rays(N_RAYS, (0, 1, 0), 1.2 * 1.5, seed=0xC0FFEE)
    return sd, rays


if __name__ == "__main__":
    ref = abi.Lib(os.path.join(ROOT, "oracle", "_ref", "libhrt_ref.so"), "ref_")
    sd, rays = inputs()
    t0 = time.time(); sc = sd.build(ref, abi.BVH_REFERENCE, SEED); t1 = time.time()
    out = sc.trace(rays, extras=False); t2 = time.time()
    print(f"reference build {t1 - t0:.1f} s, trace {t2 - t1:.1f} s, hits {int((out['prim'] != abi.MISS).sum())} of {N_RAYS}, NaN t {int(np.isnan(out['t']).sum())}")
    np.savez_compressed(os.path.join(HERE, "trace_bigmesh.npz"), prim=out["prim"], t=out["t"], quads=QUADS, n_rays=N_RAYS, seed=SEED,
                        rays_crc=np.uint32(__import__("zlib").crc32(rays.tobytes())))
