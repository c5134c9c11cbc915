#!/usr/bin/env python
"""Big-mesh parity probe on the GPU: the wide walk against the two-child walk of the same scene (must be identical), and both
against the reference visiting order; prints what the mismatching rays look like (ties in t between neighbouring triangles?)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import hobbyraytracer_b200 as pkg
from hobbyraytracer_b200 import abi, scenes
lib = pkg.load()
quads = int(sys.argv[1]) if len(sys.argv) > 1 else 2290
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
sd, cam, info = scenes.big_mesh_scene(os.path.join(ROOT, "tests", "_cache"), quads=quads, env_w=64, env_h=32)
rays = scenes.random_rays(n, (0, 1, 0), 1.8, seed=0xC0FFEE)
ref = sd.build(lib, abi.BVH_REFERENCE, 3).trace(rays, extras=False)
for name, mode in (("sah", abi.BVH_SAH), ("lbvh_gpu", abi.BVH_LBVH_GPU), ("lbvh_host", abi.BVH_LBVH_HOST)):
    sc = sd.build(lib, mode, 3)
    os.environ.pop("HRT_NO_WIDE", None); w = sc.trace(rays, extras=False)
    os.environ["HRT_NO_WIDE"] = "1"; two = sc.trace(rays, extras=False); os.environ.pop("HRT_NO_WIDE")
    d_wt = int((w["prim"] != two["prim"]).sum() + (w["t"].view(np.uint32) != two["t"].view(np.uint32)).sum())
    bad = np.nonzero((w["prim"] != ref["prim"]) | (w["t"].view(np.uint32) != ref["t"].view(np.uint32)))[0]
    print(f"{name}: wide vs two-child differences {d_wt}; wide vs reference order {len(bad)} of {n}")
    for i in bad[:6]:
        print("   ray", i, "wide", w["prim"][i], repr(float(w["t"][i])), "ref", ref["prim"][i], repr(float(ref["t"][i])), "two", two["prim"][i], repr(float(two["t"][i])))
    sc.close()
