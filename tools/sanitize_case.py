#!/usr/bin/env python
"""Small invocations of every kernel family for compute-sanitizer (one tool per gpurun call, B200_PROFILING.md):
    compute-sanitizer --tool memcheck|racecheck|initcheck python tools/sanitize_case.py
64 x 64 x 16-spp renders of a mesh-in-a-Cornell-box scene through the default pipeline, the wide-tree kernels (forced with
HRT_FORCE_WIDE_SECONDARY / HRT_FORCE_WIDE_PRIMARY, set by this script before the library loads), the megakernel, the Russian-roulette
shade pass and a progressive render; a 20 k-ray batch (wide walk + pending list + resolve); texture / scatter batches; a device build."""
import os, sys
os.environ["HRT_FORCE_WIDE_SECONDARY"] = "1"; os.environ["HRT_FORCE_WIDE_PRIMARY"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import hobbyraytracer_b200 as pkg
from hobbyraytracer_b200 import abi, scenes
lib = pkg.load()
cache = os.path.join(ROOT, "tests", "_cache")
sd, cam, _ = scenes.teapot_scene(cache, n=16, env_w=64, env_h=32)
w, h, spp = 64, 64, 16
c = scenes.camera_for(cam, w, h)
out = {}
for name, mode in (("sah", abi.BVH_SAH), ("lbvh_gpu", abi.BVH_LBVH_GPU)):
    g = sd.build(lib, mode, 3)
    out[name] = g.render(c, w, h, spp, seed=3)[0]                                        # forced wide secondary + wide primary kernels
    mk = g.render(c, w, h, spp, seed=3, flags=abi.RENDER_MEGAKERNEL)[0]
    assert np.array_equal(out[name].view(np.uint32), mk.view(np.uint32)), name
    g.render(c, w, h, spp, seed=3, flags=abi.RENDER_RR)
    g.render_progressive(c, w, h, spp, seed=3, slice_spp=5)
    rays = scenes.random_rays(20000, (0, 2.5, 0), 2.4, seed=2)
    rays[:50, 4:7] = [0, 0, 1]                                                            # zero direction components: the pending list, trace_ref
    g.trace(rays)
    g.close()
sd2, cam2, _ = scenes.cornell_smoke_scene(glass_sphere=True)
g = sd2.build(lib, abi.BVH_SAH, 3)
g.render(scenes.camera_for(cam2, w, h), w, h, spp, seed=1)
x = np.random.default_rng(0).normal(size=(1000, 14)).astype(np.float32)
g.scatter(0, x, np.ones(1000, np.uint8), seed=1)
g.texture_eval(0, np.random.default_rng(1).random((1000, 5)).astype(np.float32))
g.close()
lib.fn("trim")()
print("sanitize_case ok")
