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
# queues under pressure: 2 M camera paths through queues of 2^18 entries (the append margins and the "conservatively full" logic work hardest here),
# wavefront == megakernel bit for bit
os.environ["HRT_WAVE_CAPACITY"] = str(1 << 18)
for make in (lambda: scenes.teapot_scene(cache, n=16, env_w=64, env_h=32)[:2], lambda: scenes.cornell_smoke_scene(glass_sphere=True)[:2]):
    sdp, camp = make()
    gp = sdp.build(lib, abi.BVH_SAH, 3)
    cp = scenes.camera_for(camp, 512, 512)
    a = gp.render(cp, 512, 512, 8, seed=11)[0]
    b = gp.render(cp, 512, 512, 8, seed=11, flags=abi.RENDER_MEGAKERNEL)[0]
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "queues under pressure: wavefront != megakernel"
    gp.close()
del os.environ["HRT_WAVE_CAPACITY"]
lib.fn("trim")()
# bounds-checked variant (HRT_PRODUCT_SO=hobbyraytracer_b200/csrc/variants/libhrt_b200_bounds.so, built by tools/build_variant.sh bounds
# "-DHRT_BOUNDS_CHECK"): every indexed store of the queue machinery counted an index outside its array instead of writing; site 7 is the
# self test and must have fired once, every other counter must be zero
import ctypes as C
try:
    f = lib.dll.hrt_debug_bounds
except AttributeError:
    f = None
if f is not None:
    out = (C.c_uint * 8)()
    assert f(out) == 0
    counts = list(out)
    print("bounds counters (pending list, camera append, shade append, film, hit records, batch hit records, -, self test):", counts)
    assert counts[7] == 1 and counts[6] == 0 and not any(counts[:6]), counts
    print("bounds check ok: no index left its array")
print("sanitize_case ok")
