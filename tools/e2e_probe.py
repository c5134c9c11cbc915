#!/usr/bin/env python
"""Per-phase wall times of the end-to-end C-ABI call sequence on shiny_teapot (where does e2e lose against the device-timed value?)."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import hobbyraytracer_b200 as pkg
from hobbyraytracer_b200 import abi, scenes
lib = pkg.load()
sd, cam, info = scenes.shiny_teapot_scene(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "_cache"))
w, h, spp = 1920, 1080, int(sys.argv[1]) if len(sys.argv) > 1 else 256
c = scenes.camera_for(cam, w, h)
h_sum = torch.empty((h, w, 3), dtype=torch.float32).pin_memory().numpy()
h_rgb = torch.empty((h, w, 3), dtype=torch.uint8).pin_memory().numpy()
orig = lib.fn
def timed(name):
    f = orig(name)
    def wr(*a):
        t = time.perf_counter(); r = f(*a); dt = time.perf_counter() - t
        if dt > 5e-4: print(f"   {name}: {dt*1e3:.1f} ms")
        return r
    return wr
lib.fn = timed
for k in range(4):
    t0 = time.perf_counter()
    s2 = sd.build(lib, abi.BVH_SAH, 3)
    p = abi.RenderParams(w, h, spp, 0, 0, 50, 0.001, 1, 0, 0)
    lib.check(lib.fn("render")(s2.h, C.byref(c), C.byref(p), h_sum.ctypes.data, h_rgb.ctypes.data), "render")
    st = s2.stats(); s2.close()
    dt = time.perf_counter() - t0
    print(f"step {k}: {dt*1e3:.1f} ms total, device {st.ms_total:.1f} ms, {st.rays/dt/1e6:.0f} Mrays/s e2e")
