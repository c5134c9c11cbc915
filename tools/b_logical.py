#!/usr/bin/env python
"""SURVEY.md 8(d), supplementary figure: B_logical = bytes the walk LOGICALLY touches per ray, from the walk's own visit counts.

The counts come from the CPU simulation of the device headers (tests/hostsim: the same hrt_wide.cuh the kernels compile, one ray per thread, so
they are the counts of the algorithm, not of a warp's scheduling): wide-node steps (96 B: three 32-byte sectors), triangle leaves (96 B of the
leaf record + 32 B of the reference node's box when a triangle accepted - counted as 128), mesh entries (112 B object record + 96 B mesh
record), other world-list entries (112 B).  Beside them: SURVEY's input-determined B_alg and the DRAM bytes per ray ncu measured for the kernel
(profiles/traffic.json) - what the caches save.  Needs no GPU.

usage: python tools/b_logical.py [--quads 2290] [--rays 262144]
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from hobbyraytracer_b200 import abi, scenes

ap = argparse.ArgumentParser()
ap.add_argument("--quads", type=int, default=2290)
ap.add_argument("--rays", type=int, default=1 << 18)
args = ap.parse_args()
subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "tests", "hostsim")], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
lib = abi.Lib(os.path.join(ROOT, "tests", "hostsim", "_build", "libhrt_hostsim.so"), "hsim_")
f = lib.dll.hsim_wide_visit_counts
f.restype = C.c_int; f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_uint64, C.c_void_p]
g = lib.dll.hsim_trace_batch_wide
g.restype = C.c_int; g.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
cache = os.path.join(ROOT, "tests", "_cache")
traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))


def counts(sd, rays):
    sc = sd.build(lib, abi.BVH_SAH, 3)
    out = (C.c_uint64 * 5)()
    rays = np.ascontiguousarray(rays, np.float32)
    assert f(sc.h, rays.ctypes.data, len(rays), 0, out) == 0
    return [v / len(rays) for v in out]


def secondary_rays(sd, cam, w, h, n, seed=3):
    """rays leaving the surfaces camera rays hit, directions uniform on the sphere: the population the secondary pass sees"""
    sc = sd.build(lib, abi.BVH_SAH, 3)
    prim = np.ascontiguousarray(scenes.camera_rays(cam, w, h, 4 * n, seed=seed), np.float32)
    p = np.empty(len(prim), np.uint32); t = np.empty(len(prim), np.float32); st = C.c_uint64(0)
    assert g(sc.h, prim.ctypes.data, len(prim), p.ctypes.data, t.ctypes.data, 0, C.byref(st)) == 0
    hit = (p != abi.MISS) & np.isfinite(t)
    o = prim[hit, 0:3] + t[hit, None] * prim[hit, 4:7]
    rng = np.random.default_rng(seed)
    d = rng.normal(size=o.shape); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.empty((len(o), 8), np.float32); rays[:, 0:3] = o; rays[:, 3] = 1e-3; rays[:, 4:7] = d; rays[:, 7] = np.inf
    return rays[:n]


rows = []
sd, cam, _ = scenes.teapot_scene(cache, n=56, env_w=64, env_h=32)
rows.append(("C1 teapot_scene, secondary rays", counts(sd, secondary_rays(sd, cam, 640, 640, args.rays)), 640, "teapot_scene"))
sd, cam, _ = scenes.big_mesh_scene(cache, quads=args.quads, env_w=64, env_h=32)
rows.append((f"C5b ray batch vs {2 * args.quads * args.quads} triangles", counts(sd, scenes.random_rays(args.rays, (0, 1, 0), 1.8, seed=0xC0FFEE)), 992, "raybatch"))
rows.append((f"C5 big_mesh, secondary rays", counts(sd, secondary_rays(sd, cam, 3840, 2160, args.rays)), 992, "big_mesh"))
print(f"{'rays':44s} {'node':>6s} {'leaf':>6s} {'mesh':>6s} {'entry':>6s} {'pass 2':>7s} {'B_logical':>10s} {'B_alg':>6s} {'DRAM B/ray (ncu)':>17s}")
for name, (kn, kl, km, ko, pend), b_alg, key in rows:
    b_log = 176 + 96 * kn + 128 * kl + 208 * km + 112 * ko
    tr = traffic.get(key) or {}
    dram = ""
    if key == "raybatch" and tr.get("bytes"):
        dram = f"{tr['bytes'] / (2 << 20):.0f}"
    print(f"{name:44s} {kn:6.2f} {kl:6.2f} {km:6.2f} {ko:6.2f} {pend:7.4f} {b_log:10.0f} {b_alg:6d} {dram:>17s}")
