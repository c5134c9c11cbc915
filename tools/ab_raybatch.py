#!/usr/bin/env python
"""A/B of library builds on the incoherent ray batch (C5b): tools/ab_raybatch.py [--quads Q] [--rays N] lib1.so lib2.so ...
Each library runs in its own process (HRT_PRODUCT_SO); the mesh arrays are generated once and shared through /tmp.
Prints Mrays/s (device-timed, best of 3 after a warm-up) and a checksum of the hit ids, which must agree between builds."""
import os, subprocess, sys, time, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

def arg(name, default):
    return int(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else default

quads, nrays = arg("--quads", 2290), arg("--rays", 8 << 20)
cache = f"/tmp/ab_mesh_{quads}.npz"

def worker():
    import torch
    import hobbyraytracer_b200 as pkg
    from hobbyraytracer_b200 import abi, scenes
    lib = pkg.load()
    z = np.load(cache)
    sd = abi.SceneDesc(); m = sd.material(abi.MAT_METAL, (0.73, 0.73, 0.73), p0=0.2)
    sd.add(sd.mesh(z["pos"], z["nrm"], z["uv"], m))
    t0 = time.perf_counter(); sc = sd.build(lib, abi.BVH_SAH, 3); tb = time.perf_counter() - t0
    rays = torch.from_numpy(scenes.random_rays(nrays, (0, 1, 0), 1.8, seed=0xC0FFEE)).cuda()
    prim = torch.empty(nrays, dtype=torch.int32, device="cuda"); t = torch.empty(nrays, dtype=torch.float32, device="cuda")
    st = torch.cuda.Stream(); torch.cuda.set_stream(st)
    f = lib.fn("trace_batch_device")
    best = 1e9
    for k in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); lib.check(f(sc.h, rays.data_ptr(), nrays, prim.data_ptr(), t.data_ptr(), 0, st.cuda_stream), "trace"); e1.record(); torch.cuda.synchronize()
        if k: best = min(best, e0.elapsed_time(e1))
    crc = zlib.crc32(prim.cpu().numpy().tobytes()) ^ zlib.crc32(t.cpu().numpy().tobytes())
    print(f"{os.path.basename(os.environ.get('HRT_PRODUCT_SO', 'product')):40s} {nrays / best / 1e3:8.1f} Mrays/s  {best:7.3f} ms  crc {crc:08x}  build {tb:.1f} s", flush=True)

if "--worker" in sys.argv:
    worker(); sys.exit(0)
if not os.path.exists(cache):
    from hobbyraytracer_b200 import scenes
    pos, nrm, uv = scenes.blob_mesh(quads, quads, radius=1.0, bump=0.2, seed=7, center=(0, 1, 0))
    np.savez(cache, pos=pos, nrm=nrm, uv=uv)
libs = [a for a in sys.argv[1:] if a.endswith(".so")] or [None]
for so in libs:
    env = dict(os.environ)
    if so: env["HRT_PRODUCT_SO"] = os.path.abspath(so)
    subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", "--quads", str(quads), "--rays", str(nrays)], env=env)
