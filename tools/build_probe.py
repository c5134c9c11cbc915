#!/usr/bin/env python
"""hrt_scene_commit on the 10.5 M-triangle mesh (C5), per build mode and ingest form: wall time of the add + commit calls,
the library's own split (upload / device build), bytes host -> device.  Run under `ncu --metrics gpu__time_duration.sum`
for the launch list of the device builder (profiles/*_build_launches.csv).
usage: python tools/build_probe.py [quads=2290] [modes=lbvh_gpu,sah] [forms=indexed,soup] [repeats=3]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hobbyraytracer_b200 as pkg
from hobbyraytracer_b200 import abi, scenes

quads = int(sys.argv[1]) if len(sys.argv) > 1 else 2290
modes = (sys.argv[2] if len(sys.argv) > 2 else "lbvh_gpu,sah").split(",")
forms = (sys.argv[3] if len(sys.argv) > 3 else "indexed,soup").split(",")
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
MODES = {"reference": abi.BVH_REFERENCE, "sah": abi.BVH_SAH, "lbvh_gpu": abi.BVH_LBVH_GPU, "lbvh_host": abi.BVH_LBVH_HOST}
lib = pkg.load()
cache = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "_cache")
for form in forms:
    sd, cam, info = scenes.big_mesh_scene(cache, quads=quads, indexed=(form == "indexed"))
    for mode in modes:
        for k in range(reps):
            t0 = time.perf_counter()
            sc = sd.build(lib, MODES[mode], 3)
            dt = time.perf_counter() - t0
            st = sc.stats(); sc.close()
            print(f"{form:8s} {mode:9s} run {k}: add+commit {dt * 1e3:8.1f} ms | ms_build {st.ms_build:8.1f} (upload {st.ms_build_upload:7.1f}) | "
                  f"h2d {st.h2d_bytes / 1e6:8.1f} MB | resident {st.scene_bytes / 1e6:8.1f} MB | launches {st.launches}", flush=True)
