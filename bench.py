#!/usr/bin/env python
"""bench.py — Mrays/s of the path-tracing hot path on N B200s, beside the reference's CPU renderer.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's own CPU renderer on the host cores

A "step" is one full render of the workload (default: BASELINE.json configs[1], shiny_teapot 1920x1080 at 1024 spp)
with the scene already resident in HBM: a persistent wavefront (trace, shade + compaction + path regeneration + film
add, repeated until the image is out of samples), the spp split over the ranks, one NCCL reduce of the integer film to rank 0, film resolve (mean -> ACES -> sqrt -> 8 bit) on rank 0.
`value` = rays (closest-hit queries, = world->hit calls of main.cpp:45) of all ranks / max-over-ranks device time.
`e2e` = the same through the host-buffer C ABI: scene build + commit (host->device upload of geometry, BVHs, textures),
hrt_render / render_split, device->host read of the float film and the 8-bit image, every step.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from hobbyraytracer_b200 import abi, scenes  # noqa: E402

CACHE = os.path.join(ROOT, "tests", "_cache")
BVH_MODES = {"reference": 0, "sah": 1, "lbvh_gpu": 2, "lbvh_host": 3}      # abi.BVH_*
ORACLE_SO = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libhrt_ref.so")


_REAL_STDOUT = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout; libraries (NCCL prints its version there) get stderr instead."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n"); out.flush()


def make_workload(name, args):
    """-> (SceneDesc, camera dict, width, height, spp, n_prims-for-B_alg, prim bytes)"""
    if name == "shiny_teapot":
        sd, cam, info = scenes.shiny_teapot_scene(CACHE, n=56, env_w=4096, env_h=2048)
    elif name == "shiny_teapot_glass":      # BASELINE config 2's "rough dielectric" wording: the same scene with Dielectric(1.5, 0.1)
        sd, cam, info = scenes.shiny_teapot_scene(CACHE, n=56, env_w=4096, env_h=2048, rough_dielectric=True); info["name"] = "shiny_teapot_glass"
    elif name == "teapot_scene":
        sd, cam, info = scenes.teapot_scene(CACHE, n=56, env_w=4096, env_h=2048)
    elif name == "scattered_balls":
        sd, cam, info = scenes.scattered_balls_scene(seed=42, grid=11)
    elif name == "cornell_smoke":
        sd, cam, info = scenes.cornell_smoke_scene()
    elif name == "big_mesh":
        sd, cam, info = scenes.big_mesh_scene(CACHE, quads=args.quads, indexed=args.indexed)
    else:
        raise SystemExit(f"unknown workload {name}")
    w = args.width or info["width"]; h = args.height or info["height"]; spp = args.spp or info["spp"]
    return sd, cam, w, h, spp, info


def alg_bytes_per_ray(sd):
    """B_alg = B_state (176 B of ray / hit / path-state streams) + one root-to-leaf path of 32 B nodes
    + one primitive (48 B triangle / 16 B sphere / 32 B rect).  Input-determined: it does not shrink when the kernel
    keeps a camera path in registers instead of streaming it."""
    n_prims = 0; prim_bytes = 16
    for s in sd.shapes:
        if s[0] == "mesh":
            n_prims += s[1].shape[0]; prim_bytes = 48
        elif s[0] == "mesh_indexed":
            n_prims += s[4].shape[0] // 3; prim_bytes = 48
        elif s[0] == "box":
            n_prims += 6; prim_bytes = max(prim_bytes, 32)
        else:
            n_prims += 1
            if s[0] == "rect":
                prim_bytes = max(prim_bytes, 32)
    depth = int(np.ceil(np.log2(max(n_prims, 2))))
    return 176 + 32 * depth + prim_bytes, n_prims


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) >= 6 and r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def load_peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    except Exception:
        return 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


# ---------------------------------------------------------------------------------------------------------------------
def cpu_render_rate(sd, cam, w, h, target_s, seed=1):
    """The reference's own renderer (oracle/_ref, `reference`) or the oracle restatement (`port`) on all host cores, on a
    bounded sample of the workload: the full frame at the spp that takes about target_s seconds."""
    if os.path.exists(REF_SO):
        lib, kind = abi.Lib(REF_SO, "ref_"), "reference"
    else:
        lib, kind = abi.Lib(ORACLE_SO, "orc_"), "port"
    sc = sd.build(lib, abi.BVH_REFERENCE, 3, tagged=False) if kind == "reference" else sd.build(lib, abi.BVH_REFERENCE, 3)
    c = scenes.camera_for(cam, w, h)
    cores = lib.fn("hardware_threads")()
    # calibrate on a strip of rows: a narrower camera would change the rays, so render whole frames at low spp instead
    t0 = time.perf_counter(); _, rays1 = sc.render_f32(c, w, h, 1, seed=seed, threads=cores); t1 = time.perf_counter() - t0
    spp = max(1, min(64, int(target_s / max(t1, 1e-3))))
    t0 = time.perf_counter(); _, rays = sc.render_f32(c, w, h, spp, seed=seed + 1, threads=cores); dt = time.perf_counter() - t0
    return {"rays": rays, "seconds": dt, "spp": spp, "cores": cores, "kind": kind, "samples": w * h * spp,
            "sample": f"full {w}x{h} frame at {spp} spp ({w * h * spp / 1e6:.1f} M camera samples, {rays / 1e6:.1f} M rays, {dt:.1f} s), "
                      f"{cores} threads, thread-local rand()"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sd, cam, w, h, spp, info = make_workload(args.workload, args)
    os.makedirs(os.path.join(ROOT, "oracle", "_build"), exist_ok=True)
    if not os.path.exists(REF_SO) and not os.path.exists(ORACLE_SO):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True)
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        r = cpu_render_rate(sd, cam, w, h, per_step / 4)
    tot_rays = tot_s = tot_samples = 0
    for _ in range(args.steps):
        r = cpu_render_rate(sd, cam, w, h, per_step)
        tot_rays += r["rays"]; tot_s += r["seconds"]; tot_samples += r["samples"]
    value = tot_rays / tot_s / 1e6
    line = {"impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * tot_s / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "samples_per_s": tot_samples / tot_s,
            "config": {"workload": f"{info['name']} {w}x{h}, {info['spp']} spp (CPU arm: bounded sample per step)", "width": w, "height": h, "spp": info["spp"]},
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ---------------------------------------------------------------------------------------------------------------------
def run_own(args):
    import torch
    import hobbyraytracer_b200 as pkg
    from hobbyraytracer_b200 import multigpu

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = pkg.load()
    lib.check(lib.fn("set_device")(local), "set_device")
    sd, cam, w, h, spp, info = make_workload(args.workload, args)
    c = scenes.camera_for(cam, w, h)
    bvh = BVH_MODES[args.bvh]
    t0 = time.perf_counter(); sc = sd.build(lib, bvh, 3); t_commit = time.perf_counter() - t0
    npix = w * h
    film = multigpu.new_film(lib, w, h)
    fsum = torch.zeros(npix * 3, dtype=torch.float32, device="cuda")
    rgb8 = torch.zeros(npix * 3, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    flags = abi.RENDER_MEGAKERNEL if args.megakernel else (abi.RENDER_WAVEFRONT_V1 if args.pipeline == "v1" else 0)

    # every rank works on one explicit CUDA stream (the library's NULL stream is the scene's own, which torch does not see)
    work = torch.cuda.Stream(); torch.cuda.set_stream(work)
    stream = work.cuda_stream

    pg = None
    if world > 1 and args.reduce == "peer":
        pg = multigpu.PeerGroup(sc, w, h, rank, world, dist)
        if not pg.ok:                                   # no CUDA IPC / peer access on this box: NCCL reduce (still a GPU path)
            pg.close(); pg = None
    reduce_kind = "none (1 GPU)" if world == 1 else ("every rank sums and resolves its pixel slice from all films over NVLink peer memory (CUDA IPC) into the root's buffers; stream-ordered barriers" if pg else "NCCL reduce(SUM, int64) + resolve")

    def render_step(scene, out_sum=True):
        if pg is not None:
            pg.render(c, spp, seed=args.seed, flags=flags, want_sum=out_sum, scene=scene)
            return
        film.zero_()
        multigpu.render_split(scene, c, w, h, spp, film, rank, world, dist, seed=args.seed, stream_ptr=stream, rgb8=rgb8, total=fsum if out_sum else None, flags=flags)

    def step():
        render_step(sc, out_sum=False)

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # N > 1: the peer-memory reduction against the NCCL integer reduce on a small frame, bit for bit, before anything is timed
    # (the 2-GPU pytest of the same property is skipped on 1-GPU test boxes; this check runs wherever the bench runs)
    reduce_parity = None
    if pg is not None:
        w0, h0, spp0 = 256, 144, 8 * world + 3
        c0 = scenes.camera_for(cam, w0, h0)
        pg0 = multigpu.PeerGroup(sc, w0, h0, rank, world, dist)
        f0 = multigpu.new_film(lib, w0, h0); s0 = torch.zeros(w0 * h0 * 3, dtype=torch.float32, device="cuda"); b0 = torch.zeros(w0 * h0 * 3, dtype=torch.uint8, device="cuda")
        same = 1
        if pg0.ok:
            pg0.render(c0, spp0, seed=args.seed + 1)
            multigpu.render_split(sc, c0, w0, h0, spp0, f0, rank, world, dist, seed=args.seed + 1, stream_ptr=stream, rgb8=b0, total=s0)
            torch.cuda.synchronize()
            if rank == 0:
                ps, pb = pg0.root_views()
                same = int(torch.equal(ps.reshape(-1).view(torch.int32), s0.view(torch.int32)) and torch.equal(pb.reshape(-1), b0) and float(s0.sum()) > 0)
        else:
            same = 0
        flag = torch.tensor([same], dtype=torch.int32, device="cuda"); dist.broadcast(flag, src=0)
        reduce_parity = bool(int(flag.item()))
        dist.barrier(); pg0.close()
        if not reduce_parity:
            raise SystemExit("the peer-memory film reduction differs from the NCCL integer reduce")

    for _ in range(args.warmup):
        step()
    sync_all()
    lib.fn("reset_stats")(sc.h)
    clocks = ClockSampler(local) if rank == 0 else None
    if clocks:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    sync_all()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    st = sc.stats()
    tot = torch.tensor([float(st.rays), float(st.paths), float(st.launches)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX); dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms = float(ms.item()); rays, paths, launches = [float(x) for x in tot.tolist()]
    clk = clocks.stop() if clocks else None

    # ---- end to end through the host-buffer C ABI: build + commit (upload) + render + film read-back, every step ----
    # Steps are independent frames, pipelined two deep like a host that renders an animation would: a builder thread constructs and
    # commits (BVH build + host -> device upload) the scene of step k + 1 while step k renders; every step still builds, uploads,
    # renders and reads back everything it needs.
    import ctypes as C
    import queue
    h_sum = torch.empty((h, w, 3), dtype=torch.float32).pin_memory()
    h_rgb = torch.empty((h, w, 3), dtype=torch.uint8).pin_memory()
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    n_build = e2e_steps + 2
    ready = queue.Queue(maxsize=1)

    def builder():
        lib.check(lib.fn("set_device")(local), "set_device")
        for _ in range(n_build):
            ready.put(sd.build(lib, bvh, 3))

    def e2e_step():
        s2 = ready.get()
        if world == 1:
            p = abi.RenderParams(w, h, spp, 0, 0, 50, 0.001, args.seed, 0, flags)
            lib.check(lib.fn("render")(s2.h, C.byref(c), C.byref(p), h_sum.numpy().ctypes.data, h_rgb.numpy().ctypes.data), "render")
        else:
            render_step(s2, out_sum=True)
            if rank == 0:
                if pg is not None:
                    ps, pb = pg.root_views()
                    h_sum.copy_(ps, non_blocking=True); h_rgb.copy_(pb, non_blocking=True)
                else:
                    h_sum.view(-1).copy_(fsum, non_blocking=True); h_rgb.view(-1).copy_(rgb8, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        stt = s2.stats(); s2.close()
        return stt

    th = threading.Thread(target=builder, daemon=True); th.start()
    e2e_step(); e2e_step(); sync_all()                 # two untimed end-to-end steps: host allocator, page cache and driver pools settle
    t0 = time.perf_counter(); e2e_rays = 0.0; scene_bytes = 0
    for _ in range(e2e_steps):
        stt = e2e_step(); e2e_rays += float(stt.rays); scene_bytes = int(stt.h2d_bytes)
    sync_all()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    e2e_r = torch.tensor([e2e_rays], dtype=torch.float64, device="cuda")
    th.join()
    if dist is not None:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX); dist.all_reduce(e2e_r, op=dist.ReduceOp.SUM)

    if rank == 0:
        peak, peak_src = load_peaks()
        b_alg, n_prims = alg_bytes_per_ray(sd)
        rank_rays = float(st.rays)
        ach = rank_rays * b_alg / (st.ms_trace * 1e-3) / 1e9 if st.ms_trace > 0 else None
        traffic = traffic_src = traffic_ms = traffic_kernel = None
        try:                                    # dram__bytes of one launch of the dominant kernel from an ncu --set full capture (not measured in this run)
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            ent = tj.get(args.workload)
            if isinstance(ent, dict):
                traffic, traffic_src, traffic_ms, traffic_kernel = ent.get("bytes"), ent.get("source"), ent.get("launch_ms"), ent.get("kernel")
        except Exception:
            pass
        line = {"metric": "Mrays/s", "value": rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "samples_per_s": paths / (ms * 1e-3), "samples_per_s_per_gpu": paths / (ms * 1e-3) / world, "rays_per_sample": rays / max(paths, 1),
                "config": {"workload": f"{info['name']} {w}x{h}, {spp} spp split over {world} GPU(s)", "width": w, "height": h, "spp": spp, "max_depth": 50,
                           "bvh": args.bvh, "n_prims": n_prims, "film_reduce": reduce_kind, "pipeline": "megakernel" if args.megakernel else ("persistent wavefront, every ray queued" if args.pipeline == "v1" else "wavefront, camera paths fused into the trace pass"),
                           "cache": "camera paths are generated on the device every step (nothing to reuse); between steps the film (32 B per pixel) is re-zeroed and the ray queues (up to 5 GB) rewritten, both larger than L2 with the 33 MB scene; no flush needed"},
                "e2e": {"value": float(e2e_r.item()) / float(e2e_s.item()) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": scene_bytes,
                        "d2h_bytes_per_step": npix * 15, "steps": e2e_steps, "includes": "scene build + BVH build + upload, render, float film + 8-bit image read-back; steps pipelined two deep (a builder thread commits the scene of step k + 1 while step k renders)"},
                "gpu_launches": int(launches), "commit_s": t_commit, "build_ms": st.ms_build, "replayed_fraction": float(st.retraced) / max(float(st.rays), 1.0), "ref_walk_rays": int(st.prim_tests),
                "kernel_ms": {"trace": st.ms_trace, "shade": st.ms_shade, "other": st.ms_other, "total": st.ms_total, "note": "rank 0, CUDA events inside the timed region"},
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": (ach / peak) if ach else None, "traffic": traffic, "traffic_source": traffic_src, "traffic_launch_ms": traffic_ms,
                             "traffic_gbs": (traffic / (traffic_ms * 1e-3) / 1e9) if traffic and traffic_ms else None,
                             "kernel": "k_wave_trace" if args.pipeline == "v1" else (traffic_kernel or "k_wave2_trace"), "alg_bytes_per_ray": b_alg, "trace_launches": int(st.launches) // 2, "peak_source": peak_src,
                             "note": "achieved = rays x B_alg / device time of the traversal kernel (CUDA events); these are the input-determined bytes SURVEY 8(d) charges per ray, not bytes moved: camera paths live in registers and the scene is L1/L2 resident, so the fraction can exceed 1 on scenes dominated by camera rays that miss; traffic_gbs is the DRAM rate ncu measured for one launch of the kernel, and the kernel is issue / latency bound (DESIGN.md section 3)"},
                "clocks": clk}
        if reduce_parity is not None:
            line["reduce_parity"] = reduce_parity
        if world == 1 and not args.no_cpu:
            try:
                r = cpu_render_rate(sd, cam, w, h, args.cpu_seconds)
                line["cpu_baseline"] = {"value": r["rays"] / r["seconds"] / 1e6, "unit": "Mrays/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
                                        "samples_per_s": r["samples"] / r["seconds"]}
            except Exception as e:  # the checker is missing: say so, never substitute
                line["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": 0, "kind": "unavailable", "sample": str(e)}
        emit(line)
    if dist is not None:
        dist.barrier(); dist.destroy_process_group()


def run_raybatch(args):
    """C5(b): the incoherent-ray intersection micro-benchmark: N rays, origins uniform in a ball around the
    big mesh, directions uniform on S^2, split in contiguous slices over the ranks (no exchange step), rays resident in HBM.
    Checks on the fly: the SAH-order answers of a 1 M-ray slice equal the reference-order answers bit for bit."""
    import torch
    import hobbyraytracer_b200 as pkg
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = pkg.load(); lib.check(lib.fn("set_device")(local), "set_device")
    sd, cam, info = scenes.big_mesh_scene(CACHE, quads=args.quads, env_w=64, env_h=32)
    t0 = time.perf_counter(); sc = sd.build(lib, BVH_MODES[args.bvh], 3); t_commit = time.perf_counter() - t0
    n_total = args.rays; n = n_total // world
    rays_h = scenes.random_rays(n, (0, 1, 0), 1.2 * 1.5, seed=0xC0FFEE + rank)
    rays = torch.from_numpy(rays_h).cuda()
    prim = torch.empty(n, dtype=torch.int32, device="cuda"); t = torch.empty(n, dtype=torch.float32, device="cuda")
    tstream = torch.cuda.Stream()                               # an explicit stream: the library's own is invisible to torch events
    torch.cuda.synchronize(); torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream

    def step():
        lib.check(lib.fn("trace_batch_device")(sc.h, rays.data_ptr(), n, prim.data_ptr(), t.data_ptr(), 0, stream), "trace_batch_device")

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record(); sync_all()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    # parity property on a slice: reference visiting order == SAH order
    k = min(n, 1 << 20)
    a = sc.trace(rays_h[:k], extras=False)
    ref = sd.build(lib, abi.BVH_REFERENCE, 3).trace(rays_h[:k], extras=False)
    same = bool(np.array_equal(a["prim"], ref["prim"]) and np.array_equal(a["t"].view(np.uint32), ref["t"].view(np.uint32)))
    hit_rate = float((a["prim"] != abi.MISS).mean())
    # end to end through the host-buffer entry (hrt_trace_batch): pinned host rays in, pinned host (prim, t) out, copies inside the timed region
    h_rays = torch.from_numpy(rays_h).pin_memory()
    h_prim = torch.empty(n, dtype=torch.int32).pin_memory(); h_t = torch.empty(n, dtype=torch.float32).pin_memory()

    def e2e_step():
        lib.check(lib.fn("trace_batch")(sc.h, h_rays.data_ptr(), n, h_prim.data_ptr(), h_t.data_ptr(), None, None, None, 0), "trace_batch")

    e2e_step(); sync_all()
    t1 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    e2e_s = torch.tensor([(time.perf_counter() - t1) / args.steps], dtype=torch.float64, device="cuda")
    assert np.array_equal(h_prim.numpy()[:k].view(np.uint32), a["prim"]) and np.array_equal(h_t.numpy()[:k].view(np.uint32), a["t"].view(np.uint32))
    if dist is not None:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    if rank == 0:
        peak, peak_src = load_peaks()
        b_alg, n_prims = alg_bytes_per_ray(sd)
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("raybatch") or {}
        except Exception:
            tr = {}
        rate = n * world * args.steps / (ms * 1e-3)
        st = sc.stats()
        emit({"metric": "Mrays/s", "value": rate / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                          "config": {"workload": f"ray batch: {n_total} incoherent rays vs {n_prims}-triangle mesh, split over {world} GPU(s)", "n_prims": n_prims,
                                     "rays": n_total, "cache": "16 M rays x 32 B = 512 MB per step, larger than L2"},
                          "e2e": {"value": n * world / float(e2e_s.item()) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": n * 32, "d2h_bytes_per_step": n * 8},
                          "gpu_launches": int(st.launches), "commit_s": t_commit, "build_ms": st.ms_build, "scene_bytes": int(st.scene_bytes),
                          "parity": {"sah_equals_reference_order": same, "rays_checked": k, "hit_rate": hit_rate},
                          "roofline": {"bound": "hbm", "achieved": rate / world * b_alg / 1e9, "peak": peak, "unit": "GB/s", "frac": rate / world * b_alg / 1e9 / peak,
                                       "traffic": tr.get("bytes"), "traffic_launch_ms": tr.get("launch_ms"), "traffic_source": tr.get("source"),
                                       "traffic_gbs": (tr["bytes"] / (tr["launch_ms"] * 1e-3) / 1e9) if tr.get("bytes") and tr.get("launch_ms") else None,
                                       "kernel": "k_trace_batch", "alg_bytes_per_ray": b_alg, "peak_source": peak_src}})
    if dist is not None:
        dist.barrier(); dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--workload", default="shiny_teapot")
    ap.add_argument("--spp", type=int, default=0, help="samples per pixel per step (default: the workload's own)")
    ap.add_argument("--width", type=int, default=0); ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--quads", type=int, default=2290)
    ap.add_argument("--indexed", action="store_true", help="big_mesh: hand the mesh over as vertex / index buffers (hrt_scene_add_mesh_indexed)")
    ap.add_argument("--rays", type=int, default=16 << 20, help="--workload raybatch: total rays")
    ap.add_argument("--bvh", default="sah", choices=["sah", "reference", "lbvh_gpu", "lbvh_host"],
                    help="sah: host binned-SAH trees; lbvh_gpu: the whole per-mesh build on the device (Morton LBVH); lbvh_host: its host mirror")
    ap.add_argument("--megakernel", action="store_true")
    ap.add_argument("--reduce", default="peer", choices=["peer", "nccl"], help="N > 1: how the rank films reach rank 0")
    ap.add_argument("--pipeline", default="fused", choices=["fused", "v1"], help="fused: camera paths traced in the trace pass, only hits are queued; v1: every ray queued")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "raybatch":
        run_raybatch(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
