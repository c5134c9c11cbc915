#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/libhrt_ref.so, built from
/root/reference by oracle/Makefile).  Run in the build container:  python tests/golden/make_golden.py
The fixtures hold inputs (mesh arrays, ray batches, camera parameters, film sums, texture lookups) and the reference's
outputs; tests/test_golden.py replays them without the reference."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, HERE)

from hobbyraytracer_b200 import abi, scenes  # noqa: E402
import cases  # noqa: E402

F = np.float32
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libhrt_ref.so")


def special_rays(rng, center, n):
    """Rays the reference treats specially: axis-parallel directions (zero components -> NaN t hits in ITriangle),
    huge / tiny directions, origins on primitive surfaces, t_min above the first hit, finite t_max."""
    r = scenes.random_rays(n, center, 3.0, seed=int(rng.integers(1 << 30)))
    k = n // 8
    r[0:k, 4:7] = np.eye(3, dtype=F)[rng.integers(0, 3, k)] * rng.choice([-1, 1], (k, 1)).astype(F)
    r[k:2 * k, 4:7] *= F(1e-3); r[2 * k:3 * k, 4:7] *= F(250.0)
    r[3 * k:4 * k, 3] = F(2.0)                                  # t_min beyond near hits
    r[4 * k:5 * k, 7] = rng.random(k).astype(F) * 4             # finite t_max
    r[5 * k:6 * k, 4 + rng.integers(0, 3)] = 0                  # one zero component
    return r


def main():
    if not os.path.exists(REF_SO):
        raise SystemExit("build oracle/_ref first: make -C oracle ref")
    ref = abi.Lib(REF_SO, "ref_")
    rng = np.random.default_rng(20261104)
    pos, nrm, uv = scenes.teapot_like_mesh(8)
    fx_common = dict(mesh_pos=pos.astype(F), mesh_nrm=nrm.astype(F), mesh_uv=uv.astype(F))

    # ---- hit() on ray batches ---------------------------------------------------------------------------------
    for name, make in cases.TRACE_SCENES.items():
        sd = make(fx_common)
        sc = sd.build(ref, abi.BVH_REFERENCE, 3)
        center = (0, 0.5, -1) if name == "prims" else (0, 1, 0)
        rays = np.concatenate([scenes.random_rays(6000, center, 4.0, seed=11), special_rays(rng, center, 2000)])
        out = sc.trace(rays)
        fx = dict(fx_common) if name != "prims" else {}
        fx.update(rays=rays, prim=out["prim"], t=out["t"], normal=out["normal"], uv=out["uv"], front=out["front"])
        np.savez_compressed(os.path.join(HERE, f"trace_{name}.npz"), **fx)
        print(name, "hits", int((out["prim"] != 0xFFFFFFFF).sum()), "of", len(rays), "NaN t", int(np.isnan(out["t"]).sum()))

    # ---- camera, film, textures ------------------------------------------------------------------------------
    cam = scenes.camera_for(cases.RENDER_CAMERA, 48, 32)
    st = rng.random((4000, 2)).astype(F) * 1.1 - 0.05
    cam_out = np.empty((4000, 6), F)
    ref.check(ref.fn("camera_rays")(abi.C.byref(cam), abi._ptr(st), 4000, abi._ptr(cam_out)), "camera_rays")
    sums = (rng.random((32, 48, 3)) ** 4 * 64).astype(F) * 16
    sums[0, 0] = np.nan; sums[0, 1, 1] = np.inf; sums[0, 2] = 0; sums[0, 3] = -1.0
    rgb8 = np.empty((32, 48, 3), np.uint8)
    ref.check(ref.fn("film_resolve")(abi._ptr(sums), 48, 32, 16, abi._ptr(rgb8)), "film_resolve")

    with tempfile.TemporaryDirectory() as tmp:
        rgbe = scenes.hall_env(64, 32, 1234)
        hdr = os.path.join(tmp, "env.hdr"); scenes.write_hdr(hdr, rgbe)
        env = scenes.rgbe_to_float(rgbe)
        sd = abi.SceneDesc()
        a, b = sd.solid((0.9, 0.1, 0.2)), sd.solid((0.1, 0.2, 0.9))
        chk = sd.checker(a, b)
        img = rng.integers(0, 256, (16, 24, 3), dtype=np.uint8)
        timg = sd.image(img)
        tenv = sd.env(env, hdr)
        sd.add(sd.sphere((0, 0, 0), 1, sd.material(abi.MAT_LAMBERTIAN, albedo_tex=chk)))
        sc = sd.build(ref, abi.BVH_REFERENCE, 3)
        uvp = rng.random((3000, 5)).astype(F) * 1.4 - 0.2; uvp[:, 2:] *= 3
        tex = {f"tex_{n}": sc.texture_eval(t, uvp) for n, t in (("checker", chk), ("image", timg), ("env", tenv))}
        np.savez_compressed(os.path.join(HERE, "host_kats.npz"), cam_st=st, cam_rays=cam_out, film_sums=sums, film_spp=np.int32(16), film_rgb8=rgb8,
                            tex_uvp=uvp, tex_image=img, tex_envmap=env, **{k + "_out": v for k, v in tex.items()})

        # ---- a converged render of the reference's own rayColour loop (rand() stream, 2 seeds) ----------------------
        fx = dict(fx_common); fx["env"] = env
        sd = cases.scene_render(fx, hdr)
        sc = sd.build(ref, abi.BVH_REFERENCE, 3, tagged=False)
        c = scenes.camera_for(cases.RENDER_CAMERA, cases.RENDER_W, cases.RENDER_H)
        a, rays_a = sc.render_f32(c, cases.RENDER_W, cases.RENDER_H, cases.RENDER_SPP, seed=1, threads=8)
        b, rays_b = sc.render_f32(c, cases.RENDER_W, cases.RENDER_H, cases.RENDER_SPP, seed=2, threads=8)
        fx.update(sum_a=a, sum_b=b, rays_a=np.int64(rays_a), rays_b=np.int64(rays_b), spp=np.int32(cases.RENDER_SPP))
        np.savez_compressed(os.path.join(HERE, "render_small.npz"), **fx)
        ma, mb = a / cases.RENDER_SPP, b / cases.RENDER_SPP
        print("render: mean", float(np.nanmean(ma)), "rmse(a,b)", float(np.sqrt(np.nanmean((ma - mb) ** 2))), "rays/sample", rays_a / (cases.RENDER_W * cases.RENDER_H * cases.RENDER_SPP))


if __name__ == "__main__":
    main()
