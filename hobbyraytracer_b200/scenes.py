"""Procedural stand-ins for the BASELINE.json configs (their assets — teapot.obj, old_hall_4k.hdr — are not in the
reference repository) and small variants of them for parity tests.

Everything is seeded and built with numpy only; scenes are returned as `abi.SceneDesc` + `abi.Camera`.
"""
from __future__ import annotations

import math
import os
import struct
from typing import Optional, Tuple

import numpy as np

from . import abi
from .abi import SceneDesc

F = np.float32


# ---------------------------------------------------------------------------------------------------------------
# meshes (per-triangle-corner arrays: pos n x 9, nrm n x 9, uv n x 6 — what mesh.cpp:24-35 feeds ITriangle)
# ---------------------------------------------------------------------------------------------------------------
def grid_mesh(nu: int, nv: int, fn, flip: bool = False, indexed: bool = False):
    """Tessellate a parametric surface p(u,v), n(u,v) on an nu x nv quad grid -> 2*nu*nv triangles.
    indexed: -> (vertices [V,3], normals [V,3], uvs [V,2], indices [6*nu*nv] u32) whose flattening (mesh.cpp:24-33) is
    exactly the triangle soup the default form returns."""
    u = np.linspace(0.0, 1.0, nu + 1)
    v = np.linspace(0.0, 1.0, nv + 1)
    uu, vv = np.meshgrid(u, v, indexing="ij")
    P, N = fn(uu, vv)                       # (nu+1, nv+1, 3)
    UV = np.stack([uu, vv], -1)

    def corners(a):
        a00, a10, a01, a11 = a[:-1, :-1], a[1:, :-1], a[:-1, 1:], a[1:, 1:]
        if flip:
            t0 = np.stack([a00, a01, a10], -2); t1 = np.stack([a10, a01, a11], -2)
        else:
            t0 = np.stack([a00, a10, a01], -2); t1 = np.stack([a10, a11, a01], -2)
        t = np.stack([t0, t1], 2)           # (nu, nv, 2, 3, c)
        return t.reshape(-1, 3 * a.shape[-1])

    if indexed:
        ids = np.arange((nu + 1) * (nv + 1), dtype=np.int64).reshape(nu + 1, nv + 1, 1)
        return (P.reshape(-1, 3).astype(F), N.reshape(-1, 3).astype(F), UV.reshape(-1, 2).astype(F),
                corners(ids).reshape(-1).astype(np.uint32))
    return corners(P).astype(F), corners(N).astype(F), corners(UV).astype(F)


def _value_noise(x, y, seed, octaves=4):
    rng = np.random.default_rng(seed)
    out = np.zeros_like(x)
    amp, freq = 1.0, 4.0
    for _ in range(octaves):
        n = int(freq) + 2
        g = rng.random((n, n))
        xi, yi = x * freq, y * freq
        x0, y0 = np.floor(xi).astype(int) % (n - 1), np.floor(yi).astype(int) % (n - 1)
        fx, fy = xi - np.floor(xi), yi - np.floor(yi)
        fx, fy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
        a = g[x0, y0] * (1 - fx) + g[x0 + 1, y0] * fx
        b = g[x0, y0 + 1] * (1 - fx) + g[x0 + 1, y0 + 1] * fx
        out += amp * (a * (1 - fy) + b * fy)
        amp *= 0.5; freq *= 2.0
    return out


def blob_mesh(nu: int, nv: int, radius: float = 1.0, bump: float = 0.15, seed: int = 7, center=(0.0, 0.0, 0.0), indexed: bool = False):
    """A noise-displaced UV sphere ("teapot-scale" / "statue-scale" stand-in). 2*nu*nv triangles.
    Normals are the analytic sphere normals (smooth shading)."""
    def fn(uu, vv):
        theta = uu * 2.0 * math.pi
        phi = (vv * 0.98 + 0.01) * math.pi          # avoid degenerate pole triangles
        d = np.stack([np.sin(phi) * np.cos(theta), np.cos(phi), np.sin(phi) * np.sin(theta)], -1)
        nz = _value_noise(uu, vv, seed)
        nz[-1, :] = nz[0, :]
        r = radius * (1.0 + bump * (nz - 1.0))
        return d * r[..., None] + np.asarray(center)[None, None, :], d
    return grid_mesh(nu, nv, fn, indexed=indexed)


def teapot_like_mesh(n: int = 56, seed: int = 7):
    """~2*n*n triangles, roughly the size of the classic Utah teapot tessellation (6,320 faces at n=56),
    sitting on y=0 with height ~1.6 like teapot.obj in the reference's sample scenes."""
    pos, nrm, uv = blob_mesh(n, n, radius=0.8, bump=0.25, seed=seed, center=(0.0, 0.8, 0.0))
    return pos, nrm, uv


def write_obj(path: str, pos: np.ndarray, nrm: Optional[np.ndarray], uv: Optional[np.ndarray]) -> None:
    """Unshared-vertex OBJ (v/vt/vn per corner). uv are stored flipped back (1-v) because importers flip them."""
    pos = pos.reshape(-1, 3)
    with open(path, "w") as f:
        for p in pos:
            f.write("v %.9g %.9g %.9g\n" % tuple(p))
        if uv is not None:
            for t in uv.reshape(-1, 2):
                f.write("vt %.9g %.9g\n" % (t[0], 1.0 - t[1]))
        if nrm is not None:
            for q in nrm.reshape(-1, 3):
                f.write("vn %.9g %.9g %.9g\n" % tuple(q))
        for i in range(pos.shape[0] // 3):
            a = 3 * i + 1
            if uv is not None and nrm is not None:
                f.write("f %d/%d/%d %d/%d/%d %d/%d/%d\n" % (a, a, a, a + 1, a + 1, a + 1, a + 2, a + 2, a + 2))
            elif nrm is not None:
                f.write("f %d//%d %d//%d %d//%d\n" % (a, a, a + 1, a + 1, a + 2, a + 2))
            else:
                f.write("f %d %d %d\n" % (a, a + 1, a + 2))


# ---------------------------------------------------------------------------------------------------------------
# environment map: procedural Radiance .hdr (stand-in for old_hall_4k.hdr)
# ---------------------------------------------------------------------------------------------------------------
def float_to_rgbe(rgb: np.ndarray) -> np.ndarray:
    rgb = np.maximum(rgb.astype(np.float64), 0.0)
    m = rgb.max(-1)
    e = np.zeros(m.shape, np.int64)
    nz = m > 1e-32
    mant, ex = np.frexp(m[nz])
    scale = np.zeros_like(m)
    scale[nz] = mant * 256.0 / m[nz]
    e[nz] = ex + 128
    out = np.zeros(rgb.shape[:-1] + (4,), np.uint8)
    out[..., :3] = np.clip(rgb * scale[..., None], 0, 255).astype(np.uint8)
    out[..., 3] = np.where(nz, e, 0).astype(np.uint8)
    return out


def rgbe_to_float(rgbe: np.ndarray) -> np.ndarray:
    """stb_image's stbi__hdr_convert (stb_image.h): f = ldexp(1, e - 136); rgb = byte * f; e == 0 -> 0."""
    e = rgbe[..., 3].astype(np.int32)
    f = np.ldexp(np.float32(1.0), e - 136).astype(F)
    out = rgbe[..., :3].astype(F) * f[..., None]
    out[e == 0] = 0
    return out.astype(F)


def write_hdr(path: str, rgbe: np.ndarray) -> None:
    h, w, _ = rgbe.shape
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n")
        f.write(("-Y %d +X %d\n" % (h, w)).encode())
        f.write(rgbe.tobytes())          # flat (non-RLE) scanlines are valid Radiance and decode in stb


def hall_env(w: int = 512, h: int = 256, seed: int = 1234) -> np.ndarray:
    """Sky gradient + ground + a small very bright 'window' (HDR range ~5e3) -> rgbe[h,w,4]."""
    rng = np.random.default_rng(seed)
    v = (np.arange(h) + 0.5) / h
    u = (np.arange(w) + 0.5) / w
    uu, vv = np.meshgrid(u, v)
    sky = np.stack([0.35 + 0.4 * (1 - vv), 0.45 + 0.4 * (1 - vv), 0.6 + 0.5 * (1 - vv)], -1)
    ground = np.stack([0.25 + 0 * vv, 0.2 + 0 * vv, 0.15 + 0 * vv], -1)
    img = np.where((vv < 0.55)[..., None], sky, ground)
    img *= (0.9 + 0.2 * rng.random((h, w, 1)))
    win = (np.abs(uu - 0.3) < 0.02) & (np.abs(vv - 0.3) < 0.04)
    img[win] = np.array([5000.0, 4500.0, 3800.0])
    lamp = ((uu - 0.75) ** 2 * 4 + (vv - 0.2) ** 2) < 0.002
    img[lamp] = np.array([60.0, 50.0, 30.0])
    return float_to_rgbe(img)


def env_texture(sd: SceneDesc, cache_dir: str, w: int = 512, h: int = 256, seed: int = 1234) -> int:
    os.makedirs(cache_dir, exist_ok=True)
    path = os.path.join(cache_dir, f"hall_{w}x{h}_{seed}.hdr")
    rgbe = hall_env(w, h, seed)
    if not os.path.exists(path):
        write_hdr(path, rgbe)
    return sd.env(rgbe_to_float(rgbe), path)


# ---------------------------------------------------------------------------------------------------------------
# scenes
# ---------------------------------------------------------------------------------------------------------------
def cornell_walls(sd: SceneDesc, light_strength: float = 4.5):
    """The six rects of sampleScenes/teapot_scene.yaml:38-79 (a 5x5x5 box around the origin, light at y=4.99)."""
    white = sd.material(abi.MAT_LAMBERTIAN, (0.73, 0.73, 0.73))
    red = sd.material(abi.MAT_LAMBERTIAN, (0.65, 0.05, 0.05))
    green = sd.material(abi.MAT_LAMBERTIAN, (0.12, 0.45, 0.15))
    light = sd.material(abi.MAT_DIFFUSE_LIGHT, (0.93, 0.75, 0.43), p0=light_strength)
    sd.add(sd.rect(0, 0, 5, -2.5, 2.5, 2.5, red))
    sd.add(sd.rect(0, 0, 5, -2.5, 2.5, -2.5, green))
    sd.add(sd.rect(1, -2.5, 2.5, -2.5, 2.5, 0, white))
    sd.add(sd.rect(1, -2.5, 2.5, -2.5, 2.5, 5, white))
    sd.add(sd.rect(2, -2.5, 2.5, 0, 5, -2.5, white))
    sd.add(sd.rect(1, -1, 1, -1, 1, 4.99, light))
    return white, red, green, light


def teapot_scene(cache_dir: str, n: int = 56, env_w: int = 512, env_h: int = 256) -> Tuple[SceneDesc, abi.Camera, dict]:
    """C1: sampleScenes/teapot_scene.yaml with the procedural teapot stand-in (mesh under rotate/translate/scale)."""
    sd = SceneDesc()
    sd.background = env_texture(sd, cache_dir, env_w, env_h)
    white, *_ = cornell_walls(sd)
    pos, nrm, uv = teapot_like_mesh(n)
    m = sd.mesh(pos, nrm, uv, white)
    sd.add(m, [(abi.XF_ROTATE_EULER_DEG, (0, 180, 0)), (abi.XF_SCALE, (1.4, 1.4, 1.4)), (abi.XF_TRANSLATE, (0, 1, 0))])
    cam = dict(position=(0, 2.5, 8.5), look_at=(0, 2.5, 0), up=(0, 1, 0), vfov_deg=45, aperture=0.001, focal_distance=8.5)
    return sd, cam, dict(name="teapot_scene", width=640, height=640, spp=100)


def shiny_teapot_scene(cache_dir: str, n: int = 56, env_w: int = 4096, env_h: int = 2048, rough_dielectric: bool = False):
    """C2: sampleScenes/shiny_teapot.yaml (metal 0.2 mesh lit by the HDR environment), 1920x1080."""
    sd = SceneDesc()
    sd.background = env_texture(sd, cache_dir, env_w, env_h)
    if rough_dielectric:
        shiny = sd.material(abi.MAT_DIELECTRIC, p0=1.5, p1=0.1)
    else:
        shiny = sd.material(abi.MAT_METAL, (0.73, 0.73, 0.73), p0=0.2)
    pos, nrm, uv = teapot_like_mesh(n)
    sd.add(sd.mesh(pos, nrm, uv, shiny))
    cam = dict(position=(0, 1, 3), look_at=(0, 1, 0), up=(0, 1, 0), vfov_deg=60, aperture=0.001, focal_distance=8.5)
    return sd, cam, dict(name="shiny_teapot", width=1920, height=1080, spp=1024)


def scattered_balls_scene(seed: int = 42, grid: int = 11):
    """C3: Ray Tracing in One Weekend final scene (~485 spheres at grid=11), camera (13,2,3)->0, fov 20, sky background."""
    rng = np.random.default_rng(seed)
    sd = SceneDesc()
    sd.background = sd.solid((0.7, 0.8, 1.0))
    ground = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.sphere((0, -1000, 0), 1000, ground))
    for a in range(-grid, grid):
        for b in range(-grid, grid):
            choose = rng.random()
            c = np.array([a + 0.9 * rng.random(), 0.2, b + 0.9 * rng.random()])
            if np.linalg.norm(c - np.array([4, 0.2, 0])) > 0.9:
                if choose < 0.8:
                    m = sd.material(abi.MAT_LAMBERTIAN, rng.random(3) * rng.random(3))
                elif choose < 0.95:
                    m = sd.material(abi.MAT_METAL, 0.5 + 0.5 * rng.random(3), p0=0.5 * rng.random())
                else:
                    m = sd.material(abi.MAT_DIELECTRIC, p0=1.5, p1=0.0)
                sd.add(sd.sphere(c, 0.2, m))
    sd.add(sd.sphere((0, 1, 0), 1.0, sd.material(abi.MAT_DIELECTRIC, p0=1.5, p1=0.0)))
    sd.add(sd.sphere((-4, 1, 0), 1.0, sd.material(abi.MAT_LAMBERTIAN, (0.4, 0.2, 0.1))))
    sd.add(sd.sphere((4, 1, 0), 1.0, sd.material(abi.MAT_METAL, (0.7, 0.6, 0.5), p0=0.0)))
    cam = dict(position=(13, 2, 3), look_at=(0, 0, 0), up=(0, 1, 0), vfov_deg=20, aperture=0.1, focal_distance=10.0)
    return sd, cam, dict(name="scattered_balls", width=1200, height=800, spp=500)


def cornell_smoke_scene(light_strength: float = 7.0, glass_sphere: bool = False):
    """C4: RTIOW 'Cornell smoke' — 555-unit box, two Box -> RotateY -> Translate instances wrapped in ConstantMedium."""
    sd = SceneDesc()
    sd.background = sd.solid((0, 0, 0))
    red = sd.material(abi.MAT_LAMBERTIAN, (0.65, 0.05, 0.05))
    white = sd.material(abi.MAT_LAMBERTIAN, (0.73, 0.73, 0.73))
    green = sd.material(abi.MAT_LAMBERTIAN, (0.12, 0.45, 0.15))
    light = sd.material(abi.MAT_DIFFUSE_LIGHT, (1, 1, 1), p0=light_strength)
    sd.add(sd.rect(0, 0, 555, 0, 555, 555, green))
    sd.add(sd.rect(0, 0, 555, 0, 555, 0, red))
    sd.add(sd.rect(1, 113, 443, 127, 432, 554, light))
    sd.add(sd.rect(1, 0, 555, 0, 555, 555, white))
    sd.add(sd.rect(1, 0, 555, 0, 555, 0, white))
    sd.add(sd.rect(2, 0, 555, 0, 555, 555, white))
    b1 = sd.box_minmax((0, 0, 0), (165, 330, 165), white)
    b2 = sd.box_minmax((0, 0, 0), (165, 165, 165), white)
    sd.add_medium(b1, [(abi.XF_ROTATE_Y_DEG, (15,)), (abi.XF_TRANSLATE, (265, 0, 295))], 0.01, (0, 0, 0))
    sd.add_medium(b2, [(abi.XF_ROTATE_Y_DEG, (-18,)), (abi.XF_TRANSLATE, (130, 0, 65))], 0.01, (1, 1, 1))
    if glass_sphere:
        sd.add(sd.sphere((400, 90, 120), 70, sd.material(abi.MAT_DIELECTRIC, p0=1.5, p1=0.1)))
    cam = dict(position=(278, 278, -800), look_at=(278, 278, 0), up=(0, 1, 0), vfov_deg=40, aperture=0.0, focal_distance=10.0)
    return sd, cam, dict(name="cornell_smoke", width=1024, height=1024, spp=4096)


def big_mesh_scene(cache_dir: str, quads: int = 2290, env_w: int = 1024, env_h: int = 512, indexed: bool = False):
    """C5: one noise-displaced sphere with 2*quads^2 triangles (10,488,200 at quads=2290), metal 0.2, env-lit, 3840x2160.
    indexed: the same triangles handed over as vertex / index buffers (hrt_scene_add_mesh_indexed)."""
    sd = SceneDesc()
    sd.background = env_texture(sd, cache_dir, env_w, env_h)
    shiny = sd.material(abi.MAT_METAL, (0.73, 0.73, 0.73), p0=0.2)
    if indexed:
        v, n, t, idx = blob_mesh(quads, quads, radius=1.0, bump=0.2, seed=7, center=(0, 1, 0), indexed=True)
        sd.add(sd.mesh_indexed(v, n, t, idx, shiny)); sd.borrow = True       # the arrays stay in sd.shapes: no copy at add time
    else:
        pos, nrm, uv = blob_mesh(quads, quads, radius=1.0, bump=0.2, seed=7, center=(0, 1, 0))
        sd.add(sd.mesh(pos, nrm, uv, shiny))
    cam = dict(position=(0, 1, 3), look_at=(0, 1, 0), up=(0, 1, 0), vfov_deg=60, aperture=0.0, focal_distance=8.5)
    return sd, cam, dict(name="big_mesh", width=3840, height=2160, spp=64)


def camera_for(cam: dict, width: int, height: int) -> abi.Camera:
    aspect = np.float32(width) / np.float32(height)          # Film::getAspectRatio film.h:20-22
    return abi.make_camera(cam["position"], cam["look_at"], cam["up"], cam["vfov_deg"], aspect,
                           cam.get("aperture", 0.0), cam["focal_distance"])


# ---------------------------------------------------------------------------------------------------------------
# ray batches
# ---------------------------------------------------------------------------------------------------------------
def random_rays(n: int, center, radius: float, seed: int = 0xC0FFEE, t_min: float = 1e-3, t_max: float = np.inf) -> np.ndarray:
    """Origins uniform in a ball of `radius` around `center`, directions uniform on S^2 (SURVEY.md §8d, 16 M-ray batch)."""
    rng = np.random.default_rng(seed)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = rng.normal(size=(n, 3)); o /= np.linalg.norm(o, axis=1, keepdims=True)
    o *= (rng.random((n, 1)) ** (1.0 / 3.0)) * radius
    o += np.asarray(center)[None, :]
    rays = np.empty((n, 8), F)
    rays[:, 0:3] = o; rays[:, 3] = t_min; rays[:, 4:7] = d; rays[:, 7] = t_max
    return rays


def camera_rays(cam: dict, width: int, height: int, n: int, seed: int = 1) -> np.ndarray:
    """Unnormalised primary rays computed in float64 then rounded (inputs for hit() parity, not a camera test)."""
    rng = np.random.default_rng(seed)
    pos, at, up = (np.asarray(cam[k], np.float64) for k in ("position", "look_at", "up"))
    h = math.tan(math.radians(cam["vfov_deg"]) / 2); vh = 2 * h; vw = width / height * vh
    w = pos - at; w /= np.linalg.norm(w); u = np.cross(up, w); u /= np.linalg.norm(u); v = np.cross(w, u)
    fd = cam["focal_distance"]
    hor, ver = fd * vw * u, fd * vh * v
    llc = pos - hor / 2 - ver / 2 - fd * w
    s, t = rng.random((n, 1)), rng.random((n, 1))
    d = llc + s * hor + t * ver - pos
    rays = np.empty((n, 8), F)
    rays[:, 0:3] = pos; rays[:, 3] = 1e-3; rays[:, 4:7] = d; rays[:, 7] = np.inf
    return rays
