"""ctypes view of the C ABI declared in include/hrt.h.

`Lib(path, prefix)` binds one shared library that implements the scene-builder ABI.  The product library is
`hobbyraytracer_b200/csrc/libhrt_b200.so` with prefix ``hrt_``; the test-only oracle libraries export the same
signatures under ``orc_`` / ``ref_`` and are bound by tests with this same class (the package itself never loads
them).  `SceneDesc` is a plain description of a scene in the reference's own constructor vocabulary
(scene.cpp:127-374) that can be replayed onto any bound library.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

MISS = 0xFFFFFFFF

MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_ISOTROPIC, MAT_DIFFUSE_LIGHT, MAT_PBR, MAT_UVTEST = range(7)
XF_TRANSLATE, XF_SCALE, XF_ROTATE_QUAT, XF_ROTATE_Y, XF_ROTATE_EULER_DEG, XF_ROTATE_Y_DEG = range(6)
BVH_REFERENCE, BVH_SAH, BVH_LBVH_GPU, BVH_LBVH_HOST = range(4)
RENDER_MEGAKERNEL = 1
RENDER_WAVEFRONT_V1 = 2
RENDER_DEFOCUS = 4
RENDER_RR = 8
MESH_BORROW = 1


PROGRESS_FN = C.CFUNCTYPE(C.c_int, C.c_int, C.c_int, C.c_void_p)      # hrt_progress_fn


class Material(C.Structure):
    _fields_ = [("kind", C.c_int32), ("albedo", C.c_float * 3), ("albedo_tex", C.c_int32),
                ("p0", C.c_float), ("p0_tex", C.c_int32), ("p1", C.c_float), ("p1_tex", C.c_int32)]


class Xform(C.Structure):
    _fields_ = [("kind", C.c_int32), ("v", C.c_float * 4)]


class Camera(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("look_at", C.c_float * 3), ("up", C.c_float * 3),
                ("vfov_deg", C.c_float), ("aspect", C.c_float), ("aperture", C.c_float), ("focal_distance", C.c_float)]


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("sample_first", C.c_int32),
                ("sample_count", C.c_int32), ("max_depth", C.c_int32), ("t_min", C.c_float), ("seed", C.c_uint64),
                ("rng_mode", C.c_int32), ("flags", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("paths", C.c_uint64), ("launches", C.c_uint64), ("ms_trace", C.c_double),
                ("ms_shade", C.c_double), ("ms_other", C.c_double), ("ms_total", C.c_double),
                ("film_clamped", C.c_uint64), ("prim_tests", C.c_uint64), ("retraced", C.c_uint64),
                ("scene_bytes", C.c_uint64), ("ms_build", C.c_double),
                ("h2d_bytes", C.c_uint64), ("ms_build_upload", C.c_double)]


def make_camera(position, look_at, up, vfov_deg, aspect, aperture=0.0, focal_distance=1.0) -> Camera:
    c = Camera()
    c.position[:] = [float(x) for x in position]
    c.look_at[:] = [float(x) for x in look_at]
    c.up[:] = [float(x) for x in up]
    c.vfov_deg = float(vfov_deg)
    c.aspect = float(np.float32(aspect))
    c.aperture = float(aperture)
    c.focal_distance = float(focal_distance)
    return c


class HrtError(RuntimeError):
    pass


_vp = C.c_void_p


class Lib:
    """One shared library exporting `<prefix>scene_*` etc."""

    def __init__(self, path: str, prefix: str):
        if not os.path.exists(path):
            raise HrtError(f"shared library not found: {path}")
        self.path, self.prefix = path, prefix
        self.dll = C.CDLL(path)
        self._bind()

    def has(self, name: str) -> bool:
        return hasattr(self.dll, self.prefix + name)

    def fn(self, name: str):
        return getattr(self.dll, self.prefix + name)

    def _sig(self, name, restype, argtypes, required=True):
        if not self.has(name):
            if required:
                raise HrtError(f"{self.path} does not export {self.prefix}{name}")
            return
        f = self.fn(name)
        f.restype, f.argtypes = restype, argtypes

    def _bind(self):
        i, f32, i64, u64 = C.c_int, C.c_float, C.c_int64, C.c_uint64
        self._sig("last_error", C.c_char_p, [])
        self._sig("scene_create", _vp, [])
        self._sig("scene_destroy", None, [_vp])
        self._sig("scene_add_texture_solid", i, [_vp, _vp])
        self._sig("scene_add_texture_checker", i, [_vp, i, i])
        self._sig("scene_add_texture_image", i, [_vp, _vp, i, i])
        self._sig("scene_add_texture_env", i, [_vp, _vp, i, i, i], required=False)
        self._sig("scene_add_texture_env_file", i, [_vp, C.c_char_p], required=False)
        self._sig("scene_add_material", i, [_vp, C.POINTER(Material)])
        self._sig("scene_add_sphere", i, [_vp, _vp, f32, i])
        self._sig("scene_add_rect", i, [_vp, i, f32, f32, f32, f32, f32, i])
        self._sig("scene_add_box", i, [_vp, _vp, _vp, i])
        self._sig("scene_add_mesh", i, [_vp, _vp, _vp, _vp, i64, i])
        self._sig("scene_add_mesh_indexed", i, [_vp, _vp, _vp, _vp, i64, _vp, i64, _vp, i, i, C.c_uint32], required=False)
        self._sig("scene_prim_base", i64, [_vp, i])
        self._sig("scene_prim_count", i64, [_vp])
        self._sig("scene_add_object", i, [_vp, i, _vp, i])
        self._sig("scene_add_medium", i, [_vp, i, _vp, i, f32, _vp, i])
        self._sig("scene_set_background", i, [_vp, i])
        self._sig("scene_commit", i, [_vp, i, u64])
        self._sig("scene_set_tagged", i, [_vp, i], required=False)
        self._sig("trace_batch", i, [_vp, _vp, i64, _vp, _vp, _vp, _vp, _vp, u64])
        self._sig("trace_batch_device", i, [_vp, _vp, i64, _vp, _vp, u64, _vp], required=False)
        self._sig("trace_batch_static", i, [_vp, _vp, i64, _vp, _vp, u64, _vp], required=False)
        # product-only
        self._sig("abi_version", i, [], required=False)
        self._sig("trim", None, [], required=False)
        self._sig("device_count", i, [], required=False)
        self._sig("set_device", i, [i], required=False)
        self._sig("render", i, [_vp, C.POINTER(Camera), C.POINTER(RenderParams), _vp, _vp], required=False)
        self._sig("render_multi", i, [_vp, _vp, i, C.POINTER(Camera), C.POINTER(RenderParams), _vp, _vp], required=False)
        self._sig("render_progressive", i, [_vp, C.POINTER(Camera), C.POINTER(RenderParams), C.c_char_p, i, PROGRESS_FN, _vp, _vp, _vp], required=False)
        self._sig("checkpoint_samples", i, [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)], required=False)
        self._sig("render_accumulate_device", i, [_vp, C.POINTER(Camera), C.POINTER(RenderParams), _vp, _vp], required=False)
        self._sig("film_resolve_device", i, [_vp, _vp, i, i, i, _vp, _vp, _vp, _vp], required=False)
        self._sig("film_bytes", i64, [i, i], required=False)
        self._sig("peer_alloc", _vp, [i64], required=False)
        self._sig("peer_free", None, [_vp], required=False)
        self._sig("peer_handle", i, [_vp, _vp], required=False)
        self._sig("peer_open", _vp, [_vp, _vp], required=False)
        self._sig("film_resolve_slice", i, [_vp, _vp, i, i, i, i, i, i, _vp, _vp, _vp, _vp], required=False)
        self._sig("film_create", _vp, [i, i], required=False)
        self._sig("film_destroy", None, [_vp], required=False)
        self._sig("film_zero", i, [_vp, i, i, _vp], required=False)
        self._sig("film_ipc_handle", i, [_vp, _vp], required=False)
        self._sig("film_resolve_peers", i, [_vp, _vp, _vp, i, i, i, i, i, _vp, _vp, _vp, _vp], required=False)
        self._sig("get_stats", i, [_vp, C.POINTER(Stats)], required=False)
        self._sig("reset_stats", i, [_vp], required=False)
        # oracle-only
        self._sig("render_u8", i, [_vp, C.POINTER(Camera), i, i, i, u64, _vp], required=False)
        self._sig("render_f32", i, [_vp, C.POINTER(Camera), i, i, i, u64, i, _vp, _vp], required=False)
        self._sig("render_f32_defocus", i, [_vp, C.POINTER(Camera), i, i, i, u64, i, _vp, _vp], required=False)
        self._sig("film_resolve", i, [_vp, i, i, i, _vp], required=False)
        self._sig("camera_rays", i, [C.POINTER(Camera), _vp, i64, _vp], required=False)
        self._sig("rng_draw", i, [i, u64, u64, i64, _vp], required=False)
        if self.prefix == "hrt_":     # the product's entry also takes the rng mode of hrt_render_params
            self._sig("scatter_batch", i, [_vp, i, _vp, _vp, i64, u64, i, _vp], required=False)
        else:
            self._sig("scatter_batch", i, [_vp, i, _vp, _vp, i64, u64, _vp], required=False)
        self._sig("texture_eval", i, [_vp, i, _vp, i64, _vp], required=False)
        self._sig("mesh_bvh_nodes", i64, [_vp, i, _vp, _vp], required=False)
        self._sig("scene_mesh_array", i64, [_vp, i, i, _vp, i64], required=False)
        self._sig("hardware_threads", i, [], required=False)

    def check(self, rc: int, what: str = "") -> int:
        if rc < 0:
            msg = self.fn("last_error")()
            raise HrtError(f"{self.prefix}{what} failed ({rc}): {msg.decode() if msg else ''}")
        return rc


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data


@dataclass
class SceneDesc:
    """A scene as a list of reference constructor calls; replay with `.build(lib)`."""
    textures: list = field(default_factory=list)      # ("solid", rgb) | ("checker", even, odd) | ("image", u8[h,w,3]) | ("env", f32[h,w,c]) | ("env_file", path, f32[h,w,c])
    materials: list = field(default_factory=list)     # dict(kind=, albedo=, albedo_tex=, p0=, p0_tex=, p1=, p1_tex=)
    borrow: bool = False                              # mesh_indexed shapes: pass HRT_MESH_BORROW (the arrays live in self.shapes until build() returns)
    flatten_for_all: bool = False                     # mesh_indexed shapes: flatten in numpy even for libraries that have the indexed entry
    shapes: list = field(default_factory=list)        # ("sphere", c, r, mat) | ("rect", axis, a0,a1,b0,b1,k, mat) | ("box", center, dims, mat) | ("mesh", pos, nrm, uv, mat)
    objects: list = field(default_factory=list)       # ("object", shape, ops) | ("medium", shape, ops, density, rgb, tex)
    background: Optional[int] = None

    # -- authoring helpers -------------------------------------------------------------------------------
    def solid(self, rgb) -> int:
        self.textures.append(("solid", tuple(float(x) for x in rgb))); return len(self.textures) - 1

    def checker(self, even_tex: int, odd_tex: int) -> int:
        self.textures.append(("checker", even_tex, odd_tex)); return len(self.textures) - 1

    def image(self, u8: np.ndarray) -> int:
        self.textures.append(("image", np.ascontiguousarray(u8, dtype=np.uint8))); return len(self.textures) - 1

    def env(self, f32: np.ndarray, path: Optional[str] = None) -> int:
        self.textures.append(("env", _f32(f32), path)); return len(self.textures) - 1

    def material(self, kind, albedo=(0, 0, 0), albedo_tex=-1, p0=0.0, p0_tex=-1, p1=0.0, p1_tex=-1) -> int:
        self.materials.append(dict(kind=kind, albedo=tuple(float(x) for x in albedo), albedo_tex=albedo_tex,
                                   p0=float(p0), p0_tex=p0_tex, p1=float(p1), p1_tex=p1_tex))
        return len(self.materials) - 1

    def sphere(self, c, r, mat) -> int:
        self.shapes.append(("sphere", tuple(float(x) for x in c), float(r), mat)); return len(self.shapes) - 1

    def rect(self, axis, a0, a1, b0, b1, k, mat) -> int:
        self.shapes.append(("rect", int(axis), float(a0), float(a1), float(b0), float(b1), float(k), mat)); return len(self.shapes) - 1

    def box(self, center, dims, mat) -> int:
        self.shapes.append(("box", tuple(float(x) for x in center), tuple(float(x) for x in dims), mat)); return len(self.shapes) - 1

    def box_minmax(self, mn, mx, mat) -> int:
        """Box::minMaxBox (box.h:11-14): centre = (min+max)/2, dimensions = max-min, in fp32."""
        mn, mx = np.asarray(mn, np.float32), np.asarray(mx, np.float32)
        return self.box((mn + mx) / np.float32(2.0), mx - mn, mat)

    def mesh(self, pos, nrm, uv, mat) -> int:
        pos = _f32(pos).reshape(-1, 9)
        nrm = None if nrm is None else _f32(nrm).reshape(-1, 9)
        uv = None if uv is None else _f32(uv).reshape(-1, 6)
        self.shapes.append(("mesh", pos, nrm, uv, mat)); return len(self.shapes) - 1

    def mesh_indexed(self, vertices, normals, uvs, indices, mat, submeshes=None) -> int:
        """hrt_scene_add_mesh_indexed: per-vertex arrays + an index buffer as Mesh::assimpLoadFile concatenates them
        (mesh.cpp:63-115); submeshes = [(first_index, n_index, base_vertex), ...] applies the base-vertex offset the reference
        leaves out (mesh.cpp:111-114).  Libraries without the entry point (oracle, reference harness) get the triangles
        flattened here exactly like Mesh::Mesh's loop (mesh.cpp:24-33)."""
        v = _f32(vertices).reshape(-1, 3)
        n = None if normals is None else _f32(normals).reshape(-1, 3)
        t = None if uvs is None else _f32(uvs).reshape(-1, 2)
        idx = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        sub = None if submeshes is None else np.ascontiguousarray(submeshes, dtype=np.int64).reshape(-1, 3)
        self.shapes.append(("mesh_indexed", v, n, t, idx, sub, mat)); return len(self.shapes) - 1

    @staticmethod
    def flatten_indexed(v, n, t, idx, sub):
        """numpy restatement of Mesh::Mesh's gather (mesh.cpp:24-33) -> (pos [T,9], nrm [T,9], uv [T,6])"""
        idx = idx.astype(np.int64)
        if sub is not None:
            idx = idx.copy()
            for first, count, base in sub:
                idx[first:first + count] += base
        if n is None:
            n = np.zeros_like(v)
        if t is None:
            t = np.zeros((v.shape[0], 2), np.float32)
        return (np.ascontiguousarray(v[idx].reshape(-1, 9)), np.ascontiguousarray(n[idx].reshape(-1, 9)),
                np.ascontiguousarray(t[idx].reshape(-1, 6)))

    def add(self, shape: int, ops: Sequence[Tuple] = ()) -> None:
        self.objects.append(("object", shape, list(ops)))

    def add_medium(self, shape: int, ops: Sequence[Tuple], density: float, rgb=(1, 1, 1), tex: int = -1) -> None:
        self.objects.append(("medium", shape, list(ops), float(density), tuple(float(x) for x in rgb), tex))

    # -- replay --------------------------------------------------------------------------------------------
    def build(self, lib: Lib, bvh_mode: int = BVH_REFERENCE, seed: int = 1, tagged: Optional[bool] = None) -> "Scene":
        h = lib.fn("scene_create")()
        if not h:
            raise HrtError(f"{lib.prefix}scene_create failed")
        sc = Scene(lib, h)
        if tagged is not None and lib.has("scene_set_tagged"):
            lib.fn("scene_set_tagged")(h, 1 if tagged else 0)
        for t in self.textures:
            if t[0] == "solid":
                a = (C.c_float * 3)(*t[1]); lib.check(lib.fn("scene_add_texture_solid")(h, a), "add_texture_solid")
            elif t[0] == "checker":
                lib.check(lib.fn("scene_add_texture_checker")(h, t[1], t[2]), "add_texture_checker")
            elif t[0] == "image":
                img = t[1]; hh, ww = (img.shape[0], img.shape[1]) if img.size else (0, 0)
                lib.check(lib.fn("scene_add_texture_image")(h, _ptr(img) if img.size else None, ww, hh), "add_texture_image")
            elif t[0] == "env":
                data, path = t[1], t[2]
                if lib.has("scene_add_texture_env"):
                    hh, ww, ch = data.shape
                    lib.check(lib.fn("scene_add_texture_env")(h, _ptr(data), ww, hh, ch), "add_texture_env")
                else:   # the unmodified reference can only load environment maps from files
                    if path is None:
                        raise HrtError("this library needs an .hdr file path for environment maps")
                    lib.check(lib.fn("scene_add_texture_env_file")(h, path.encode()), "add_texture_env_file")
        for m in self.materials:
            mm = Material(); mm.kind = m["kind"]; mm.albedo[:] = m["albedo"]; mm.albedo_tex = m["albedo_tex"]
            mm.p0 = m["p0"]; mm.p0_tex = m["p0_tex"]; mm.p1 = m["p1"]; mm.p1_tex = m["p1_tex"]
            lib.check(lib.fn("scene_add_material")(h, C.byref(mm)), "add_material")
        for s in self.shapes:
            if s[0] == "sphere":
                c = (C.c_float * 3)(*s[1]); lib.check(lib.fn("scene_add_sphere")(h, c, s[2], s[3]), "add_sphere")
            elif s[0] == "rect":
                lib.check(lib.fn("scene_add_rect")(h, *s[1:]), "add_rect")
            elif s[0] == "box":
                c = (C.c_float * 3)(*s[1]); d = (C.c_float * 3)(*s[2]); lib.check(lib.fn("scene_add_box")(h, c, d, s[3]), "add_box")
            elif s[0] == "mesh":
                _, pos, nrm, uv, mat = s
                lib.check(lib.fn("scene_add_mesh")(h, _ptr(pos), _ptr(nrm), _ptr(uv), pos.shape[0], mat), "add_mesh")
            elif s[0] == "mesh_indexed":
                _, v, n, t, idx, sub, mat = s
                if lib.has("scene_add_mesh_indexed") and not self.flatten_for_all:
                    lib.check(lib.fn("scene_add_mesh_indexed")(h, _ptr(v), _ptr(n), _ptr(t), v.shape[0], _ptr(idx), idx.shape[0],
                                                                _ptr(sub), 0 if sub is None else sub.shape[0], mat,
                                                                MESH_BORROW if self.borrow else 0), "add_mesh_indexed")   # self.shapes keeps the arrays alive
                else:
                    pos, nrm, uv = self.flatten_indexed(v, n, t, idx, sub)
                    lib.check(lib.fn("scene_add_mesh")(h, _ptr(pos), _ptr(nrm), _ptr(uv), pos.shape[0], mat), "add_mesh")
        for o in self.objects:
            ops = o[2]
            arr = (Xform * max(1, len(ops)))()
            for k, op in enumerate(ops):
                arr[k].kind = op[0]
                vals = list(op[1]) + [0.0] * (4 - len(op[1]))
                arr[k].v[:] = [float(x) for x in vals]
            if o[0] == "object":
                lib.check(lib.fn("scene_add_object")(h, o[1], arr, len(ops)), "add_object")
            else:
                rgb = (C.c_float * 3)(*o[4])
                lib.check(lib.fn("scene_add_medium")(h, o[1], arr, len(ops), o[3], rgb, o[5]), "add_medium")
        if self.background is not None:
            lib.check(lib.fn("scene_set_background")(h, self.background), "set_background")
        lib.check(lib.fn("scene_commit")(h, bvh_mode, seed), "scene_commit")
        return sc


class Scene:
    """A built scene on one library."""

    def __init__(self, lib: Lib, handle):
        self.lib, self.h = lib, handle

    def close(self):
        if self.h:
            self.lib.fn("scene_destroy")(self.h); self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def prim_base(self, shape: int) -> int:
        return int(self.lib.fn("scene_prim_base")(self.h, shape))

    def prim_count(self) -> int:
        return int(self.lib.fn("scene_prim_count")(self.h))

    def trace(self, rays: np.ndarray, rng_seed: int = 0, extras: bool = True):
        """world->hit on n rays (n x 8: o, t_min, d, t_max). Returns dict(prim, t[, normal, uv, front])."""
        rays = _f32(rays).reshape(-1, 8)
        n = rays.shape[0]
        prim = np.empty(n, np.uint32); t = np.empty(n, np.float32)
        nrm = np.empty((n, 3), np.float32) if extras else None
        uv = np.empty((n, 2), np.float32) if extras else None
        fr = np.empty(n, np.uint8) if extras else None
        self.lib.check(self.lib.fn("trace_batch")(self.h, _ptr(rays), n, _ptr(prim), _ptr(t), _ptr(nrm), _ptr(uv), _ptr(fr), rng_seed), "trace_batch")
        out = dict(prim=prim, t=t)
        if extras:
            out.update(normal=nrm, uv=uv, front=fr)
        return out

    # product
    def render(self, cam: Camera, width: int, height: int, spp: int, seed: int = 1, sample_first: int = 0,
               sample_count: int = 0, max_depth: int = 50, t_min: float = 0.001, rng_mode: int = 0, flags: int = 0,
               want_sum: bool = True, want_rgb8: bool = True):
        p = RenderParams(width, height, spp, sample_first, sample_count, max_depth, t_min, seed, rng_mode, flags)
        s = np.empty((height, width, 3), np.float32) if want_sum else None
        b = np.empty((height, width, 3), np.uint8) if want_rgb8 else None
        self.lib.check(self.lib.fn("render")(self.h, C.byref(cam), C.byref(p), _ptr(s), _ptr(b)), "render")
        return s, b

    def render_multi(self, devices, cam: Camera, width: int, height: int, spp: int, seed: int = 1, max_depth: int = 50, t_min: float = 0.001,
                     rng_mode: int = 0, flags: int = 0):
        """hrt_render_multi: one process, the GPUs in `devices` (devices[0] = the scene's own) -> (sum, rgb8)"""
        p = RenderParams(width, height, spp, 0, 0, max_depth, t_min, seed, rng_mode, flags)
        d = (C.c_int * len(devices))(*devices)
        s = np.empty((height, width, 3), np.float32); b = np.empty((height, width, 3), np.uint8)
        self.lib.check(self.lib.fn("render_multi")(self.h, d, len(devices), C.byref(cam), C.byref(p), _ptr(s), _ptr(b)), "render_multi")
        return s, b

    def render_progressive(self, cam: Camera, width: int, height: int, spp: int, seed: int = 1, checkpoint: Optional[str] = None,
                           slice_spp: int = 0, on_slice=None, max_depth: int = 50, t_min: float = 0.001, rng_mode: int = 0, flags: int = 0):
        """hrt_render_progressive -> (rc, sum, rgb8); rc 1 = stopped by on_slice(samples_done, spp) returning non-zero."""
        p = RenderParams(width, height, spp, 0, 0, max_depth, t_min, seed, rng_mode, flags)
        s = np.empty((height, width, 3), np.float32); b = np.empty((height, width, 3), np.uint8)
        cb = PROGRESS_FN(lambda done, total, user: int(on_slice(done, total) or 0)) if on_slice else PROGRESS_FN(0)
        rc = self.lib.fn("render_progressive")(self.h, C.byref(cam), C.byref(p), checkpoint.encode() if checkpoint else None,
                                               slice_spp, cb, None, _ptr(s), _ptr(b))
        self.lib.check(rc, "render_progressive")
        return rc, s, b

    def mesh_array(self, mesh_index: int, which: int) -> np.ndarray:
        """One per-mesh array of the committed scene as float32 words (hrt_scene_mesh_array): 0 reference-topology nodes,
        1 fast-tree nodes, 2 leaf records, 3 triangle positions, 4 triangle attributes."""
        f = self.lib.fn("scene_mesh_array")
        n = f(self.h, mesh_index, which, None, 0)
        self.lib.check(n, "scene_mesh_array")
        out = np.empty(n // 4, np.float32)
        if n:
            self.lib.check(f(self.h, mesh_index, which, _ptr(out), n), "scene_mesh_array")
        return out

    def stats(self) -> Stats:
        st = Stats(); self.lib.check(self.lib.fn("get_stats")(self.h, C.byref(st)), "get_stats"); return st

    # oracle
    def render_u8(self, cam: Camera, width: int, height: int, spp: int, seed: int = 1) -> np.ndarray:
        b = np.empty((height, width, 3), np.uint8)
        self.lib.check(self.lib.fn("render_u8")(self.h, C.byref(cam), width, height, spp, seed, _ptr(b)), "render_u8")
        return b

    def render_f32(self, cam: Camera, width: int, height: int, spp: int, seed: int = 1, threads: int = 0):
        s = np.empty((height, width, 3), np.float32); rays = C.c_uint64(0)
        self.lib.check(self.lib.fn("render_f32")(self.h, C.byref(cam), width, height, spp, seed, threads, _ptr(s), C.byref(rays)), "render_f32")
        return s, int(rays.value)

    def render_f32_defocus(self, cam: Camera, width: int, height: int, spp: int, seed: int = 1, threads: int = 0):
        s = np.empty((height, width, 3), np.float32); rays = C.c_uint64(0)
        self.lib.check(self.lib.fn("render_f32_defocus")(self.h, C.byref(cam), width, height, spp, seed, threads, _ptr(s), C.byref(rays)), "render_f32_defocus")
        return s, int(rays.value)

    def scatter(self, material: int, inputs: np.ndarray, front: np.ndarray, seed: int = 0, rng_mode: int = 0) -> np.ndarray:
        inputs = _f32(inputs).reshape(-1, 14); front = np.ascontiguousarray(front, np.uint8)
        out = np.empty((inputs.shape[0], 10), np.float32)
        if self.lib.prefix == "hrt_":
            self.lib.check(self.lib.fn("scatter_batch")(self.h, material, _ptr(inputs), _ptr(front), inputs.shape[0], seed, rng_mode, _ptr(out)), "scatter_batch")
        else:
            self.lib.check(self.lib.fn("scatter_batch")(self.h, material, _ptr(inputs), _ptr(front), inputs.shape[0], seed, _ptr(out)), "scatter_batch")
        return out

    def texture_eval(self, tex: int, uvp: np.ndarray) -> np.ndarray:
        uvp = _f32(uvp).reshape(-1, 5); out = np.empty((uvp.shape[0], 3), np.float32)
        self.lib.check(self.lib.fn("texture_eval")(self.h, tex, _ptr(uvp), uvp.shape[0], _ptr(out)), "texture_eval")
        return out
