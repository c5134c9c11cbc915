"""Host-side scene loader: the reference's YAML schema (Scene::loadScene, scene.cpp:127-374) replayed into the
C ABI's scene builder.  It mirrors the loader statement by statement — same required keys, same lookup rules (an albedo that
is not a sequence is a texture name, else an image path, scene.cpp:77-99; a background that is not a sequence is a texture
name, else an .hdr path, scene.cpp:215-237), same wrapper order rotate -> scale -> translate (scene.cpp:335-353), unknown
materials skip the object (scene.cpp:281-290) — so that the SAME .yaml file can be handed to the reference's own loader
(oracle/ref_harness.cpp: ref_scene_load_yaml) and to the CUDA library, and their hit() answers compared bit for bit
(tests/test_yaml_dropin.py).

Extension (keys the shipped loader does not know, for the classes it cannot reach):
  materials: type dielectric (ior, roughness) | isotropic | pbr (roughness, metalness) | uv_test
  objects:   type box (center, dimensions) | triangle-less meshes as in the reference; `rotate_y: deg` in transform (applied
             first, like RTIOW); `medium: {density: f, albedo: [r,g,b]}` wraps the transformed object in a ConstantMedium.
Decoding of image / .hdr / .obj files stays on the host as in the reference (stb_image / Assimp there; PIL, a Radiance
reader and a small OBJ reader here, equivalent for the files the tests use)."""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple

import numpy as np
import yaml

from . import abi

F = np.float32


class SceneError(Exception):
    """Scene::loadScene's failure paths (it prints and returns -1, scene.cpp:366-373)."""


def _req(node: dict, name: str):
    if not isinstance(node, dict) or name not in node or node[name] is None:
        raise SceneError(f"Could not find required property: {name}")        # scene.cpp:22
    return node[name]


def _vec(node: dict, name: str, n: int):
    v = _req(node, name)
    if not isinstance(v, (list, tuple)):
        raise SceneError(f"Invalid value for vector {n}: {name}")
    if len(v) != n:
        raise SceneError(f"Invalid size for vector {n}: {name}")
    return tuple(float(F(x)) for x in v)


# ---- file decoders (host side, like stb_image / Assimp in the reference) ----------------------------------------------------
def read_hdr(path: str) -> Optional[np.ndarray]:
    """Radiance .hdr -> float32 [h, w, 3] exactly like stbi_loadf (texture.cpp:99-115): byte * 2^(e-136), flat or new-RLE scanlines."""
    try:
        data = open(path, "rb").read()
    except OSError:
        return None
    pos = data.find(b"\n\n")
    if pos < 0 or not data.startswith(b"#?"):
        return None
    head_end = pos + 2
    line_end = data.index(b"\n", head_end)
    tok = data[head_end:line_end].split()
    if len(tok) != 4 or tok[0] != b"-Y" or tok[2] != b"+X":
        return None
    h, w = int(tok[1]), int(tok[3])
    p = line_end + 1
    out = np.zeros((h, w, 4), np.uint8)
    raw = np.frombuffer(data, np.uint8)
    rle = 8 <= w < 32768 and raw[p] == 2 and raw[p + 1] == 2 and not (raw[p + 2] & 0x80)     # stbi__hdr_load's test of the first scanline
    if not rle:
        out[:] = raw[p:p + w * h * 4].reshape(h, w, 4)
    else:
        for y in range(h):
            if raw[p] != 2 or raw[p + 1] != 2 or (raw[p + 2] & 0x80) or ((int(raw[p + 2]) << 8) | int(raw[p + 3])) != w:
                return None                                                       # "invalid decoded scanline length"
            p += 4
            for c in range(4):
                x = 0
                while x < w:
                    cnt = int(raw[p]); p += 1
                    if cnt > 128:
                        cnt -= 128; out[y, x:x + cnt, c] = raw[p]; p += 1
                    else:
                        out[y, x:x + cnt, c] = raw[p:p + cnt]; p += cnt
                    x += cnt
    e = out[..., 3].astype(np.int32)
    f = np.ldexp(F(1.0), e - 136).astype(F)
    rgb = out[..., :3].astype(F) * f[..., None]
    rgb[e == 0] = 0
    return rgb.astype(F)


def read_image(path: str) -> np.ndarray:
    """8-bit RGB [h, w, 3] like stbi_load(path, ..., 3) (texture.cpp:30-51); an unreadable file gives the empty (cyan) texture."""
    try:
        from PIL import Image
        return np.ascontiguousarray(np.asarray(Image.open(path).convert("RGB"), dtype=np.uint8))
    except Exception:
        return np.zeros((0, 0, 3), np.uint8)


def read_obj(path: str) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """What mesh.cpp:53-120 gets from Assimp's OBJ importer under aiProcess_Triangulate | aiProcess_FlipUVs: unshared corners,
    fan triangulation, v -> 1 - v, zero normals / uvs when the file has none (same rules as oracle/obj_reader.h)."""
    V, N, T, corners = [], [], [], []
    with open(path, "r") as f:
        for line in f:
            p = line.split()
            if not p:
                continue
            if p[0] == "v":
                V.append([float(x) for x in (p[1:4] + ["0", "0", "0"])[:3]])
            elif p[0] == "vn":
                N.append([float(x) for x in (p[1:4] + ["0", "0", "0"])[:3]])
            elif p[0] == "vt":
                T.append([float(x) for x in (p[1:3] + ["0", "0"])[:2]])
            elif p[0] == "f":
                face = []
                for tok in p[1:]:
                    q = (tok.split("/") + ["", ""])[:3]
                    v, t, n = (int(x) if x else 0 for x in q)
                    if v < 0: v = len(V) + v + 1
                    if t < 0: t = len(T) + t + 1
                    if n < 0: n = len(N) + n + 1
                    face.append((v, t, n))
                for k in range(1, len(face) - 1):
                    corners += [face[0], face[k], face[k + 1]]
    Va, Na, Ta = np.asarray(V, F).reshape(-1, 3), np.asarray(N, F).reshape(-1, 3), np.asarray(T, F).reshape(-1, 2)
    c = np.asarray(corners, np.int64).reshape(-1, 3)
    any_n, any_t = bool((c[:, 2] > 0).any()), bool((c[:, 1] > 0).any())
    pos = np.zeros((len(c), 3), F); nrm = np.zeros((len(c), 3), F); uv = np.zeros((len(c), 2), F)
    ok = (c[:, 0] >= 1) & (c[:, 0] <= len(Va)); pos[ok] = Va[c[ok, 0] - 1]
    if any_n:
        ok = (c[:, 2] >= 1) & (c[:, 2] <= len(Na)); nrm[ok] = Na[c[ok, 2] - 1]
    if any_t:
        ok = (c[:, 1] >= 1) & (c[:, 1] <= len(Ta)); uv[ok] = Ta[c[ok, 1] - 1]; uv[ok, 1] = F(1.0) - uv[ok, 1]
    return pos.reshape(-1, 9), nrm.reshape(-1, 9), uv.reshape(-1, 6)


# ---- Scene::loadScene -------------------------------------------------------------------------------------------------------
class LoadedScene:
    def __init__(self):
        self.sd = abi.SceneDesc()
        self.film: Dict = {}
        self.camera: Dict = {}
        self.textures: Dict[str, int] = {}
        self.materials: Dict[str, int] = {}
        self.messages = []                     # what the reference prints to stdout while loading

    def camera_struct(self) -> "abi.Camera":
        w, h = self.film["width"], self.film["height"]
        aspect = np.float32(w) / np.float32(h)                                    # Film::getAspectRatio film.h:20-22
        c = self.camera
        return abi.make_camera(c["position"], c["look_at"], c["up"], c["vfov_deg"], aspect, c["aperture"], c["focal_distance"])


def load_scene(path: str) -> LoadedScene:
    """scene.cpp:127-374.  Relative asset paths resolve against the process cwd, as in the reference."""
    try:
        root = yaml.safe_load(open(path, "r"))
    except Exception as e:                                                       # scene.cpp:366-369
        raise SceneError(str(e))
    ls = LoadedScene(); sd = ls.sd
    film = root.get("film") if isinstance(root, dict) else None
    if not film:
        raise SceneError("Must specify film descriptor!")                         # scene.cpp:150-154
    ls.film = dict(width=int(_req(film, "width")), height=int(_req(film, "height")), samples=int(_req(film, "samples")), output=str(_req(film, "output")))
    cam = root.get("camera")
    if not cam:
        raise SceneError("Must specify camera descriptor!")                       # scene.cpp:168-172
    ls.camera = dict(position=_vec(cam, "position", 3), look_at=_vec(cam, "look_at", 3), up=_vec(cam, "up", 3), vfov_deg=float(_req(cam, "fov")),
                     aperture=float(_req(cam, "aperture")), focal_distance=float(_req(cam, "focal_distance")))

    def image_texture(name):                                                      # textures[name] = make_shared<ImageTexture>(name)
        ls.textures[name] = sd.image(read_image(name)); return ls.textures[name]

    def env_texture(name):                                                        # make_shared<EnvironmentMap>(path), texture.cpp:99-115
        data = read_hdr(name)
        ls.textures[name] = sd.env(data if data is not None else np.zeros((0, 0, 3), F), name); return ls.textures[name]

    for t in root.get("textures") or []:                                          # scene.cpp:174-213
        name = str(_req(t, "name"))
        if name in ls.textures:
            raise SceneError("Texture name already exists!")
        kind = str(_req(t, "type"))
        if kind == "solid":
            ls.textures[name] = sd.solid(_vec(t, "colour", 3))
        if kind == "image":
            ls.textures[name] = sd.image(read_image(str(_req(t, "path"))))
        if kind == "checkered":
            ls.textures[name] = sd.checker(sd.solid(_vec(t, "even", 3)), sd.solid(_vec(t, "odd", 3)))
        if kind == "environment":
            p = str(_req(t, "path")); data = read_hdr(p)
            ls.textures[name] = sd.env(data if data is not None else np.zeros((0, 0, 3), F), p)

    bg = _req(cam, "background")                                                  # scene.cpp:215-237
    if isinstance(bg, (list, tuple)):
        sd.background = sd.solid(_vec(cam, "background", 3))
    else:
        sd.background = ls.textures[str(bg)] if str(bg) in ls.textures else env_texture(str(bg))

    def mat_vec3(node, name):                                                     # getProperty<MatVec3> scene.cpp:77-99
        v = _req(node, name)
        if isinstance(v, (list, tuple)):
            return _vec(node, name, 3), -1
        return (0.0, 0.0, 0.0), (ls.textures[str(v)] if str(v) in ls.textures else image_texture(str(v)))

    def mat_scalar(node, name):                                                   # getProperty<MatScalar> scene.cpp:101-125
        v = _req(node, name)
        if isinstance(v, (int, float)) and not isinstance(v, bool):
            return float(v), -1
        return 0.0, (ls.textures[str(v)] if str(v) in ls.textures else image_texture(str(v)))

    mats = root.get("materials")
    if mats:                                                                      # scene.cpp:240-266
        for m in mats:
            name = str(_req(m, "name")); albedo, atex = mat_vec3(m, "albedo"); kind = str(_req(m, "type"))
            if kind == "diffuse_light":
                s, stex = mat_scalar(m, "strength"); ls.materials[name] = sd.material(abi.MAT_DIFFUSE_LIGHT, albedo, atex, s, stex)
            elif kind == "lambertian":
                ls.materials[name] = sd.material(abi.MAT_LAMBERTIAN, albedo, atex)
            elif kind == "metal":
                r, rtex = mat_scalar(m, "roughness"); ls.materials[name] = sd.material(abi.MAT_METAL, albedo, atex, r, rtex)
            elif kind == "dielectric":                                            # extension: Dielectric(ior, roughness) material.h:202
                i, itex = mat_scalar(m, "ior"); r, rtex = mat_scalar(m, "roughness") if "roughness" in m else (0.0, -1)
                ls.materials[name] = sd.material(abi.MAT_DIELECTRIC, albedo, atex, i, itex, r, rtex)
            elif kind == "isotropic":                                             # extension: Isotropic(albedo) material.h:73-89
                ls.materials[name] = sd.material(abi.MAT_ISOTROPIC, albedo, atex)
            elif kind == "pbr":                                                   # extension: PBR(albedo, metalness, roughness) material.cpp:4-28
                r, rtex = mat_scalar(m, "roughness"); q, qtex = mat_scalar(m, "metalness")
                ls.materials[name] = sd.material(abi.MAT_PBR, albedo, atex, r, rtex, q, qtex)
            elif kind == "uv_test":
                ls.materials[name] = sd.material(abi.MAT_UVTEST)
    else:
        ls.messages.append("Couldn't find any material descriptors!")

    objs = root.get("objects")
    if isinstance(objs, list):                                                    # scene.cpp:271-359
        for o in objs:
            key = str(_req(o, "material"))
            if key not in ls.materials:
                ls.messages.append(f"Material {key} does not exist!"); continue  # scene.cpp:286-289
            m = ls.materials[key]; kind = str(_req(o, "type")); shape = None
            if kind == "mesh":
                pos, nrm, uv = read_obj(str(_req(o, "path"))); shape = sd.mesh(pos, nrm, uv, m)
            if kind == "sphere":
                shape = sd.sphere(_vec(o, "center", 3), float(F(_req(o, "radius"))), m)
            if kind == "yz_rect":
                a, b = _vec(o, "y", 2), _vec(o, "z", 2); shape = sd.rect(0, a[0], a[1], b[0], b[1], float(F(_req(o, "k"))), m)
            if kind == "xz_rect":
                a, b = _vec(o, "x", 2), _vec(o, "z", 2); shape = sd.rect(1, a[0], a[1], b[0], b[1], float(F(_req(o, "k"))), m)
            if kind == "xy_rect":
                a, b = _vec(o, "x", 2), _vec(o, "y", 2); shape = sd.rect(2, a[0], a[1], b[0], b[1], float(F(_req(o, "k"))), m)
            if kind == "box":                                                     # extension: Box(center, dimensions) box.h:41-44
                shape = sd.box(_vec(o, "center", 3), _vec(o, "dimensions", 3), m)
            if shape is None:
                raise SceneError(f"unknown object type {kind} (the reference adds a null object and crashes on the first hit, scene.cpp:279,356)")
            ops = []
            tr = o.get("transform")
            if tr:                                                                # scene.cpp:335-353, innermost first
                if "rotate_y" in tr: ops.append((abi.XF_ROTATE_Y_DEG, (float(tr["rotate_y"]),)))      # extension: RotateY rotateY.h:8
                if "rotate" in tr: ops.append((abi.XF_ROTATE_EULER_DEG, _vec(tr, "rotate", 3)))
                if "scale" in tr: ops.append((abi.XF_SCALE, _vec(tr, "scale", 3)))
                if "translate" in tr: ops.append((abi.XF_TRANSLATE, _vec(tr, "translate", 3)))
            med = o.get("medium")
            if med:                                                               # extension: ConstantMedium(boundary, density, albedo)
                sd.add_medium(shape, ops, float(_req(med, "density")), _vec(med, "albedo", 3))
            else:
                sd.add(shape, ops)
    return ls
