"""Drop-in check at the reference's own front door: ONE .yaml scene file (the schema of scene.cpp:127-374) is loaded by
the UNMODIFIED reference loader (Scene::loadScene through oracle/_ref) and by hobbyraytracer_b200.yaml_scene into the oracle
(CPU) and the CUDA library (-m gpu); world->hit on the same rays must agree bit for bit (hit flag, t, normal)."""
import ctypes as C
import os

import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes, yaml_scene

SCENE = """# teapot-in-a-room in the reference's schema: every object / material / texture / transform kind the shipped loader knows
film: {width: 64, height: 48, samples: 4, output: out.png}
camera:
    position: [0, 1.5, 5]
    look_at: [0, 1, 0]
    up: [0, 1, 0]
    fov: 50
    aperture: 0.001
    focal_distance: 5
    background: sky
textures:
  - {name: sky, type: environment, path: sky.hdr}
  - {name: tiles, type: checkered, even: [0.8, 0.8, 0.8], odd: [0.1, 0.1, 0.3]}
  - {name: red, type: solid, colour: [0.65, 0.05, 0.05]}
  - {name: photo, type: image, path: photo.png}
materials:
  - {name: shiny, type: metal, albedo: [0.73, 0.73, 0.73], roughness: 0.2}
  - {name: floor, type: lambertian, albedo: tiles}
  - {name: wall, type: lambertian, albedo: red}
  - {name: poster, type: lambertian, albedo: photo}
  - {name: lamp, type: diffuse_light, albedo: [1, 0.9, 0.8], strength: 4}
  - {name: rough, type: metal, albedo: [0.9, 0.6, 0.2], roughness: photo}
objects:
  - {type: mesh, path: pot.obj, material: shiny, transform: {rotate: [0, 35, 10], scale: [1.2, 1.2, 1.2], translate: [0.2, 0, -0.3]}}
  - {type: mesh, path: pot.obj, material: rough, transform: {translate: [-2.5, 0, -1]}}
  - {type: xz_rect, x: [-6, 6], z: [-6, 6], k: 0, material: floor}
  - {type: yz_rect, y: [0, 4], z: [-6, 6], k: -4, material: wall}
  - {type: xy_rect, x: [-4, 4], y: [0, 4], k: -5, material: poster}
  - {type: xz_rect, x: [-1, 1], z: [-1, 1], k: 3.9, material: lamp}
  - {type: sphere, center: [2, 0.7, 0.5], radius: 0.7, material: rough}
  - {type: sphere, center: [0, 0, 0], radius: 0.5, material: shiny, transform: {scale: [1, 2, 1], translate: [3, 1, -2]}}
  - {type: sphere, center: [9, 9, 9], radius: 1, material: nosuchmaterial}
"""


@pytest.fixture()
def scene_dir(tmp_path):
    """The scene file and its assets; relative paths resolve against the cwd in the reference (and here)."""
    pos, nrm, uv = scenes.teapot_like_mesh(10)
    scenes.write_obj(str(tmp_path / "pot.obj"), pos, nrm, uv)
    scenes.write_hdr(str(tmp_path / "sky.hdr"), scenes.hall_env(64, 32, 7))
    from PIL import Image
    rng = np.random.default_rng(1)
    Image.fromarray(rng.integers(0, 256, (24, 32, 3), dtype=np.uint8)).save(str(tmp_path / "photo.png"))
    (tmp_path / "scene.yaml").write_text(SCENE)
    old = os.getcwd(); os.chdir(tmp_path)
    try:
        yield tmp_path
    finally:
        os.chdir(old)


def ref_answers(ref, rays):
    ref._sig("scene_load_yaml", C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)])
    ref._sig("yaml_trace_batch", C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p])
    h = ref.fn("scene_create")()
    w, hh, spp = C.c_int(), C.c_int(), C.c_int()
    ref.check(ref.fn("scene_load_yaml")(h, b"scene.yaml", 3, C.byref(w), C.byref(hh), C.byref(spp)), "scene_load_yaml")
    n = len(rays)
    t = np.empty(n, np.float32); nrm = np.empty((n, 3), np.float32); hit = np.empty(n, np.uint8)
    ref.check(ref.fn("yaml_trace_batch")(h, rays.ctypes.data, n, t.ctypes.data, nrm.ctypes.data, hit.ctypes.data), "yaml_trace_batch")
    ref.fn("scene_destroy")(h)
    return (w.value, hh.value, spp.value), hit, t, nrm


def bits(a):
    a = np.ascontiguousarray(a, np.float32); b = a.view(np.uint32).copy(); b[np.isnan(a)] = 0x7FC00000; return b


def rays_for(ls):
    cam = dict(ls.camera); cam["focal_distance"] = ls.camera["focal_distance"]
    a = scenes.camera_rays(cam, ls.film["width"], ls.film["height"], 6000, seed=2)
    b = scenes.random_rays(6000, (0, 1, 0), 4.0, seed=5)
    return np.ascontiguousarray(np.concatenate([a, b]), np.float32)


def compare(lib, mode, ls, rays, hit, t, nrm):
    out = ls.sd.build(lib, mode, 3).trace(rays)
    got = out["prim"] != abi.MISS
    assert np.array_equal(got, hit.astype(bool)), f"{(got != hit.astype(bool)).sum()} rays disagree on hit / miss"
    assert np.array_equal(bits(out["t"][got]), bits(t[got]))
    assert np.array_equal(bits(out["normal"][got]), bits(nrm[got]))
    assert got.mean() > 0.5


def test_same_yaml_into_reference_loader_and_oracle(ref, orc, scene_dir):
    ls = yaml_scene.load_scene("scene.yaml")
    assert ls.messages == ["Material nosuchmaterial does not exist!"] and len(ls.sd.objects) == 8      # scene.cpp:286-289: skipped
    rays = rays_for(ls)
    film, hit, t, nrm = ref_answers(ref, rays)
    assert film == (ls.film["width"], ls.film["height"], ls.film["samples"])
    compare(orc, abi.BVH_REFERENCE, ls, rays, hit, t, nrm)


def test_loader_failures_match_the_reference(ref, scene_dir):
    """loadScene prints and returns -1 (scene.cpp:150-172, 366-373); the mirror raises SceneError for the same files."""
    ref._sig("scene_load_yaml", C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)])
    bad = {"nofilm.yaml": SCENE.replace("film: {width: 64, height: 48, samples: 4, output: out.png}\n", ""),
           "nofov.yaml": SCENE.replace("    fov: 50\n", ""),
           "vec2.yaml": SCENE.replace("position: [0, 1.5, 5]", "position: [0, 1.5]"),
           "dup.yaml": SCENE.replace("{name: red, type: solid", "{name: sky, type: solid")}
    for name, text in bad.items():
        open(name, "w").write(text)
        h = ref.fn("scene_create")()
        rc = ref.fn("scene_load_yaml")(h, name.encode(), 3, None, None, None)
        ref.fn("scene_destroy")(h)
        assert rc < 0, name
        with pytest.raises(yaml_scene.SceneError):
            yaml_scene.load_scene(name)


def test_extension_keys_reach_the_unreachable_classes(orc, scene_dir):
    """dielectric / isotropic / pbr materials, box, rotate_y, constant medium through the same schema."""
    text = SCENE.replace("materials:\n", "materials:\n  - {name: glass, type: dielectric, albedo: [1, 1, 1], ior: 1.5, roughness: 0.1}\n"
                                         "  - {name: fog, type: isotropic, albedo: [1, 1, 1]}\n  - {name: mix, type: pbr, albedo: tiles, roughness: 0.3, metalness: photo}\n")
    text += ("  - {type: box, center: [1, 1, 1], dimensions: [1, 2, 1], material: glass, transform: {rotate_y: 15, translate: [0, 0, 1]}}\n"
             "  - {type: box, center: [-1, 1, 2], dimensions: [1, 1, 1], material: fog, transform: {rotate_y: -18}, medium: {density: 0.5, albedo: [1, 1, 1]}}\n"
             "  - {type: sphere, center: [0, 3, 0], radius: 0.4, material: mix}\n")
    open("ext.yaml", "w").write(text)
    ls = yaml_scene.load_scene("ext.yaml")
    assert len(ls.sd.objects) == 11 and ls.sd.objects[-2][0] == "medium"
    sc = ls.sd.build(orc, abi.BVH_REFERENCE, 3)
    s, rays = sc.render_f32(ls.camera_struct(), 32, 24, 4, seed=1, threads=2)
    assert np.isfinite(s).all() and s.mean() > 0 and rays > 32 * 24 * 4


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [abi.BVH_REFERENCE, abi.BVH_SAH])
def test_same_yaml_into_reference_loader_and_gpu(ref, hrt, scene_dir, mode):
    ls = yaml_scene.load_scene("scene.yaml")
    rays = rays_for(ls)
    _, hit, t, nrm = ref_answers(ref, rays)
    compare(hrt, mode, ls, rays, hit, t, nrm)
    s, rgb = ls.sd.build(hrt, mode, 3).render(ls.camera_struct(), ls.film["width"], ls.film["height"], 16, seed=1)
    assert np.isfinite(s).all() and s.mean() > 0 and rgb.shape == (48, 64, 3)
