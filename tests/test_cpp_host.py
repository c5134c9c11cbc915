"""host/_build/hobbyraytracer_b200: the reference's own host — Scene::loadScene + yaml-cpp, Mesh + Assimp, stb, Film::outputFilm,
compiled UNMODIFIED from /root/reference (host/Makefile) — in front of libhrt_b200.so (host/hrt_host.cpp replaces main.cpp).

CPU: the binary loads a scene file with every kind the shipped schema knows (and the SURVEY §8(f)1 extension keys), replays it into
the C ABI and stops at "no CUDA device" with the reference's failure code.  GPU: the PNG it writes is compared with hrt_render through
ctypes on the same scene (bit for bit where the primitive order is the file order; within the noise floor where meshes are involved)
and with the CPU oracle."""
import os
import subprocess

import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes, yaml_scene
from conftest import ROOT
from test_yaml_dropin import SCENE, scene_dir  # noqa: F401

HOST = os.path.join(ROOT, "host", "_build", "hobbyraytracer_b200")

PLAIN = """film: {width: 96, height: 64, samples: 32, output: plain.png}
camera: {position: [0, 2, 7], look_at: [0, 1, 0], up: [0, 1, 0], fov: 45, aperture: 0.0, focal_distance: 7, background: [0.6, 0.7, 0.9]}
textures:
  - {name: tiles, type: checkered, even: [0.8, 0.8, 0.8], odd: [0.1, 0.1, 0.3]}
materials:
  - {name: floor, type: lambertian, albedo: tiles}
  - {name: shiny, type: metal, albedo: [0.8, 0.7, 0.6], roughness: 0.1}
  - {name: lamp, type: diffuse_light, albedo: [1, 0.9, 0.8], strength: 5}
  - {name: matte, type: lambertian, albedo: [0.7, 0.2, 0.2]}
objects:
  - {type: xz_rect, x: [-6, 6], z: [-6, 6], k: 0, material: floor}
  - {type: sphere, center: [-1.2, 1, 0], radius: 1, material: shiny}
  - {type: sphere, center: [1.2, 0.8, 0.5], radius: 0.8, material: matte, transform: {rotate: [0, 30, 0], scale: [1, 1.3, 1], translate: [0.2, 0, -0.4]}}
  - {type: xy_rect, x: [-3, 3], y: [0, 3], k: -3, material: matte}
  - {type: xz_rect, x: [-1, 1], z: [-1, 1], k: 4, material: lamp}
"""

EXT = SCENE.replace("materials:\n", "materials:\n  - {name: glass, type: dielectric, albedo: [1, 1, 1], ior: 1.5, roughness: 0.1}\n"
                    "  - {name: fog, type: isotropic, albedo: [1, 1, 1]}\n  - {name: mix, type: pbr, albedo: tiles, roughness: 0.3, metalness: photo}\n") + (
    "  - {type: box, center: [1, 1, 1], dimensions: [1, 2, 1], material: glass, transform: {rotate_y: 15, translate: [0, 0, 1]}}\n"
    "  - {type: box, center: [-1, 1, 2], dimensions: [1, 1, 1], material: fog, transform: {rotate_y: -18}, medium: {density: 0.5, albedo: [1, 1, 1]}}\n"
    "  - {type: sphere, center: [0, 3, 0], radius: 0.4, material: mix}\n")


@pytest.fixture(scope="session")
def host_bin():
    if os.path.isdir("/root/reference/hobbyraytracer"):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "hobbyraytracer_b200", "csrc")], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "host")], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if not os.path.exists(HOST):
        pytest.skip("host/_build/hobbyraytracer_b200 not built and /root/reference absent")
    return HOST


def run(host_bin, *args):
    return subprocess.run([host_bin, *args], capture_output=True, text=True, timeout=600)


def png(path):
    from PIL import Image
    return np.asarray(Image.open(path).convert("RGB"))


def test_host_loads_and_flattens_without_a_gpu(host_bin, scene_dir):
    """No GPU in this container: everything up to hrt_scene_commit must work (reference loader, Assimp OBJ import, stb decoding, the
    walk over the object graph, every hrt_scene_add_* call) and the program must end like the reference does on a failure (-1)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("this is the no-GPU leg")
    (scene_dir / "ext.yaml").write_text(EXT)
    for name in ("scene.yaml", "ext.yaml"):
        r = run(host_bin, name, "--quiet")
        assert "Loaded mesh: pot.obj" in r.stdout and "Indexed file: pot.obj" in r.stdout          # mesh.cpp:40,118: the real Assimp path ran
        assert "Material nosuchmaterial does not exist!" in r.stdout                                # scene.cpp:288
        assert "Loaded scene: " + name in r.stdout and "no CUDA device" in r.stdout, r.stdout
        assert r.returncode == 255                                                                   # main.cpp:156: return -1
    (scene_dir / "bad.yaml").write_text(SCENE.replace("    fov: 50\n", ""))
    r = run(host_bin, "bad.yaml")
    assert r.returncode == 255 and "fov" in r.stdout and "Loaded scene" not in r.stdout
    (scene_dir / "badext.yaml").write_text(EXT.replace("dimensions: [1, 2, 1]", "dimensions: [1, 2]"))
    r = run(host_bin, "badext.yaml")
    assert r.returncode == 255 and "dimensions" in r.stdout


@pytest.mark.gpu
def test_host_png_equals_hrt_render(host_bin, hrt, tmp_path):
    """Spheres, rects, instance wrappers, textures: the world list reaches the library in file order from both hosts, so the
    8-bit image the C++ host writes through Film::outputFilm is the one hrt_render returns through ctypes, byte for byte."""
    os.chdir(tmp_path)
    (tmp_path / "plain.yaml").write_text(PLAIN)
    r = run(host_bin, "plain.yaml", "--quiet", "--seed", "5", "--sum", "sum.f32")
    assert r.returncode == 1, r.stdout + r.stderr                                                   # stb's success code, main.cpp:194
    ls = yaml_scene.load_scene("plain.yaml")
    s, rgb = ls.sd.build(hrt, abi.BVH_SAH, 5).render(ls.camera_struct(), 96, 64, 32, seed=5)
    assert np.array_equal(png("plain.png"), rgb)
    assert np.array_equal(np.fromfile("sum.f32", np.float32).reshape(64, 96, 3).view(np.uint32), s.view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["scene.yaml", "ext.yaml"])
def test_host_mesh_and_extension_scenes(host_bin, hrt, orc, scene_dir, name):
    """The teapot-in-a-room file (meshes through Assimp) and the same file with the extension keys: the C++ host's float sums against
    hrt_render on the Python loader's scene (triangles in another order: same image up to ties) and against the CPU oracle (noise floor)."""
    (scene_dir / "ext.yaml").write_text(EXT)
    w, h, spp = 64, 48, 256
    r = run(host_bin, name, "--quiet", "--spp", str(spp), "--seed", "3", "--sum", "sum.f32")
    assert r.returncode == 1, r.stdout + r.stderr
    got = np.fromfile("sum.f32", np.float32).reshape(h, w, 3) / spp
    assert png("out.png").shape == (h, w, 3)
    ls = yaml_scene.load_scene(name)
    cam = ls.camera_struct()
    s, _ = ls.sd.build(hrt, abi.BVH_SAH, 3).render(cam, w, h, spp, seed=3)
    py = s / spp
    o = ls.sd.build(orc, abi.BVH_REFERENCE, 3)
    a, _ = o.render_f32(cam, w, h, spp, seed=11, threads=0); b, _ = o.render_f32(cam, w, h, spp, seed=12, threads=0)
    a /= spp; b /= spp
    lim = 60.0                                                           # the lamp / HDR window: compare below the fireflies
    rm = lambda x, y: float(np.sqrt(np.mean((np.minimum(x, lim) - np.minimum(y, lim)) ** 2)))
    floor = rm(a, b)
    assert rm(got, a) <= 1.25 * floor and rm(got, py) <= 1.25 * floor, (rm(got, a), rm(got, py), floor)
    assert abs(np.minimum(got, lim).mean() - np.minimum(a, lim).mean()) <= 3 * abs(np.minimum(a, lim).mean() - np.minimum(b, lim).mean()) + 0.02 * np.minimum(a, lim).mean()


@pytest.mark.gpu
def test_host_multi_gpu_is_bit_identical(host_bin, hrt, tmp_path):
    if hrt.fn("device_count")() < 2:
        pytest.skip("needs 2 GPUs")
    os.chdir(tmp_path)
    (tmp_path / "plain.yaml").write_text(PLAIN)
    assert run(host_bin, "plain.yaml", "--quiet", "--seed", "5").returncode == 1
    one = png("plain.png").copy()
    assert run(host_bin, "plain.yaml", "--quiet", "--seed", "5", "--gpus", "2").returncode == 1
    assert np.array_equal(png("plain.png"), one)
