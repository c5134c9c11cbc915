"""Scenes of the golden fixtures (tests/golden/*.npz).  Shared by make_golden.py (which runs the UNMODIFIED reference
through oracle/_ref and records its answers) and tests/test_golden.py (which replays them against the oracle on the
CPU and against the CUDA library on the GPU).  Everything a scene needs that is not a literal below (mesh arrays, ray
batches) is stored inside the fixture, so the replay does not depend on numpy's generators or libm."""
import numpy as np

from hobbyraytracer_b200 import abi

F = np.float32


def scene_prims(fx):
    """Sphere / rect / box arithmetic and list-order ties (sphere.cpp:20-49, aarect.h, box.h:27-39, hittableList.cpp:4-21)."""
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    sd.add(sd.sphere((0, 0, -1), 0.5, m))
    sd.add(sd.sphere((0, -100.5, -1), 100, m))
    sd.add(sd.sphere((1.5, 0.25, -2), 0.75, m))
    sd.add(sd.sphere((1.5, 0.25, -2), 0.75, m))            # identical twin: the later one wins (closed interval)
    sd.add(sd.rect(0, -1, 1, -3, 1, -2.5, m))              # YZRect x = -2.5
    sd.add(sd.rect(1, -3, 3, -4, 1, 2.0, m))               # XZRect y = 2
    sd.add(sd.rect(2, -3, 3, -1, 2, -4.0, m))              # XYRect z = -4
    sd.add(sd.box((-1.5, 0.5, -2.5), (0.8, 1.0, 0.6), m))
    return sd


def scene_mesh(fx):
    """ITriangle + BVHNode (triangle.cpp:57-131, bvh.cpp:69-78) with a floor rect behind it."""
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_METAL, (0.7, 0.7, 0.7), p0=0.2)
    sd.add(sd.mesh(fx["mesh_pos"], fx["mesh_nrm"], fx["mesh_uv"], m))
    sd.add(sd.rect(1, -4, 4, -4, 4, -0.25, m))
    return sd


def scene_instances(fx):
    """Translate / Scale / RotateQuat / RotateY chains in the loader's nesting order (scene.cpp:335-353)."""
    sd = abi.SceneDesc()
    m = sd.material(abi.MAT_LAMBERTIAN, (0.5, 0.5, 0.5))
    mesh = sd.mesh(fx["mesh_pos"], fx["mesh_nrm"], fx["mesh_uv"], m)
    sd.add(mesh, [(abi.XF_ROTATE_EULER_DEG, (20, 35, -10)), (abi.XF_SCALE, (1.5, 0.75, 1.25)), (abi.XF_TRANSLATE, (0.5, 0.25, -0.5))])
    sd.add(sd.sphere((0, 0, 0), 1.0, m), [(abi.XF_SCALE, (3, 3, 3)), (abi.XF_TRANSLATE, (6, 1, 0))])
    sd.add(sd.box((0, 0, 0), (1, 2, 1), m), [(abi.XF_ROTATE_Y_DEG, (15,)), (abi.XF_TRANSLATE, (-4, 1, 1))])
    sd.add(sd.sphere((0, 0, 0), 0.5, m), [(abi.XF_ROTATE_EULER_DEG, (0, 45, 0)), (abi.XF_TRANSLATE, (0, 3, 0))])
    sd.add(sd.rect(1, -10, 10, -10, 10, -1.0, m))
    return sd


def scene_render(fx, env_path=None):
    """Small env-lit render: metal mesh + lambertian floor + light rect under the procedural sky (rayColour main.cpp:38-79)."""
    sd = abi.SceneDesc()
    sd.background = sd.env(fx["env"], env_path)      # the reference can only read environment maps from .hdr files
    shiny = sd.material(abi.MAT_METAL, (0.73, 0.73, 0.73), p0=0.2)
    floor = sd.material(abi.MAT_LAMBERTIAN, (0.4, 0.5, 0.3))
    light = sd.material(abi.MAT_DIFFUSE_LIGHT, (1.0, 0.9, 0.8), p0=3.0)
    sd.add(sd.mesh(fx["mesh_pos"], fx["mesh_nrm"], fx["mesh_uv"], shiny), [(abi.XF_TRANSLATE, (0, 0, 0))])
    sd.add(sd.rect(1, -4, 4, -4, 4, 0.0, floor))
    sd.add(sd.rect(2, -1, 1, 2.5, 3.5, -2.5, light))
    sd.add(sd.sphere((1.6, 0.5, 0.6), 0.5, sd.material(abi.MAT_LAMBERTIAN, (0.7, 0.2, 0.2))))
    return sd


RENDER_CAMERA = dict(position=(0, 1.5, 4), look_at=(0, 1, 0), up=(0, 1, 0), vfov_deg=50, aperture=0.0, focal_distance=4.0)
RENDER_W, RENDER_H, RENDER_SPP = 48, 32, 512

TRACE_SCENES = {"prims": scene_prims, "mesh": scene_mesh, "instances": scene_instances}
