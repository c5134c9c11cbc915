"""No-GPU checks of the drop-in boundary: the CUDA library loads, exports every symbol include/hrt.h declares, and its
argument / state errors follow the header's conventions (no compute call is made here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from hobbyraytracer_b200 import abi
from conftest import ROOT, PRODUCT_SO


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "hrt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hrt_[a-z0-9_]+)\s*\(", text)))


def test_exports_every_declared_symbol(hrt):
    syms = declared_symbols()
    assert len(syms) >= 28
    missing = [s for s in syms if not hasattr(hrt.dll, s)]
    assert not missing, f"libhrt_b200.so does not export {missing}"
    assert hrt.fn("abi_version")() == 7


def test_builder_errors_and_state(hrt):
    h = hrt.fn("scene_create")()
    assert h
    rgb = (C.c_float * 3)(0.5, 0.5, 0.5)
    assert hrt.fn("scene_add_texture_solid")(h, rgb) == 0
    assert hrt.fn("scene_add_texture_checker")(h, 0, 7) == -1            # unknown texture -> HRT_ERR_ARG
    assert b"unknown texture" in hrt.fn("last_error")()
    m = abi.Material(); m.kind = 99
    assert hrt.fn("scene_add_material")(h, C.byref(m)) == -1
    m.kind = abi.MAT_LAMBERTIAN; m.albedo_tex = -1; m.p0_tex = -1; m.p1_tex = -1
    assert hrt.fn("scene_add_material")(h, C.byref(m)) == 0
    c = (C.c_float * 3)(0, 0, 0)
    assert hrt.fn("scene_add_sphere")(h, c, 1.0, 5) == -1                  # unknown material
    sh = hrt.fn("scene_add_sphere")(h, c, 1.0, 0)
    assert sh == 0 and hrt.fn("scene_prim_base")(h, sh) == 0 and hrt.fn("scene_prim_count")(h) == 1
    assert hrt.fn("scene_add_rect")(h, 3, 0, 1, 0, 1, 0, 0) == -1           # bad axis
    assert hrt.fn("scene_add_object")(h, 42, None, 0) == -1
    ops = (abi.Xform * 9)()
    assert hrt.fn("scene_add_object")(h, sh, ops, 9) == -1                  # more than 8 wrappers
    assert hrt.fn("scene_add_object")(h, sh, None, 0) == 0
    # compute entry points before commit: HRT_ERR_STATE, never a crash
    rays = np.zeros((1, 8), np.float32); prim = np.zeros(1, np.uint32); t = np.zeros(1, np.float32)
    assert hrt.fn("trace_batch")(h, rays.ctypes.data, 1, prim.ctypes.data, t.ctypes.data, None, None, None, 0) == -2
    cam = abi.make_camera((0, 0, 3), (0, 0, 0), (0, 1, 0), 40, 1.0)
    p = abi.RenderParams(8, 8, 1, 0, 0, 50, 0.001, 1, 0, 0)
    assert hrt.fn("render")(h, C.byref(cam), C.byref(p), None, None) == -2
    assert b"not committed" in hrt.fn("last_error")()
    hrt.fn("scene_destroy")(h)
    hrt.fn("scene_destroy")(None)


def test_no_cpu_fallback_in_product_sources():
    """The product must not reach into oracle/ or tests/hostsim (the checkers)."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "hobbyraytracer_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in src and "hostsim.cpp" not in src and "libhrt_ref" not in src, f
