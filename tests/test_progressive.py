"""hrt_render_progressive — render() in slices with checkpoints and a progress callback (the
reference prints progress from a reporter thread, main.cpp:95-109, and has no way to resume a render).

The film is integer and Philox is keyed by (pixel, sample index), so slicing, stopping and resuming must not change a
single bit of the float sums or of the 8-bit image.
"""
import os

import numpy as np
import pytest

from hobbyraytracer_b200 import abi, scenes

W, H, SPP = 96, 64, 12


@pytest.fixture(scope="module")
def scene(hrt, cache_dir):
    sd, cam, _ = scenes.teapot_scene(cache_dir, n=16, env_w=64, env_h=32)
    return sd.build(hrt, abi.BVH_SAH, 3), scenes.camera_for(cam, W, H)


def test_checkpoint_file_errors_without_a_device(hrt, tmp_path):
    """Pure host code: runs in the CPU suite too."""
    w = abi.C.c_int(0)
    assert hrt.fn("checkpoint_samples")(str(tmp_path / "missing.ckpt").encode(), abi.C.byref(w), None, None) == -4
    junk = tmp_path / "junk.ckpt"; junk.write_bytes(b"not a film" * 20)
    assert hrt.fn("checkpoint_samples")(str(junk).encode(), abi.C.byref(w), None, None) == -4
    assert b"not a checkpoint" in hrt.fn("last_error")()


@pytest.mark.gpu
@pytest.mark.parametrize("slice_spp", [1, 5, 12, 100])
def test_slices_do_not_change_the_image(scene, slice_spp):
    sc, cam = scene
    want_sum, want_rgb = sc.render(cam, W, H, SPP, seed=9)
    seen = []
    rc, s, b = sc.render_progressive(cam, W, H, SPP, seed=9, slice_spp=slice_spp, on_slice=lambda done, total: seen.append((done, total)))
    assert rc == 0 and np.array_equal(s.view(np.uint32), want_sum.view(np.uint32)) and np.array_equal(b, want_rgb)
    assert seen[-1] == (SPP, SPP) and len(seen) == -(-SPP // min(slice_spp, SPP))


@pytest.mark.gpu
def test_stop_and_resume_from_the_checkpoint(scene, hrt, tmp_path):
    sc, cam = scene
    want_sum, want_rgb = sc.render(cam, W, H, SPP, seed=4)
    ck = str(tmp_path / "film.ckpt")
    rc, _, _ = sc.render_progressive(cam, W, H, SPP, seed=4, checkpoint=ck, slice_spp=5, on_slice=lambda done, total: done >= 10)
    assert rc == 1                                                           # HRT_STOPPED after two slices
    w, h, spp = abi.C.c_int(0), abi.C.c_int(0), abi.C.c_int(0)
    assert hrt.fn("checkpoint_samples")(ck.encode(), abi.C.byref(w), abi.C.byref(h), abi.C.byref(spp)) == 10
    assert (w.value, h.value, spp.value) == (W, H, SPP) and not os.path.exists(ck + ".tmp")
    seen = []
    rc, s, b = sc.render_progressive(cam, W, H, SPP, seed=4, checkpoint=ck, slice_spp=5, on_slice=lambda done, total: seen.append(done))
    assert rc == 0 and seen == [12]                                          # only the missing two samples were rendered
    assert np.array_equal(s.view(np.uint32), want_sum.view(np.uint32)) and np.array_equal(b, want_rgb)
    # a finished checkpoint resolves again without rendering anything
    rays_before = sc.stats().rays
    rc, s, b = sc.render_progressive(cam, W, H, SPP, seed=4, checkpoint=ck, slice_spp=5)
    assert rc == 0 and sc.stats().rays == rays_before and np.array_equal(b, want_rgb)


@pytest.mark.gpu
def test_a_checkpoint_of_another_render_is_refused(scene, tmp_path):
    sc, cam = scene
    ck = str(tmp_path / "film.ckpt")
    sc.render_progressive(cam, W, H, SPP, seed=4, checkpoint=ck, slice_spp=6)
    for kwargs in (dict(seed=5), dict(seed=4, max_depth=7), dict(seed=4, rng_mode=1)):
        with pytest.raises(abi.HrtError, match="different render"):
            sc.render_progressive(cam, W, H, SPP, checkpoint=ck, slice_spp=6, **kwargs)
    with pytest.raises(abi.HrtError, match="different render"):
        sc.render_progressive(cam, W, H, SPP + 1, seed=4, checkpoint=ck)
    with open(ck, "r+b") as f:
        f.truncate(os.path.getsize(ck) // 2)
    with pytest.raises(abi.HrtError, match="truncated"):
        sc.render_progressive(cam, W, H, SPP, seed=4, checkpoint=ck)


@pytest.mark.gpu
def test_a_checkpoint_of_an_edited_scene_is_refused(hrt, tmp_path):
    """Same primitive count, same camera, same parameters - but one albedo differs: the checkpoint carries a hash of the scene's content
    (objects, wrappers, materials, textures, geometry), so the edited scene does not resume into the other one's film."""
    def build(albedo):
        sd = abi.SceneDesc()
        sd.add(sd.sphere((0, 0, -1), 0.5, sd.material(abi.MAT_LAMBERTIAN, albedo)))
        sd.add(sd.sphere((0, -100.5, -1), 100, sd.material(abi.MAT_LAMBERTIAN, (0.8, 0.8, 0.0))))
        sd.background = sd.solid((0.6, 0.7, 0.9))
        return sd.build(hrt, abi.BVH_SAH, 3)
    cam = scenes.camera_for(dict(position=(0, 0, 0), look_at=(0, 0, -1), up=(0, 1, 0), vfov_deg=60, focal_distance=1.0), 64, 48)
    ck = str(tmp_path / "film.ckpt")
    a = build((0.7, 0.3, 0.3))
    a.render_progressive(cam, 64, 48, 8, seed=2, checkpoint=ck, slice_spp=4)
    with pytest.raises(abi.HrtError, match="different render"):
        build((0.3, 0.7, 0.3)).render_progressive(cam, 64, 48, 8, seed=2, checkpoint=ck, slice_spp=4)
    _, s2, _ = build((0.7, 0.3, 0.3)).render_progressive(cam, 64, 48, 8, seed=2, checkpoint=ck, slice_spp=4)   # the same scene built again resumes
    s1, _ = a.render(cam, 64, 48, 8, seed=2)
    assert np.array_equal(s1.view(np.uint32), s2.view(np.uint32))
