import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import hobbyraytracer_b200 as pkg
from hobbyraytracer_b200 import abi, scenes
lib = pkg.load()
def check(name, sd, cam, w=72, h=72, spp=8):
    c = scenes.camera_for(cam, w, h)
    g = sd.build(lib, abi.BVH_SAH, 3)
    a, _ = g.render(c, w, h, spp, seed=3)
    b, _ = g.render(c, w, h, spp, seed=3, flags=abi.RENDER_MEGAKERNEL)
    d = (a.view(np.uint32) != b.view(np.uint32)).any(axis=2)
    print(name, "pixels differing", int(d.sum()), "of", w*h, "mean", float(a.mean()), float(b.mean()), flush=True)
sd, cam, _ = scenes.scattered_balls_scene(grid=3); check("balls", sd, cam)
sd, cam, _ = scenes.cornell_smoke_scene(); check("smoke", sd, cam)
sd, cam, _ = scenes.shiny_teapot_scene("tests/_cache", n=12, env_w=64, env_h=32); check("shiny", sd, cam)
sd, cam, _ = scenes.teapot_scene("tests/_cache", n=12, env_w=64, env_h=32); check("teapot", sd, cam)
