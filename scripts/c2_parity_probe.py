"""How many pixels of a C2 frame differ between the GPU path and the reference (oracle/_ref or the C restatement)?
The one operation of the path that is not bit-defined is powf in schlick() (material.h:25)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes
from util import best_oracle, oracle_scene

sc = scenes.c2_spheres()
ctx = crt.Context(0)
ctx.set_scene(sc)
ctx.build()
o = oracle_scene(best_oracle(), sc)
tot = 0
for nx, ny, spp, seed in ((150, 100, 4, 2), (150, 100, 4, 7), (300, 200, 8, 1), (300, 200, 8, 2), (240, 160, 16, 3)):
    cam = crt.camera_for(sc, nx, ny)
    img = ctx.render(cam, nx, ny, spp, 8, crt.SHADE_PATH, seed=seed)
    ref = o.render(cam, nx, ny, spp, 8, crt.SHADE_PATH, seed)
    d = int((img != ref).any(axis=2).sum())
    tot += d
    print("%dx%d x %d spp seed %d: %d pixels differ (%d paths)" % (nx, ny, spp, seed, d, nx * ny * spp), flush=True)
print("total", tot)
