"""Frame time against slice count for smaller frames of C3 (development aid): slices are capped by the 2 M-samples rule."""
import sys
sys.path.insert(0, ".")
import torch
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes
sc = scenes.c3_mesh()
ctx = crt.Context(0)
ctx.set_scene(sc); ctx.build()
for nx, ny, spp in ((1920, 1080, 16), (960, 1080, 16), (960, 540, 16), (480, 540, 16), (480, 270, 16), (1024, 512, 4)):
    cam = crt.camera_for(sc, nx, ny)
    acc = torch.zeros(ny, nx, 3, device="cuda")
    row = []
    for lanes in (1, 2, 3, 4):
        ctx.set_option(crt.OPT_LANES, lanes)
        for i in range(2):
            ctx.render_accum_device(cam, nx, ny, 0, spp, 8, crt.SHADE_PATH, 1 + i, 0, 1, acc.data_ptr())
        ms = []
        for i in range(5):
            ctx.render_accum_device(cam, nx, ny, 0, spp, 8, crt.SHADE_PATH, 3 + i, 0, 1, acc.data_ptr())
            ms.append(ctx.counters()["render_ms"])
        row.append("%d:%.3f" % (ctx.counters()["lanes"], sorted(ms)[2]))
    print("%dx%dx%d (%.1f M samples): %s" % (nx, ny, spp, nx * ny * spp / 1e6, "  ".join(row)))
