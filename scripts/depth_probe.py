"""Development aid: C3 frame time against max_depth (what the last, nearly empty bounces cost)."""
import sys
import torch
sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes

sc = scenes.c3_mesh()
ctx = crt.Context(0)
ctx.set_scene(sc)
ctx.build()
nx, ny, spp = sc["nx"], sc["ny"], sc["spp"]
cam = crt.camera_for(sc)
acc = torch.zeros(ny, nx, 3, dtype=torch.float32, device="cuda")
st = torch.cuda.ExternalStream(ctx.stream())
for lanes in (4, 1):
    ctx.set_option(crt.OPT_LANES, lanes)
    for depth in (8, 4, 3, 2, 1, 0):
        for _ in range(3):
            ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 1, 0, 1, acc.data_ptr())
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for i in range(10):
            ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 10 + i, 0, 1, acc.data_ptr())
        e1.record(st)
        ctx.sync()
        print("lanes %d depth %d: %.3f ms/frame, rays %d" % (lanes, depth, e0.elapsed_time(e1) / 10, ctx.counters()["rays"]))
        ctx.reset_counters()
