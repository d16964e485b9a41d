"""C4 on one GPU (10 M triangles, 4K, 64 spp = 531 M camera samples, 16 x CRT_OPT_BATCH_PATHS): how the frame is cut
into jobs. CRT_OPT_BATCH_SAMPLES = 0 is the old cut (16 batches of 4 samples over every tile), n > 0 cuts it into
groups of tiles rendered with up to n samples per pixel. Prints ms per frame for each setting and checks that the
accumulators are the same bit for bit.  python scripts/c4_batch_ab.py [--steps 3] [--settings 0,64,32,16]"""
import argparse
import sys
import time

import torch

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--settings", default="0,64,32,16")
ap.add_argument("--lanes", default="3")
ap.add_argument("--builder", type=int, default=0, help="CRT_OPT_BUILDER: 0 LBVH (default), 1 PLOC")
a = ap.parse_args()
t0 = time.time()
sc = scenes.c4_instanced()
nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
cam = crt.camera_for(sc)
ctx = crt.Context(0)
ctx.set_option(crt.OPT_BUILDER, a.builder)
ctx.set_scene(sc)
ctx.build()
print("builder %d" % a.builder)
print("scene + build: %.1f s, %d triangles" % (time.time() - t0, len(sc["tri"])), flush=True)
dev = torch.device("cuda", 0)
accum = torch.zeros(ny, nx, 3, dtype=torch.float32, device=dev)
ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
ref = None
for lanes in [int(x) for x in a.lanes.split(",")]:
    ctx.set_option(crt.OPT_LANES, lanes)
    for bs in [int(x) for x in a.settings.split(",")]:
        ctx.set_option(crt.OPT_BATCH_SAMPLES, bs)
        for i in range(2):
            ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 7, 0, 1, accum.data_ptr())
        ctx.sync()
        if ref is None:
            ref = accum.clone()
        same = bool(torch.equal(ref, accum))
        ctx.reset_counters()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        for i in range(a.steps):
            ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 100 + i, 0, 1, accum.data_ptr())
        e1.record(ext)
        ctx.sync()
        c = ctx.counters()
        ms = e0.elapsed_time(e1) / a.steps
        print("lanes %d batch_samples %3d: %8.2f ms/frame, %7.0f Mrays/s, %d launches/frame, same image as the first setting: %s" % (
            lanes, bs, ms, c["rays"] / e0.elapsed_time(e1) / 1e3, c["launches"] // a.steps, same), flush=True)
del accum, ref, ext
ctx.close()
