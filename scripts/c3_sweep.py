"""Run-time knobs on the C3 frame (or one rank's share of it), one process: for each setting 3 warm frames + K timed.
  python scripts/c3_sweep.py [--world 1] [--rank 0] [--steps 20]
Settings are (lanes, batch_paths, wide_below, slice_min); the first line is the default."""
import argparse
import sys

import torch

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes

ap = argparse.ArgumentParser()
ap.add_argument("--world", type=int, default=1)
ap.add_argument("--rank", type=int, default=0)
ap.add_argument("--steps", type=int, default=20)
a = ap.parse_args()
sc = scenes.c3_mesh()
nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
cam = crt.camera_for(sc)
ctx = crt.Context(0)
ctx.set_scene(sc)
ctx.build()
dev = torch.device("cuda", 0)
accum = torch.zeros(ny, nx, 3, dtype=torch.float32, device=dev)
ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
D = dict(lanes=3, batch=1 << 25, wide=32768, smin=2_000_000)
if a.world == 1:
    settings = [{}, dict(lanes=4), dict(lanes=2), dict(batch=1 << 24), dict(batch=1 << 23), dict(batch=1 << 24, lanes=2), dict(batch=1 << 23, lanes=4),
                dict(wide=16384), dict(wide=65536), dict(wide=131072), {}]
else:
    settings = [{}, dict(lanes=1), dict(lanes=3, smin=1_000_000), dict(lanes=4, smin=1_000_000), dict(wide=16384), dict(wide=65536), dict(wide=131072),
                dict(batch=1 << 21), dict(batch=1 << 20), {}]
ref = None
for st in settings:
    o = dict(D)
    o.update(st)
    ctx.set_option(crt.OPT_LANES, o["lanes"])
    ctx.set_option(crt.OPT_BATCH_PATHS, o["batch"])
    ctx.set_option(crt.OPT_WIDE_BELOW, o["wide"])
    ctx.set_option(crt.OPT_SLICE_MIN, o["smin"])
    for i in range(3):
        ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 7, a.rank, a.world, accum.data_ptr())
    ctx.sync()
    if ref is None:
        ref = accum.clone()
    same = bool(torch.equal(ref, accum))
    ctx.reset_counters()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for i in range(a.steps):
        ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 100 + i, a.rank, a.world, accum.data_ptr())
    e1.record(ext)
    ctx.sync()
    c = ctx.counters()
    print("world %d rank %d %-60s %.3f ms/frame, %d slices, %d launches/frame, same image: %s" % (
        a.world, a.rank, str(st or "default"), e0.elapsed_time(e1) / a.steps, c["lanes"], c["launches"] // a.steps, same), flush=True)
del accum, ref, ext
ctx.close()
