"""C4 at full size on one GPU (10 M triangles, 3840x2160, 64 spp): the accumulators of one frame rendered under different
batch sizes, lane counts and tail thresholds must agree bit for bit (VERDICT r1 item 6 i; the 8-GPU bench line of run D
reported 1 differing pixel). Prints where they do not."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes

sc = scenes.c4_instanced()
nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
if len(sys.argv) > 1:
    spp = int(sys.argv[1])
cam = crt.camera_for(sc)
ctx = crt.Context(0)
ctx.set_scene(sc)
ctx.build()
dev = torch.device("cuda", 0)


def frame(name, opts, world=1):
    for k, v in dict(lanes=4, batch=1 << 25, wide=32768, packet=1).items():
        opts.setdefault(k, v)
    ctx.set_option(crt.OPT_LANES, opts["lanes"])
    ctx.set_option(crt.OPT_BATCH_PATHS, opts["batch"])
    ctx.set_option(crt.OPT_WIDE_BELOW, opts["wide"])
    ctx.set_option(crt.OPT_PACKET_PRIMARY, opts["packet"])
    total = torch.zeros(ny, nx, 3, dtype=torch.float32, device=dev)
    part = torch.zeros(ny, nx, 3, dtype=torch.float32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    e0.record(ext)
    for r in range(world):
        ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 7, r, world, part.data_ptr())
        ctx.sync()
        total += part
    e1.record(ext)
    ctx.sync()
    print("%-34s %.1f ms" % (name, e0.elapsed_time(e1)), flush=True)
    return total


base = frame("default", {})
cases = [("default again", {}, 1), ("wide_below 0", dict(wide=0), 1), ("wide_below 2^31-1", dict(wide=(1 << 31) - 1), 1),
         ("batch 2^27", dict(batch=1 << 27), 1), ("batch 2^23", dict(batch=1 << 23), 1), ("lanes 1", dict(lanes=1), 1),
         ("no packets", dict(packet=0), 1), ("8 tile parts", {}, 8)]
for name, opts, world in cases:
    img = frame(name, dict(opts), world)
    d = (img.double() - base.double()).abs().amax(dim=-1)
    bad = torch.nonzero(d > 0)
    print("   vs default: %d pixels differ" % len(bad), flush=True)
    for y, x in bad[:5].tolist():
        print("     pixel (%d, %d): %s vs %s" % (y, x, img[y, x].tolist(), base[y, x].tolist()), flush=True)
ctx.close()
