"""One rank's share of the C3 (or C4) frame on ONE GPU: renders tiles rank, rank + world, ... of the frame K times and
prints the device time per frame. Development aid for the strong-scaling work (what a rank of an N-GPU run executes,
without the reduce); run it under `ncu --metrics gpu__time_duration.sum` for the per-kernel list.
  python scripts/frame_part.py [--world 8] [--rank 0] [--lanes 4] [--slice-min N] [--steps 10] [--workload c3|c4] [--timing]"""
import argparse
import sys

import torch

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes

ap = argparse.ArgumentParser()
ap.add_argument("--world", type=int, default=8)
ap.add_argument("--rank", type=int, default=0)
ap.add_argument("--lanes", type=int, default=4)
ap.add_argument("--slice-min", type=int, default=0)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--wide-below", type=int, default=-1)
ap.add_argument("--workload", default="c3")
ap.add_argument("--segments", type=int, default=0, help="c3 only: tessellation (default 1000 -> 998,002 triangles)")
ap.add_argument("--tail-hint", type=int, default=1)
ap.add_argument("--packet-defer", type=int, default=-1)
ap.add_argument("--packet-defer-below", type=int, default=-1)
ap.add_argument("--batch-samples", type=int, default=-1, help="CRT_OPT_BATCH_SAMPLES (0: sample batches over all tiles)")
ap.add_argument("--ranks", default="", help="comma-separated ranks to time one after the other on the same scene (default: --rank)")
ap.add_argument("--timing", action="store_true", help="CRT_OPT_TIMING: per-bounce lines on stderr (needs CRT_TRACE_TIMING=1 in the environment)")
a = ap.parse_args()
sc = scenes.c4_instanced() if a.workload == "c4" else (scenes.c3_mesh(n_theta=a.segments, n_phi=a.segments // 2) if a.segments else scenes.c3_mesh())
nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
cam = crt.camera_for(sc)
ctx = crt.Context(0)
ctx.set_option(crt.OPT_LANES, a.lanes)
ctx.set_option(crt.OPT_TAIL_HINT, a.tail_hint)
if a.packet_defer >= 0:
    ctx.set_option(crt.OPT_PACKET_DEFER, a.packet_defer)
if a.packet_defer_below >= 0:
    ctx.set_option(crt.OPT_PACKET_DEFER_BELOW, a.packet_defer_below)
if a.slice_min:
    ctx.set_option(crt.OPT_SLICE_MIN, a.slice_min)
if a.batch_samples >= 0:
    ctx.set_option(crt.OPT_BATCH_SAMPLES, a.batch_samples)
if a.wide_below >= 0:
    ctx.set_option(crt.OPT_WIDE_BELOW, a.wide_below)
ctx.set_scene(sc)
ctx.build()
ctx.set_option(crt.OPT_COUNTERS, 1)
ctx.reset_counters()
_acc = torch.zeros(ny, nx, 3, dtype=torch.float32, device="cuda:0")
ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 1, a.rank, a.world, _acc.data_ptr())
ctx.sync()
w = ctx.counters()
ctx.set_option(crt.OPT_COUNTERS, 0)
print("   nodes/ray %.4f tris/ray %.4f node fetches/ray %.4f tri fetches/ray %.4f" % (w["nodes_visited"] / w["rays"], w["tris_tested"] / w["rays"],
      w["node_fetches"] / w["rays"], w["tri_fetches"] / w["rays"]), flush=True)
del _acc
dev = torch.device("cuda", 0)
accum = torch.zeros(ny, nx, 3, dtype=torch.float32, device=dev)
ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
if a.timing:
    ctx.set_option(crt.OPT_TIMING, 1)
for rk in ([int(r) for r in a.ranks.split(",")] if a.ranks else [a.rank]):
    for i in range(3):
        ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 1 + i, rk, a.world, accum.data_ptr())
    ctx.sync()
    ctx.reset_counters()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for i in range(a.steps):
        ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 100 + i, rk, a.world, accum.data_ptr())
    e1.record(ext)
    ctx.sync()
    c = ctx.counters()
    print("world %d rank %d lanes %d(%d in use) slice_min %d: %.3f ms/frame, %.0f Mrays/s, %d launches/frame" % (
        a.world, rk, a.lanes, c["lanes"], a.slice_min, e0.elapsed_time(e1) / a.steps, c["rays"] / e0.elapsed_time(e1) / 1e3, c["launches"] // a.steps), flush=True)
del accum, e0, e1, ext
ctx.close()
