"""One warm C3 frame inside a cudaProfilerStart/Stop range, rendered exactly as bench.py times its roofline
(one slice in flight, CRT_OPT_TIMING = 1: every bounce launched in order). Run plain first, then under
`ncu --profile-from-start off` (scripts/gpu_r2_n.sh)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes

sc = scenes.c3_mesh()
nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
cam = crt.camera_for(sc)
ctx = crt.Context(0)
ctx.set_option(crt.OPT_LANES, 1)
ctx.set_option(crt.OPT_TIMING, 1)
ctx.set_scene(sc)
ctx.build()
acc = torch.zeros(ny, nx, 3, dtype=torch.float32, device="cuda:0")
for i in range(3):
    ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 1 + i, 0, 1, acc.data_ptr())
ctx.sync()
ctx.reset_counters()
torch.cuda.profiler.start()
ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 7, 0, 1, acc.data_ptr())
ctx.sync()
torch.cuda.profiler.stop()
c = ctx.counters()
print("frame: %.3f ms render, closest-hit launches %.3f ms, %d rays" % (c["render_ms"], c["trace_queue_ms"], c["rays"]))
ctx.close()
