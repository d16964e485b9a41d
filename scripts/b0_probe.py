"""Camera-ray launch alone (LambertShade frame of C3): time and work counters per packet mode (development aid)."""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes
sc = scenes.c3_mesh()
ctx = crt.Context(0)
ctx.set_option(crt.OPT_LANES, 1)
ctx.set_scene(sc); ctx.build()
cam = crt.camera_for(sc)
nx, ny, spp = sc["nx"], sc["ny"], sc["spp"]
acc = torch.zeros(ny, nx, 3, device="cuda")
for mode in (2, 1, 0):
    ctx.set_option(crt.OPT_PACKET_PRIMARY, mode)
    ctx.set_option(crt.OPT_COUNTERS, 1); ctx.reset_counters()
    ctx.render_accum_device(cam, nx, ny, 0, spp, 8, crt.SHADE_LAMBERT, 1, 0, 1, acc.data_ptr())
    w = ctx.counters(); ctx.set_option(crt.OPT_COUNTERS, 0)
    ctx.render_accum_device(cam, nx, ny, 0, spp, 8, crt.SHADE_LAMBERT, 1, 0, 1, acc.data_ptr())
    ctx.reset_counters()
    for i in range(3):
        ctx.render_accum_device(cam, nx, ny, 0, spp, 8, crt.SHADE_LAMBERT, 2 + i, 0, 1, acc.data_ptr())
    c = ctx.counters()
    print("packet mode %d: trace %.3f ms  node tests %.1fM fetches %.2fM  tri tests %.1fM fetches %.2fM  rays %.1fM" % (
        mode, c["trace_queue_ms"] / 3, w["nodes_visited"] / 1e6, w["node_fetches"] / 1e6, w["tris_tested"] / 1e6, w["tri_fetches"] / 1e6, w["rays"] / 1e6))
