"""Quick device-side throughput probe (development aid, not the bench):
builds a scene, times crt_trace_device on primary + incoherent rays and one
render, prints counters. Usage: python scripts/quick_perf.py [c3|c1|c2|c5:N] [leaf_max]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c3"
    leaf_max = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    refill = int(sys.argv[3]) if len(sys.argv) > 3 else 8
    lthr = int(sys.argv[4]) if len(sys.argv) > 4 else 16
    t0 = time.time()
    if which == "c3":
        sc = scenes.c3_mesh()
    elif which == "c1":
        sc = scenes.c1_bunny()
    elif which == "c2":
        sc = scenes.c2_spheres()
    else:
        sc = scenes.c5_triangles(int(which.split(":")[1]))
    print("scene %s: %d tris %d spheres (%.1fs to generate)" % (sc["name"], len(sc["tri"]), len(sc["sph"]), time.time() - t0))
    ctx = crt.Context(0)
    ctx.set_option(crt.OPT_LEAF_MAX, leaf_max)
    ctx.set_option(crt.OPT_REFILL, refill)
    ctx.set_option(crt.OPT_LEAF_THRESH, lthr)
    if os.environ.get("CRT_BUILDER"):
        ctx.set_option(crt.OPT_BUILDER, int(os.environ["CRT_BUILDER"]))
    if os.environ.get("CRT_LANES"):
        ctx.set_option(crt.OPT_LANES, int(os.environ["CRT_LANES"]))
    t0 = time.time()
    ctx.set_scene(sc)
    print("upload %.3fs" % (time.time() - t0))
    for _ in range(3):
        ctx.build()
    print("build_ms %.3f" % ctx.counters()["build_ms"], ctx.accel_info())
    ctx.refit()
    print("refit_ms %.3f" % ctx.counters()["refit_ms"])
    nx, ny = sc["nx"], sc["ny"]
    cam = crt.camera_for(sc)
    lo, hi = scenes.scene_bounds(sc)
    inc = scenes.c5_incoherent_rays(1 << 21, seed=1)
    inc[:, :3] = lo + inc[:, :3] * (hi - lo)
    for name, rays in (("primary", scenes.primary_rays(cam, nx, ny)), ("incoherent", inc)):
        d_rays = torch.from_numpy(rays).cuda()
        d_hits = torch.empty(len(rays), 4, dtype=torch.float32, device="cuda")
        ctx.set_option(crt.OPT_COUNTERS, 1)
        ctx.reset_counters()
        ctx.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr())
        ctx.sync()
        c = ctx.counters()
        ctx.set_option(crt.OPT_COUNTERS, 0)
        st = torch.cuda.ExternalStream(ctx.stream())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            ctx.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr())
        ctx.sync()
        reps = 10
        e0.record(st)
        for _ in range(reps):
            ctx.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr())
        e1.record(st)
        ctx.sync()
        ms = e0.elapsed_time(e1) / reps
        n = len(rays)
        hitf = float((d_hits[:, 1].view(torch.int32) != -1).float().mean())
        byts = 128 * c["nodes_visited"] + 48 * c["tris_tested"] + 16 * c["spheres_tested"] + 40 * n
        print("%-10s %8d rays %.3f ms  %.1f Mrays/s  hit %.2f  nodes/ray %.1f tris/ray %.1f  alg %.1f GB/s" % (
            name, n, ms, n / ms / 1e3, hitf, c["nodes_visited"] / n, c["tris_tested"] / n, byts / ms / 1e6))
    for mode, nm in ((crt.SHADE_PATH, "path"), (crt.SHADE_LAMBERT, "lambert")):
        spp = min(sc["spp"], 16)
        ctx.render(cam, nx, ny, spp, 8, mode, 1)
        ctx.reset_counters()
        t0 = time.time()
        ctx.render(cam, nx, ny, spp, 8, mode, 1)
        wall = time.time() - t0
        c = ctx.counters()
        print("render %-8s %dx%d %dspp: device %.2f ms (wall %.1f ms)  paths %d rays %d  %.1f Mrays/s %.1f Mpaths/s" % (
            nm, nx, ny, spp, c["render_ms"], wall * 1e3, c["paths"], c["rays"], c["rays"] / c["render_ms"] / 1e3,
            c["paths"] / c["render_ms"] / 1e3))


if __name__ == "__main__":
    main()
