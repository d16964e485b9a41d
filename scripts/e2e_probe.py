"""Where the end-to-end frame time goes (development aid): wall clock per frame of the C3 frame through every
form of the per-frame call, on the same box, medians over N frames."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes


def med(f, n=12, warm=3):
    for i in range(warm):
        f(i)
    ts = []
    for i in range(n):
        t0 = time.perf_counter()
        f(100 + i)
        ts.append(1e3 * (time.perf_counter() - t0))
    return float(np.median(ts)), float(np.min(ts))


def main():
    sc = scenes.c3_mesh()
    nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
    cam = crt.camera_for(sc)
    ctx = crt.Context(0)
    ctx.set_scene(sc)
    ctx.build()
    ctx.set_indices(sc["tri_idx"], len(sc["points"]))
    part = (np.arange(len(sc["tri"])) * 8 // len(sc["tri"])).astype(np.uint32)
    ctx.set_parts(part, 8)
    T = np.zeros((8, 3), np.float32)
    host_rgb = torch.empty(ny, nx, 3, dtype=torch.float32).pin_memory()
    host_verts = torch.from_numpy(np.ascontiguousarray(sc["tri"][:, :9])).pin_memory()
    host_points = torch.from_numpy(sc["points"]).pin_memory()
    accum = torch.zeros(ny, nx, 3, dtype=torch.float32, device="cuda")
    out = host_rgb.numpy()

    def dev(seed):
        ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, seed, 0, 1, accum.data_ptr())
        ctx.sync()

    rows = [
        ("device-resident accumulate + sync", dev),
        ("crt_render to a pinned host buffer", lambda s: ctx.render(cam, nx, ny, spp, depth, crt.SHADE_PATH, s, out=out)),
        ("crt_render_frame, no pose (refit only)", lambda s: ctx.render_frame(None, cam, nx, ny, spp, depth, crt.SHADE_PATH, s, out=out)),
        ("crt_render_frame_pose PARTS (96 B)", lambda s: ctx.render_frame_pose(crt.POSE_PARTS, T, cam, nx, ny, spp, depth, crt.SHADE_PATH, s, out=out)),
        ("crt_render_frame_pose POINTS (6 MB)", lambda s: ctx.render_frame_pose(crt.POSE_POINTS, host_points.numpy(), cam, nx, ny, spp, depth, crt.SHADE_PATH, s, out=out)),
        ("crt_render_frame vertices (36 MB)", lambda s: ctx.render_frame(host_verts.numpy(), cam, nx, ny, spp, depth, crt.SHADE_PATH, s, out=out)),
    ]
    for rep in range(2):
        for name, f in rows:
            m, lo = med(f)
            print("%-44s median %.3f ms  min %.3f ms  (refit_ms %.3f)" % (name, m, lo, ctx.counters()["refit_ms"]), flush=True)


if __name__ == "__main__":
    main()
