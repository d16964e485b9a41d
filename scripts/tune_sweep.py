"""Sweep of the run-time tuning options of the per-ray traversal (CRT_OPT_REFILL, CRT_OPT_LEAF_THRESH,
CRT_OPT_LEAF_MAX) on the C3 frame: device-resident frame time, median of 5 (development aid)."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes


def main():
    sc = scenes.c3_mesh()
    nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
    cam = crt.camera_for(sc)
    ctx = crt.Context(0)
    ctx.set_scene(sc)
    accum = torch.zeros(ny, nx, 3, dtype=torch.float32, device="cuda")

    def frame_ms(n=5):
        ts = []
        for i in range(n + 2):
            ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, 1 + i, 0, 1, accum.data_ptr())
            ctx.sync()
            ts.append(ctx.counters()["render_ms"])
        return float(np.median(ts[2:]))

    if len(sys.argv) > 1 and sys.argv[1] == "one":      # just the default configuration
        ctx.build()
        print("defaults: %.3f ms" % frame_ms(9), flush=True)
        return
    for leaf_max in (2, 3, 4):
        ctx.set_option(crt.OPT_LEAF_MAX, leaf_max)
        ctx.build()
        for refill in (4, 8, 12, 16):
            for lt in (4, 8, 12, 16, 24):
                ctx.set_option(crt.OPT_REFILL, refill)
                ctx.set_option(crt.OPT_LEAF_THRESH, lt)
                print("leaf_max %d refill %2d leaf_thresh %2d: %.3f ms" % (leaf_max, refill, lt, frame_ms()), flush=True)


if __name__ == "__main__":
    main()
