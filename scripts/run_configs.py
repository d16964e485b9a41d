"""Runs BASELINE.json's configs C1..C5 on one GPU through the C ABI and prints one
table row per config. Development/report aid, not the bench.
Usage: python scripts/run_configs.py [c1 c2 c3 c4 c5] [--max-tris N]
Under torchrun (N ranks) only c5 runs: every rank builds the same LBVH and traces its own 1/N of the rays (SURVEY 8e:
"rays are split evenly across GPUs against a replicated BVH; no collective"); time = max over ranks, rate = all rays / it:
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/run_configs.py c5"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes


def render_row(name, sc, mode=crt.SHADE_PATH, reps=3):
    ctx = crt.Context(0)
    ctx.set_scene(sc)
    ctx.build()
    ctx.build()
    build_ms = ctx.counters()["build_ms"]
    nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
    cam = crt.camera_for(sc)
    dev = torch.device("cuda", 0)
    accum = torch.zeros(ny, nx, 3, dtype=torch.float32, device=dev)
    ctx.set_option(crt.OPT_COUNTERS, 1)
    ctx.reset_counters()
    ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 1, 0, 1, accum.data_ptr())
    w = ctx.counters()
    ctx.set_option(crt.OPT_COUNTERS, 0)
    ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 2, 0, 1, accum.data_ptr())
    ctx.reset_counters()
    ms = []
    for i in range(reps):
        ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 3 + i, 0, 1, accum.data_ptr())
        ms.append(ctx.counters()["render_ms"])
    c = ctx.counters()
    t = float(np.median(ms))
    rays = c["rays"] / reps
    lanes = int(c["lanes"])
    # the closest-hit launches timed alone (one slice in flight), as bench.py does for the roofline
    ctx.set_option(crt.OPT_LANES, 1)
    ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 2, 0, 1, accum.data_ptr())
    ctx.reset_counters()
    for i in range(reps):
        ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 3 + i, 0, 1, accum.data_ptr())
    c = ctx.counters()
    row = dict(config=name, prims=len(sc["tri"]) + len(sc["sph"]), nx=nx, ny=ny, spp=spp, ms_per_frame=t,
               mrays_s=rays / t / 1e3, mpaths_s=nx * ny * spp / t / 1e3, rays_per_path=rays / (nx * ny * spp),
               nodes_per_ray=w["nodes_visited"] / max(1, w["rays"]), tris_per_ray=w["tris_tested"] / max(1, w["rays"]),
               spheres_per_ray=w["spheres_tested"] / max(1, w["rays"]), build_ms=build_ms, lanes=lanes,
               trace_ms=c["trace_queue_ms"] / reps,
               alg_gbs=(128 * w["node_fetches"] + 48 * w["tri_fetches"] + 16 * w["sphere_fetches"] + 48 * w["rays"]) / (c["trace_queue_ms"] / reps) / 1e6)
    ctx.close()
    del accum
    torch.cuda.empty_cache()
    return row


WORLD = int(os.environ.get("WORLD_SIZE", "1"))
RANK = int(os.environ.get("RANK", "0"))
LOCAL = int(os.environ.get("LOCAL_RANK", "0"))


def trace_rows(n_tri, reps=5):
    t0 = time.time()
    sc = scenes.c5_triangles(n_tri)
    gen = time.time() - t0
    ctx = crt.Context(LOCAL)
    ctx.set_scene(sc)
    ctx.build()
    ctx.build()
    build_ms = ctx.counters()["build_ms"]
    cam = crt.camera_for(sc)
    st = torch.cuda.ExternalStream(ctx.stream())
    rows = []
    prim = scenes.primary_rays(cam, 1920, 1080)
    for kind, rays, accel in (("coherent", prim, crt.ACCEL_LBVH), ("coherent, packets", prim, crt.ACCEL_LBVH_PACKET),
                              ("incoherent", scenes.c5_incoherent_rays(1 << 21), crt.ACCEL_LBVH)):
        # the launch is replicated 8x so that a persistent grid of 1184 blocks has work beyond its ramp-up
        rays = np.ascontiguousarray(np.tile(rays, (8, 1)))
        n_all = len(rays)
        if WORLD > 1:      # this rank's contiguous share
            rays = np.ascontiguousarray(rays[RANK * n_all // WORLD:(RANK + 1) * n_all // WORLD])
        d_rays = torch.from_numpy(rays).cuda(LOCAL)
        d_hits = torch.empty(len(rays), 4, dtype=torch.float32, device="cuda:%d" % LOCAL)
        ctx.set_option(crt.OPT_COUNTERS, 1)
        ctx.reset_counters()
        ctx.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr(), accel=accel)
        ctx.sync()
        w = ctx.counters()
        ctx.set_option(crt.OPT_COUNTERS, 0)
        for _ in range(2):
            ctx.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr(), accel=accel)
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(reps):
            ctx.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr(), accel=accel)
        e1.record(st)
        ctx.sync()
        ms = e0.elapsed_time(e1) / reps
        n = len(rays)
        if WORLD > 1:
            import torch.distributed as dist
            tmax = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            tot = torch.tensor([float(n), float(w["nodes_visited"]), float(w["tris_tested"]), float(w["node_fetches"]),
                                float(w["tri_fetches"])], dtype=torch.float64, device="cuda")
            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
            rows.append(dict(config="C5 %s" % kind, gpus=WORLD, prims=n_tri, rays=int(tot[0]), ms=float(tmax[0]),
                             mrays_s=float(tot[0]) / float(tmax[0]) / 1e3, build_ms=build_ms,
                             nodes_per_ray=float(tot[1] / tot[0]), tris_per_ray=float(tot[2] / tot[0]),
                             alg_gbs=float(128 * tot[3] + 48 * tot[4] + 48 * tot[0]) / float(tmax[0]) / 1e6))
            del d_rays, d_hits
            continue
        rows.append(dict(config="C5 %s" % kind, prims=n_tri, rays=n, ms=ms, mrays_s=n / ms / 1e3, build_ms=build_ms,
                         hit_rate=float((d_hits[:, 1].view(torch.int32) != -1).float().mean()),
                         nodes_per_ray=w["nodes_visited"] / n, tris_per_ray=w["tris_tested"] / n, scene_gen_s=gen,
                         alg_gbs=(128 * w["node_fetches"] + 48 * w["tri_fetches"] + 48 * n) / ms / 1e6))
        del d_rays, d_hits
    ctx.close()
    torch.cuda.empty_cache()
    return rows


def main():
    which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c1", "c2", "c3", "c4", "c5"]
    max_tris = 10 ** 7
    if "--max-tris" in sys.argv:
        max_tris = int(float(sys.argv[sys.argv.index("--max-tris") + 1]))
    out = []
    if WORLD > 1:
        import torch.distributed as dist
        torch.cuda.set_device(LOCAL)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", LOCAL))
        which = [w for w in which if w == "c5"]
    if "c1" in which:
        out.append(render_row("C1 bunny LambertShade", scenes.c1_bunny(), crt.SHADE_LAMBERT, reps=10))
        out.append(render_row("C1 bunny shade()", scenes.c1_bunny(), crt.SHADE_PATH, reps=10))
    if "c2" in which:
        out.append(render_row("C2 spheres", scenes.c2_spheres()))
    if "c3" in which:
        out.append(render_row("C3 1M-tri mesh", scenes.c3_mesh()))
    if "c4" in which:
        out.append(render_row("C4 10M-tri instanced, 1 GPU", scenes.c4_instanced(), reps=2))
    if "c5" in which:
        n = 1000
        if "--min-tris" in sys.argv:
            n = int(float(sys.argv[sys.argv.index("--min-tris") + 1]))
        while n <= max_tris:
            out.extend(trace_rows(n))
            n *= 10
    if RANK == 0:
        for r in out:
            print(json.dumps(r), flush=True)
    if WORLD > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
