#!/usr/bin/env python
"""bench.py — headline benchmark of the render hot path (contract: see the task brief).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun ... bench.py --gpus N ...      (one rank per GPU; the exchange step is the library's own NCCL communicator)

Workload (BASELINE.json configs[2], "C3"): synthetic 998,002-triangle mesh,
1920x1080, 16 spp, depth 8, diffuse+metal, the reference's shade() path tracer.
A step is one full frame: 33,177,600 camera samples carried to termination.

  value   Mrays/s with the scene, LBVH and framebuffers resident in HBM
          (crt_render_accum_device + crt_reduce_accum_device + resolve), CUDA
          events on the context's stream, max over ranks.
          N > 1, --scaling strong (the default: BASELINE.json's "1920x1080 at 16 spp ... 1/2/4/8 B200"): the SAME
          16-spp frame, its 32x32 tiles dealt to the ranks, one ncclReduce of the accumulators onto rank 0; the
          reduced frame is compared bit for bit with rank 0's own one-GPU render of the same seed (`parity_n`).
          --scaling weak: every rank renders 16 samples of every pixel on its own sample range (16 N spp combined).
  e2e     the reference's per-frame loop (core/render.h:207-234) through the C
          ABI with HOST buffers: pose upload (the mesh's point array, as the
          reference's updateFBXObj uploads it; --e2e-api frame uploads every
          triangle's vertices instead) -> LBVH refit -> render -> image download,
          wall clock around the calls.
  roofline  the closest-hit launches (k_trace_packet + k_trace_queue): bytes their own load/store instructions move
          (96 B per node visit, 64 B per primitive record, 48 B per ray; fetch counts from an instrumented run of the
          same frame) / the launches' device time (CUDA events around each launch, one slice in flight).
  cpu_baseline  the reference's own hit/scatter code (oracle/_ref, unmodified headers, OpenMP) through its shipping
          accelerator, BVHNode (hitable/bvh.h:76-190, what kernel.cu:24-39 renders through), built host-side (build
          time reported separately), on a bounded sample of the same frame.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s"
WORKLOAD = "C3: 998,002-triangle displaced sphere + ground, 1920x1080, 16 spp, max_depth 8, shade() path tracer"
# --workload c4: BASELINE.json configs[3], not the headline line (the driver never passes it)
WORKLOAD_C4 = "C4: 10 x 10 instances of a 100,000-triangle displaced sphere + ground (10.0 M triangles, flattened), 3840x2160, 64 spp, max_depth 8, shade() path tracer"

# Bytes the closest-hit kernels move per unit of work: what their own load/store instructions read and
# write, not the stride of the records.
BYTES_NODE = 96     # crt_traverse.cuh:load_node, three 256-bit loads of a 4-wide node (128 B stride)
BYTES_REC = 64      # two 256-bit loads of a primitive record (triangle or sphere)
BYTES_RAY = 48      # queue words 0-1 in (32 B) + hit record out (16 B)
FLOPS_NODE, FLOPS_TRI, FLOPS_SPH = 96, 50, 27     # SURVEY §8d, 4 boxes per node


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


def alg_bytes(w):
    return (BYTES_NODE * w["node_fetches"] + BYTES_REC * (w["tri_fetches"] + w["sphere_fetches"]) + BYTES_RAY * w["rays"])


def alg_flops(w):
    return FLOPS_NODE * w["nodes_visited"] + FLOPS_TRI * w["tris_tested"] + FLOPS_SPH * w["spheres_tested"]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        if os.environ.get("CRT_BENCH_SAMPLER", "smi") == "off":      # development aid: what the sampler itself costs
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def wait_first(self, timeout=5.0):
        """nvidia-smi needs a few hundred ms to start: block until it has printed one row."""
        t_end = time.perf_counter() + timeout
        while self.proc and not self.rows and time.perf_counter() < t_end:
            time.sleep(0.01)

    def stop(self, t0=None, t1=None):
        """Rows that arrived inside [t0, t1] (the timed region); when the region is shorter than three sampling
        periods, every row since start() is used instead (warm-up + timed + e2e, all under the same load)."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        rows, window = [r for _, r in self.rows], "sampler lifetime"
        if t0 is not None:
            inside = [r for t, r in self.rows if t0 <= t <= t1]
            if len(inside) >= 3:
                rows, window = inside, "timed region"
            else:
                window = "warm-up + timed region + e2e loop (timed region shorter than 3 sampling periods)"
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            parts = [p.strip() for p in r.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "window": window}


def build_ref_scene(lib, sc):
    s = lib.scene(True)
    for m in sc["mats"]:
        s.add_material(*m)
    nt, ns = len(sc["tri"]), len(sc["sph"])
    if nt:
        s.add_triangles(sc["tri"], sc["tri_mat"], np.arange(nt, dtype=np.uint32))
    if ns:
        s.add_spheres(sc["sph"], sc["sph_mat"], np.arange(nt, nt + ns, dtype=np.uint32))
    return s.finalize()


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU reference is meant to run on every host core this
    process may use, so the OpenMP runtime it shares with this process is told so explicitly."""
    import ctypes
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(int(n))
    except OSError:
        pass


def cpu_reference_setup(sc, want_bvh=True):
    """The reference's CPU implementation of the path: oracle/_ref (unmodified headers) when the prebuilt library is
    there, else the C restatement. With oracle/_ref the reference's own accelerator (BVHNode, hitable/bvh.h:76-125,
    one thread as create_BVH<<<1,1>>> is) is built; `accel` says what ref_render then walks."""
    import oracle
    kind = "ref" if oracle.available("ref") else "port"
    if not oracle.available(kind):
        oracle.build(ref=False, port=True)
    lib = oracle.load(kind)
    use_all_host_threads()
    sh = build_ref_scene(lib, sc)
    info = {"kind": "reference" if kind == "ref" else "port", "accel": "HitableList scan", "bvh_build_s": None}
    max_build = float(os.environ.get("CRT_BENCH_REF_BVH", "1"))
    if kind == "ref" and want_bvh and max_build > 0:
        info["bvh_build_s"] = sh.bvh_build(1)
        info["accel"] = "BVHNode (hitable/bvh.h:76-190, median split on a random axis, both children always visited)"
    else:
        info["accel_reason"] = ("the C restatement has no BVHNode" if kind != "ref" else "BVHNode switched off (CRT_BENCH_REF_BVH=0)")
    return lib, sh, info


def cpu_sample_shape(scene_h, cam_fn, sc, budget_s, max_nx=None):
    """Pick a reduced frame (same scene, same camera, same aspect, 1 spp) whose
    render through the current accelerator takes about budget_s on this host."""
    import cudaraytracer_b200 as crt
    nx0, ny0 = 16, 9
    scene_h.render(cam_fn(nx0, ny0), nx0, ny0, 1, sc["max_depth"], crt.SHADE_PATH, 1)
    per_px = max(scene_h.last_seconds / (nx0 * ny0), 1e-7)
    k = max(1.0, (budget_s / per_px / (nx0 * ny0)) ** 0.5)
    k = min(k, (max_nx or sc["nx"]) / nx0)
    return max(nx0, int(nx0 * k)), max(ny0, int(ny0 * k))


def run_reference(args):
    """--impl reference: the reference's own CPU hit/scatter code on the host cores through its shipping
    accelerator (BVHNode), every step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import cudaraytracer_b200 as crt
    from cudaraytracer_b200 import scenes
    if args.workload != "c3":
        print(json.dumps({"impl": "reference", "unavailable": "the reference's single-thread BVHNode build over 10 M triangles takes minutes and its list scan manages about one ray per second; the reference arm is measured on C3 only"}), flush=True)
        return
    seg = int(os.environ.get("CRT_BENCH_C3_SEGMENTS", "0"))          # tests shrink the mesh (and say so in the line)
    sc = scenes.c3_mesh(n_theta=seg, n_phi=seg // 2) if seg else scenes.c3_mesh()
    lib, sh, info = cpu_reference_setup(sc)
    bvh = sh.use_bvh(True) if info["bvh_build_s"] is not None else False
    cam_fn = lambda nx, ny: crt.camera_for(sc, nx, ny)
    steps, warm = args.steps, args.warmup
    budget = max(2.0, min(20.0, 120.0 / max(1, steps + warm)))
    budget = float(os.environ.get("CRT_BENCH_REF_BUDGET", budget))   # tests shorten it
    nx, ny = cpu_sample_shape(sh, cam_fn, sc, budget)
    cam = cam_fn(nx, ny)
    for _ in range(warm):
        sh.render(cam, nx, ny, 1, sc["max_depth"], crt.SHADE_PATH, 1)
    secs, rays, paths = 0.0, 0, 0
    for i in range(steps):
        sh.render(cam, nx, ny, 1, sc["max_depth"], crt.SHADE_PATH, 1 + i)
        secs += sh.last_seconds
        rays += sh.last_rays
        paths += nx * ny
    value = rays / secs / 1e6
    sample = "%dx%d px x 1 spp of the 1920x1080x16 frame per step (same scene, camera, depth); %s over %d triangles" % (
        nx, ny, "the reference's BVHNode" if bvh else "list scan", len(sc["tri"]))
    cb = {"value": value, "unit": "Mrays/s", "cores": lib.num_threads(), "kind": info["kind"], "sample": sample,
          "accel": info["accel"], "bvh_build_s": info["bvh_build_s"]}
    if bvh:
        npr, ppr = sh.bvh_stats(scenes.primary_rays(cam_fn(192, 108), 192, 108))
        cb["box_tests_per_camera_ray"], cb["prims_scanned_per_camera_ray"] = npr, ppr
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": 1e3 * secs / max(1, steps), "higher_is_better": True,
        "scaling": args.scaling or "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "paths_per_s": paths / secs,
        "config": {"workload": WORKLOAD if not seg else "TEST MESH of %d triangles (CRT_BENCH_C3_SEGMENTS)" % len(sc["tri"]),
                   "sample": sample, "accelerator": info["accel"], "bvh_build_s": info["bvh_build_s"]},
        "cpu_baseline": cb,
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def other_configs(crt, scenes, torch, dev, local, peaks):
    """BASELINE.json's other single-GPU configurations, a few frames each, so that they carry a driver-run number:
    C1 (the reference's main(): bunny, 1024x512, 4 spp, LambertShade and shade()), C2 (spheres 1200x800x64),
    C5 at 1e6 triangles (coherent camera rays x8 and 16.8 M incoherent rays through crt_trace_device)."""
    out = {}

    def render_cfg(name, sc, mode, reps):
        ctx = crt.Context(local)
        try:
            ctx.set_scene(sc)
            ctx.build()
            nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
            cam = crt.camera_for(sc)
            accum = torch.zeros(ny, nx, 3, dtype=torch.float32, device=dev)
            ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
            ctx.set_option(crt.OPT_COUNTERS, 1)
            ctx.reset_counters()
            ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 1, 0, 1, accum.data_ptr())
            w = ctx.counters()
            ctx.set_option(crt.OPT_COUNTERS, 0)
            for i in range(3):
                ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 2 + i, 0, 1, accum.data_ptr())
            ctx.sync()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext)
            for i in range(reps):
                ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 10 + i, 0, 1, accum.data_ptr())
            e1.record(ext)
            ctx.sync()
            ms = e0.elapsed_time(e1) / reps
            ctx.set_option(crt.OPT_LANES, 1)
            ctx.set_option(crt.OPT_TIMING, 1)
            ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 2, 0, 1, accum.data_ptr())
            ctx.reset_counters()
            for i in range(reps):
                ctx.render_accum_device(cam, nx, ny, 0, spp, depth, mode, 10 + i, 0, 1, accum.data_ptr())
            tq = ctx.counters()["trace_queue_ms"] / reps
            out[name] = {"ms_per_frame": ms, "mrays_per_s": w["rays"] / ms / 1e3, "paths_per_s": nx * ny * spp / ms * 1e3,
                         "prims": len(sc["tri"]) + len(sc["sph"]), "nx": nx, "ny": ny, "spp": spp,
                         "rays_per_path": w["rays"] / (nx * ny * spp),
                         "roofline_frac": alg_bytes(w) / (tq * 1e-3) / 1e9 / peaks["hbm_gbs"] if tq > 0 else None,
                         "closest_hit_ms": tq}
            del accum, e0, e1, ext
        finally:
            ctx.close()

    def trace_cfg(n_tri, reps=5):
        sc = scenes.c5_triangles(n_tri)
        ctx = crt.Context(local)
        try:
            ctx.set_scene(sc)
            ctx.build()
            ctx.build()
            build_ms = ctx.counters()["build_ms"]
            cam = crt.camera_for(sc)
            ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
            prim = scenes.primary_rays(cam, 1920, 1080)
            for kind, rays in (("coherent", prim), ("incoherent", scenes.c5_incoherent_rays(1 << 21))):
                rays = np.ascontiguousarray(np.tile(rays, (8, 1)))     # 8 x so that the persistent grid has work beyond its ramp-up
                d_rays = torch.from_numpy(rays).to(dev)
                d_hits = torch.empty(len(rays), 4, dtype=torch.float32, device=dev)
                ctx.set_option(crt.OPT_COUNTERS, 1)
                ctx.reset_counters()
                ctx.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr())
                ctx.sync()
                w = ctx.counters()
                ctx.set_option(crt.OPT_COUNTERS, 0)
                for _ in range(2):
                    ctx.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr())
                ctx.sync()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(ext)
                for _ in range(reps):
                    ctx.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr())
                e1.record(ext)
                ctx.sync()
                ms = e0.elapsed_time(e1) / reps
                n = len(rays)
                w["rays"] = n
                out["C5_%.0e_%s" % (n_tri, kind)] = {
                    "ms_per_launch": ms, "mrays_per_s": n / ms / 1e3, "rays": n, "prims": n_tri, "lbvh_build_ms": build_ms,
                    "nodes_per_ray": w["nodes_visited"] / n, "tris_per_ray": w["tris_tested"] / n,
                    "roofline_frac": alg_bytes(w) / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"]}
                del d_rays, d_hits, e0, e1
            del ext
        finally:
            ctx.close()

    render_cfg("C1_bunny_1024x512x4_LambertShade", scenes.c1_bunny(), crt.SHADE_LAMBERT, 20)
    render_cfg("C1_bunny_1024x512x4_shade", scenes.c1_bunny(), crt.SHADE_PATH, 20)
    render_cfg("C2_spheres_1200x800x64", scenes.c2_spheres(), crt.SHADE_PATH, 5)
    trace_cfg(10 ** 6)
    torch.cuda.empty_cache()
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    import cudaraytracer_b200 as crt
    from cudaraytracer_b200 import scenes, multi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    scaling = args.scaling or "strong"      # the configured frame is fixed as N grows (total work fixed)
    weak = scaling == "weak"

    if args.workload == "c4":
        sc = scenes.c4_instanced()
        args.no_cpu_baseline = True      # the reference's single-thread BVH build over 10 M triangles takes minutes
        args.no_configs = True
        args.e2e_api = "parts" if args.e2e_api == "points" else args.e2e_api   # C4's pose: one translation per instance
    else:
        sc = scenes.c3_mesh()
    workload = WORKLOAD_C4 if args.workload == "c4" else WORKLOAD
    nx, ny, spp, depth = sc["nx"], sc["ny"], sc["spp"], sc["max_depth"]
    cam = crt.camera_for(sc)
    ctx = crt.Context(local)
    ctx.set_option(crt.OPT_LANES, args.lanes)
    if args.packet_defer >= 0:
        ctx.set_option(crt.OPT_PACKET_DEFER, args.packet_defer)
    if args.slice_min:
        ctx.set_option(crt.OPT_SLICE_MIN, args.slice_min)
    if args.batch_samples >= 0:
        ctx.set_option(crt.OPT_BATCH_SAMPLES, args.batch_samples)
    if args.builder == "ploc":
        ctx.set_option(crt.OPT_BUILDER, crt.BUILDER_PLOC)
    ctx.set_scene(sc)
    ctx.build()
    build_cold_ms = ctx.counters()["build_ms"]      # first build: includes every allocation
    ctx.build()
    build_ms = ctx.counters()["build_ms"]           # the reference's "build" column, warm
    ctx.refit()
    refit_ms = ctx.counters()["refit_ms"]           # the reference's "update" column (Update_BVH), warm
    info = ctx.accel_info()
    if world > 1:
        multi.init_comm(ctx)                        # the library's own NCCL communicator (csrc/comm.cu)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    accum = torch.zeros(ny, nx, 3, dtype=torch.float32, device=dev)
    rgb = torch.zeros(ny, nx, 3, dtype=torch.float32, device=dev)
    host_rgb = torch.empty(ny, nx, 3, dtype=torch.float32).pin_memory()
    host_verts = torch.from_numpy(np.ascontiguousarray(sc["tri"][:, :9])).pin_memory()
    host_points = None
    if "points" in sc:
        host_points = torch.from_numpy(sc["points"]).pin_memory()
        ctx.set_indices(sc["tri_idx"], len(sc["points"]))
    host_parts = None
    if args.e2e_api == "parts":
        # the instanced scene's pose form: every instance (and the ground) moves as one part, a frame uploads one
        # translation per part (the reference's nowTransform, createScene.h:221-307) instead of 360 MB of vertices.
        # The translations are zero, so the e2e loop renders the frame the device-resident loop renders
        if "part_of_tri" not in sc:
            raise SystemExit("--e2e-api parts needs a scene with a part per triangle (--workload c4)")
        ctx.set_parts(sc["part_of_tri"], int(sc["n_parts"]))
        host_parts = torch.zeros(int(sc["n_parts"]), 3, dtype=torch.float32).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    gather = (not weak) and args.combine == "gather"

    def step(seed, marks=None):
        if marks is not None:
            marks[0].record(ext)
        if weak:
            # every rank renders the named frame's full work: its own range of 16 samples of every pixel
            ctx.render_accum_device(cam, nx, ny, spp * rank, spp * (rank + 1), depth, crt.SHADE_PATH, seed, 0, 1,
                                    accum.data_ptr())
        else:
            # the frame is fixed; ranks split its 32x32 tiles (bit-identical to the 1-GPU image)
            ctx.render_accum_device(cam, nx, ny, 0, spp, depth, crt.SHADE_PATH, seed, rank, world, accum.data_ptr())
        if marks is not None:
            marks[1].record(ext)
        if world > 1:
            if gather:
                # every pixel was rendered by one rank: its tiles go straight to rank 0 (1/N of the frame per rank)
                ctx.gather_tiles_device(accum.data_ptr(), nx, ny, 0)
            else:
                ctx.reduce_accum_device(accum.data_ptr(), nx * ny * 3, 0)
        if marks is not None:
            marks[2].record(ext)
        if rank == 0:
            ctx.resolve_device(accum.data_ptr(), nx * ny, spp * world if weak else spp, rgb.data_ptr())

    def e2e_step(seed):
        if world == 1 and args.e2e_api == "parts":
            ctx.render_frame_pose(crt.POSE_PARTS, host_parts.numpy(), cam, nx, ny, spp, depth, crt.SHADE_PATH, seed,
                                  out=host_rgb.numpy())
        elif world == 1 and args.e2e_api == "points":
            # the reference's loop body (render.h:206-233) as one call, the pose in the form updateFBXObj uploads it:
            # the point array (createScene.h:125-132); the face indices were set once
            ctx.render_frame_pose(crt.POSE_POINTS, host_points.numpy(), cam, nx, ny, spp, depth, crt.SHADE_PATH, seed,
                                  out=host_rgb.numpy())
        elif world == 1 and args.e2e_api == "frame":
            # the reference's loop body (render.h:206-233) as one call: upload + refit overlap ray generation,
            # finished rows are read back while the rest of the frame renders
            ctx.render_frame(host_verts.numpy(), cam, nx, ny, spp, depth, crt.SHADE_PATH, seed, out=host_rgb.numpy())
        elif world == 1:
            ctx.update_vertices(host_verts.numpy())
            ctx.refit()
            ctx.render(cam, nx, ny, spp, depth, crt.SHADE_PATH, seed, out=host_rgb.numpy())
        else:
            # one process per GPU: every rank queues its pose upload + refit (crt_scene_pose_async) and renders its
            # share; the camera rays are generated while the pose lands
            if args.e2e_api == "parts":
                ctx.pose_async(crt.POSE_PARTS, host_parts.numpy())
            elif args.e2e_api == "points":
                ctx.pose_async(crt.POSE_POINTS, host_points.numpy())
            elif args.e2e_api == "frame":
                ctx.pose_async(crt.POSE_VERTICES, host_verts.numpy())
            else:
                ctx.update_vertices(host_verts.numpy())
                ctx.refit()
            step(seed)
            if rank == 0:
                with torch.cuda.stream(ext):
                    host_rgb.copy_(rgb, non_blocking=True)
            ctx.sync()

    # ---- instrumented frame: work counters for the roofline (timing off) ----
    ctx.set_option(crt.OPT_COUNTERS, 1)
    ctx.reset_counters()
    step(1)
    ctx.sync()
    work = ctx.counters()
    ctx.set_option(crt.OPT_COUNTERS, 0)

    # ---- device-resident timing ----
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        sampler.wait_first()
    for i in range(args.warmup):
        step(1 + i)
    ctx.sync()
    ctx.reset_counters()
    barrier()
    t_wall0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for i in range(args.steps):
        step(100 + i)
    e1.record(ext)
    barrier()
    t_wall1 = time.perf_counter()
    ms = e0.elapsed_time(e1)
    c = ctx.counters()
    lanes, launches = int(c["lanes"]), int(c["launches"])
    # the roofline kernels timed alone: the same frames with one slice in flight and an event pair around each
    # closest-hit launch (in the timed region above up to CRT_OPT_LANES slices overlap, so an event pair would also
    # span the other slices' kernels)
    ctx.set_option(crt.OPT_LANES, 1)
    ctx.set_option(crt.OPT_TIMING, 1)
    step(50)
    ctx.sync()
    ctx.reset_counters()
    for i in range(args.steps):
        step(100 + i)
    ctx.sync()
    alone_ms = ctx.counters()["trace_queue_ms"]
    ctx.set_option(crt.OPT_TIMING, 0)
    ctx.set_option(crt.OPT_LANES, args.lanes)
    stats = torch.tensor([ms, float(c["rays"]), float(c["paths"]), alone_ms, float(launches)], dtype=torch.float64, device=dev)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms, rays, paths, bounce_ms, launches = float(mx[0]), float(sm[1]), float(sm[2]), float(mx[3]), int(sm[4])
    else:
        rays, paths, bounce_ms = float(c["rays"]), float(c["paths"]), alone_ms
    value = rays / (ms * 1e-3) / 1e6

    # ---- N > 1: where a step's time goes on every rank (one extra frame with events around its parts) ----
    per_rank = None
    if world > 1:
        marks = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        barrier()
        step(60, marks)
        ctx.sync()
        mine = torch.tensor([marks[0].elapsed_time(marks[1]), marks[1].elapsed_time(marks[2])], dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"render_ms": [round(float(t[0]), 3) for t in allr],
                    "exchange_ms": [round(float(t[1]), 3) for t in allr],
                    "note": "one frame after a barrier; exchange_ms includes waiting for the slowest rank's render"}

    # ---- N > 1: is the combined frame the right image? ----
    parity_n = None
    if world > 1:
        step(7)
        ctx.sync()
        barrier()
        if rank == 0:
            combined = accum.clone()
            ctx.render_accum_device(cam, nx, ny, 0, spp * world if weak else spp, depth, crt.SHADE_PATH, 7, 0, 1, accum.data_ptr())
            ctx.sync()
            d = (combined.double() - accum.double()).abs()
            parity_n = {"against": "rank 0's own one-GPU render of the same frame and seed (%d spp), accumulators before resolve" % (spp * world if weak else spp),
                        "pixels_differing": int((d.amax(dim=-1) > 0).sum().item()), "max_abs_diff": float(d.max().item()),
                        "expected": "bit-identical (each pixel is non-zero on one rank; adding zeros is exact)" if not weak else
                                    "equal up to the order of float additions (the sum over ranks adds sample groups)"}
            del combined
        barrier()

    # ---- end to end through host buffers ----
    for i in range(args.warmup):
        e2e_step(1 + i)
    ctx.reset_counters()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        e2e_step(100 + i)
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    ce = ctx.counters()
    e2e_stats = torch.tensor([e2e_s, float(ce["rays"])], dtype=torch.float64, device=dev)
    if world > 1:
        a = e2e_stats.clone()
        dist.all_reduce(a, op=dist.ReduceOp.MAX)
        b = e2e_stats.clone()
        dist.all_reduce(b, op=dist.ReduceOp.SUM)
        e2e_s, e2e_rays = float(a[0]), float(b[1])
    else:
        e2e_rays = float(ce["rays"])
    e2e_value = e2e_rays / e2e_s / 1e6

    if rank == 0:
        peaks, peak_src = measured_peaks()
        n_r = work["rays"]
        bytes_frame = alg_bytes(work)
        flops_frame = alg_flops(work)
        bounce_launches = depth + 1
        per_launch_ms = bounce_ms / max(1, args.steps * bounce_launches)
        achieved = bytes_frame / bounce_launches / (per_launch_ms * 1e-3) / 1e9
        traffic, traffic_source, limiter = None, None, None
        tpath = os.path.join(ROOT, "profiles", "closest_hit_ncu.json")
        if os.path.exists(tpath) and args.workload == "c3":
            with open(tpath) as f:
                prof = json.load(f)
            if world == 1 or weak:
                traffic = prof.get("dram_bytes_per_launch")
            traffic_source = prof.get("source")
            limiter = prof.get("limiter")
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "paths_per_s": paths / (ms * 1e-3),
            "config": {"workload": workload, "triangles": int(info["n_prims"]), "builder": args.builder, "lbvh_depth": info["max_depth"],
                       "lbvh_build_ms": build_ms, "lbvh_build_cold_ms": build_cold_ms, "lbvh_refit_ms": refit_ms,
                       "reference_csv_columns": "rendering = ms_per_step, update = lbvh_refit_ms (+ the pose upload, inside e2e), build = lbvh_build_ms (kernel.cu:61, core/render.h:212-230)",
                       "lanes": lanes, "parallelism": ("single" if world == 1 else
                                       "samples%d: every rank renders 16 spp of every pixel (sample range [16r, 16r+16)), %d spp in the combined frame" % (world, spp * world)
                                       if weak else "tiles%d: the fixed %d-spp frame, 32x32 tiles, tile index mod N" % (world, spp)),
                       "combine": "none" if world == 1 else
                                  ("crt_gather_tiles_device: every rank sends the tiles it rendered (%.1f MB, 1/N of the frame) to rank 0 with ncclSend/ncclRecv, issued by libcrt_b200.so on its own communicator" % (nx * ny * 12 / 1e6 / world)) if gather else
                                  ("crt_reduce_accum_device: ncclReduce(sum) of the %.1f MB accumulators onto rank 0, issued by libcrt_b200.so on its own communicator" % (nx * ny * 12 / 1e6)),
                       "per_rank": per_rank,
                       "l2": "per-step working set (queues %.1f GB + LBVH %.2f GB) exceeds the 126 MB L2; no explicit flush" % (
                           min(nx * ny * spp, 1 << 25) * 128 / 1e9, info["bytes"] / 1e9),
                       "rays_per_path": rays / max(1.0, paths)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": 1e3 * e2e_s / args.steps,
                    "h2d_bytes_per_step": int((host_parts.numel() if args.e2e_api == "parts" else host_points.numel() if args.e2e_api == "points" else host_verts.numel()) * 4 + 88),
                    "d2h_bytes_per_step": int(nx * ny * 12),
                    "what": (("crt_render_frame_pose(CRT_POSE_PARTS): upload of one translation per part (%d instances + the ground; the reference's nowTransform, createScene.h:221-307) + move of every triangle with its part + refit + render + read-back, overlapped, per frame, pinned host buffers, wall clock" % (host_parts.shape[0] - 1))
                             if world == 1 and args.e2e_api == "parts" else
                             ("crt_render_frame_pose(CRT_POSE_POINTS): upload of the %d mesh points (the form the reference's updateFBXObj uploads, createScene.h:125-132) + gather into the triangles + refit + render + read-back, overlapped, per frame, pinned host buffers, wall clock" % (host_points.numel() // 3))
                             if world == 1 and args.e2e_api == "points" else
                             "crt_render_frame (upload of all triangle vertices + refit + render + read-back, overlapped) per frame, pinned host buffers, wall clock"
                             if world == 1 and args.e2e_api == "frame" else
                             "crt_scene_update_vertices + crt_accel_refit + crt_render per frame, host buffers, wall clock"
                             if world == 1 else
                             "per rank: %s, render of the rank's share, tile gather (or reduce); rank 0: resolve + read-back to a pinned host buffer; wall clock, max over ranks" % (
                                 "crt_scene_update_vertices + crt_accel_refit" if args.e2e_api == "calls" else
                                 "crt_scene_pose_async(%s) (upload + refit beside ray generation)" % (
                                     "one translation per part" if args.e2e_api == "parts" else "points" if args.e2e_api == "points" else "vertices")))},
            "gpu_launches": launches,
            "roofline": {"kernel": "closest hit over the 4-wide LBVH: k_trace_packet (camera rays) + k_trace_queue (bounces), %d launches/step" % bounce_launches,
                         "bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved / peaks["hbm_gbs"], "traffic": traffic, "traffic_source": traffic_source,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": bytes_frame / bounce_launches,
                         "algorithmic_flops_per_launch": flops_frame / bounce_launches,
                         "avg_launch_ms": per_launch_ms,
                         "timing": "CUDA events around each closest-hit launch, on its stream, over %d frames rendered with one slice in flight (CRT_OPT_LANES=1, CRT_OPT_TIMING=1); the timed region of `value` keeps %d slices in flight" % (args.steps, lanes),
                         "nodes_per_ray": work["nodes_visited"] / max(1, n_r), "tris_per_ray": work["tris_tested"] / max(1, n_r),
                         "node_fetches_per_ray": work["node_fetches"] / max(1, n_r), "tri_fetches_per_ray": work["tri_fetches"] / max(1, n_r),
                         "rays_per_launch": n_r / bounce_launches,
                         "accounting": "bytes = %d B x node fetches (the three 256-bit loads of load_node) + %d B x primitive-record fetches (two 256-bit loads) + %d B x rays (32 B queue entry in, 16 B hit out); camera rays share one fetch (and, with a common origin, one interval box test) per packet of <= 32 rays, every other bounce fetches and tests per ray; flops = 96 x node tests + 50 x triangle tests" % (BYTES_NODE, BYTES_REC, BYTES_RAY),
                         "limiter": limiter,
                         "fp32_tflops_achieved": flops_frame / bounce_launches / (per_launch_ms * 1e-3) / 1e12},
        }
        if parity_n is not None:
            line["parity_n"] = parity_n
        if world == 1 and not args.no_configs:
            try:
                line["configs"] = other_configs(crt, scenes, torch, dev, local, peaks)
            except Exception as e:                       # the headline line must survive a failure here
                line["configs"] = {"error": repr(e)}
        if world == 1 and not args.no_cpu_baseline:
            lib, sh, cinfo = cpu_reference_setup(sc)
            cam_fn = lambda a, b: crt.camera_for(sc, a, b)
            # (1) parity at full scene size: a sub-frame through the reference's list scan (the parity oracle: the
            #     GPU path reproduces its hits bit for bit) and through the GPU path
            pnx, pny = cpu_sample_shape(sh, cam_fn, sc, 6.0)
            img_cpu = np.asarray(sh.render(cam_fn(pnx, pny), pnx, pny, 1, depth, crt.SHADE_PATH, 1), np.float32)
            list_rate = sh.last_rays / sh.last_seconds / 1e6
            img_gpu = ctx.render(cam_fn(pnx, pny), pnx, pny, 1, depth, crt.SHADE_PATH, 1)
            diff = img_gpu.astype(np.float64) - img_cpu.astype(np.float64)
            cb = {"rmse_gpu_vs_reference": float(np.sqrt(np.mean(diff * diff))), "max_abs_diff": float(np.abs(diff).max()),
                  "pixels_differing": int((np.abs(diff).max(axis=-1) > 0).sum()),
                  "parity_sample": "%dx%d px x 1 spp, the reference's HitableList scan over %d triangles (%.1f s, %.6f Mrays/s)" % (
                      pnx, pny, len(sc["tri"]), sh.last_seconds, list_rate)}
            # (2) the timed baseline: the reference's shipping accelerator
            if cinfo["bvh_build_s"] is not None and sh.use_bvh(True):
                snx, sny = cpu_sample_shape(sh, cam_fn, sc, 12.0)
                img_bvh = np.asarray(sh.render(cam_fn(snx, sny), snx, sny, 1, depth, crt.SHADE_PATH, 1), np.float32)
                cb.update({"value": sh.last_rays / sh.last_seconds / 1e6, "unit": "Mrays/s", "cores": lib.num_threads(),
                           "kind": cinfo["kind"], "accel": cinfo["accel"], "bvh_build_s": cinfo["bvh_build_s"],
                           "sample": "%dx%d px x 1 spp of the 1920x1080x16 frame (same scene, camera, depth), %.1f s, through the reference's BVHNode over %d triangles"
                                     % (snx, sny, sh.last_seconds, len(sc["tri"]))})
                npr, ppr = sh.bvh_stats(scenes.primary_rays(cam_fn(192, 108), 192, 108))
                cb["box_tests_per_camera_ray"], cb["prims_scanned_per_camera_ray"] = npr, ppr
                # how far the reference's two accelerators are apart on the parity sub-frame (its AABB::hit rejects a
                # box of zero thickness, core/aabb.h:38, so axis-aligned flat triangles such as the ground quad are
                # invisible to its BVH)
                img_b2 = np.asarray(sh.render(cam_fn(pnx, pny), pnx, pny, 1, depth, crt.SHADE_PATH, 1), np.float32)
                cb["reference_bvh_vs_its_list_scan_pixels_differing"] = int((np.abs(img_b2 - img_cpu).max(axis=-1) > 0).sum())
                del img_bvh
            else:
                cb.update({"value": list_rate, "unit": "Mrays/s", "cores": lib.num_threads(), "kind": cinfo["kind"],
                           "accel": cinfo["accel"], "accel_reason": cinfo.get("accel_reason"), "sample": cb["parity_sample"]})
            line["cpu_baseline"] = cb
        print(json.dumps(line), flush=True)
    # torch tensors that were used on the context's stream must go before the stream does
    barrier()
    del accum, rgb, host_rgb, host_verts, e0, e1
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    if world > 1:
        ctx.comm_destroy()
        dist.destroy_process_group()
    del ext
    ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1 / C2 / C5 sub-records of the N = 1 line")
    ap.add_argument("--lanes", type=int, default=3, help="CRT_OPT_LANES of the timed region (1 for profiling: one slice, launches in bounce order)")
    ap.add_argument("--slice-min", type=int, default=0, help="CRT_OPT_SLICE_MIN (tuning)")
    ap.add_argument("--builder", default="lbvh", choices=["lbvh", "ploc"], help="CRT_OPT_BUILDER (tree topology; images do not depend on it)")
    ap.add_argument("--batch-samples", type=int, default=-1, help="CRT_OPT_BATCH_SAMPLES (A/B runs; default: the library's)")
    ap.add_argument("--packet-defer", type=int, default=-1, help="CRT_OPT_PACKET_DEFER (A/B runs; default: the library's)")
    ap.add_argument("--e2e-api", default="points", choices=["points", "frame", "calls", "parts"],
                    help="end-to-end loop: points = the pose as the mesh's point array (crt_render_frame_pose, one overlapped "
                         "call per frame at N = 1); parts = one translation per instance (--workload c4's default); frame = all triangle vertices (crt_render_frame); calls = vertices through "
                         "the three separate calls")
    ap.add_argument("--workload", default="c3", choices=["c3", "c4"],
                    help="c3 (default): the configuration the metric is quoted on; c4: the 10 M-triangle instanced scene at "
                         "4K x 64 spp (BASELINE.json configs[3], tile partition + NCCL reduce under torchrun)")
    ap.add_argument("--combine", default="gather", choices=["gather", "reduce"],
                    help="N > 1, tile partition: gather = each rank's tiles to rank 0 (default); reduce = ncclReduce of the full frames")
    ap.add_argument("--scaling", default=None, choices=["weak", "strong"],
                    help="N > 1: strong (default) = the configured 16-spp frame split by tiles; weak = every rank renders the "
                         "full 16-spp frame's work on its own sample range")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
