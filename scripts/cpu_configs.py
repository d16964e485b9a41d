"""CPU reference (oracle/_ref: the unmodified reference headers, list scan, OpenMP over rows) on a bounded sub-frame of
C1, C2 and C3, for the BASELINE.md table. Runs without a GPU: `python scripts/cpu_configs.py [budget_s]`."""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402
import cudaraytracer_b200 as crt  # noqa: E402
from cudaraytracer_b200 import scenes  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
for name, sc, mode in (("C1 bunny LambertShade", scenes.c1_bunny(), crt.SHADE_LAMBERT), ("C1 bunny shade()", scenes.c1_bunny(), crt.SHADE_PATH),
                       ("C2 spheres", scenes.c2_spheres(), crt.SHADE_PATH), ("C3 1M-tri mesh", scenes.c3_mesh(), crt.SHADE_PATH)):
    lib, sh, kind = bench.cpu_reference_setup(sc)
    cam_fn = lambda a, b: crt.camera_for(sc, a, b)
    nx, ny = bench.cpu_sample_shape(sh, cam_fn, sc, budget)
    sh.render(cam_fn(nx, ny), nx, ny, 1, sc["max_depth"], mode, 1)
    print(json.dumps(dict(config=name, kind=kind, cores=lib.num_threads(), sample="%dx%d px x 1 spp" % (nx, ny),
                          seconds=sh.last_seconds, rays=sh.last_rays, mrays_s=sh.last_rays / sh.last_seconds / 1e6)), flush=True)
