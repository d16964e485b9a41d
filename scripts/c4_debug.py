"""Development aid: the camera rays of the C4 pixels on which the packet and the per-ray traversal disagreed
(profiles/r02_f_c4_invariance.txt), traced one by one through every closest-hit kernel against the brute-force scan."""
import sys

import numpy as np

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
import oracle
from cudaraytracer_b200 import scenes

sc = scenes.c4_instanced()
nx, ny, spp = sc["nx"], sc["ny"], 64
cam = crt.camera_for(sc)
o = oracle.load("ref" if oracle.available("ref") else "port")
pixels = [(591, 2138), (673, 3060), (781, 3295), (1034, 2709), (590, 3046), (771, 3659)]
rays, tags = [], []
for (y, x) in pixels:
    pix = y * nx + x
    for s in range(spp):
        key = scenes.rng_key(7, np.uint32(pix), np.uint32(s))
        r0 = scenes.rng_uniform(key, np.uint32(0))
        r1 = scenes.rng_uniform(key, np.uint32(1))
        u = np.float32(np.float32(np.float32(x) + r0) / np.float32(nx))
        v = np.float32(np.float32(np.float32(y) + r1) / np.float32(ny))
        ray, _ = o.get_ray(cam, float(u), float(v), 7, pix, s, 2)
        rays.append(ray)
        tags.append((y, x, s))
rays = np.ascontiguousarray(np.stack(rays), np.float32)
with crt.Context(0) as c:
    c.set_scene(sc)
    c.build()
    br = c.trace(rays, accel=crt.ACCEL_BRUTE)
    lb = c.trace(rays, accel=crt.ACCEL_LBVH)
    pk = c.trace(rays, accel=crt.ACCEL_LBVH_PACKET)
    # packets as the renderer forms them with 4 samples per batch: 8 neighbouring pixels x 4 samples are not available
    # here, so also one ray per packet (31 copies of a far-away miss ray beside it would widen the interval; use copies)
    solo = np.repeat(rays, 32, axis=0)
    ps = c.trace(solo, accel=crt.ACCEL_LBVH_PACKET)
    ps = tuple(a[::32] for a in ps)
    for name, h in (("lbvh", lb), ("packet(64 samples of a pixel in 2 packets)", pk), ("packet(32 copies)", ps)):
        bad = np.nonzero((h[1] != br[1]) | (h[0] != br[0]))[0]
        print("%s: %d of %d rays differ from brute" % (name, len(bad), len(rays)))
        for i in bad[:10]:
            print("   pixel %s sample %d: id %d t %.9g  vs brute id %d t %.9g   ray o %s d %s" % (
                tags[i][:2], tags[i][2], h[1][i], h[0][i], br[1][i], br[0][i], rays[i, :3], rays[i, 3:]))
    # second bounce: the scattered rays are not reproduced here; report the primary hits of the odd pixels instead
    for (y, x) in pixels[:4]:
        sel = [i for i, t in enumerate(tags) if t[:2] == (y, x)]
        print("pixel", (y, x), "brute ids", sorted(set(br[1][sel].tolist()))[:12])
