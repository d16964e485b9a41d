"""Development aid: for rays on which the LBVH misses a hit the brute-force scan finds, walk the 4-wide tree on the host
from the root towards the leaf that holds the missed triangle and report where the chain breaks."""
import sys

import numpy as np

sys.path.insert(0, ".")
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes

sc = scenes.c4_instanced()
rays = np.asarray([[0, 900, 2600, 3.8421068, -4.565735, -9.001709],
                   [0, 900, 2600, 0.73611736, -4.8258667, -8.911621],
                   [0, 900, 2600, 2.6622705, -3.414795, -9.4001465]], np.float32)
f32 = np.float32
with crt.Context(0) as c:
    c.set_scene(sc)
    c.build()
    n = c.accel_info()["n_prims"]
    br = c.trace(rays, accel=crt.ACCEL_BRUTE)
    lb = c.trace(rays, accel=crt.ACCEL_LBVH)
    print("brute", br[1], br[0], "lbvh", lb[1], lb[0])
    nodes, root = c.debug_read(crt.DBG_NODES, np.uint32, 32 * (n - 1))
    rng_, _ = c.debug_read(crt.DBG_RANGE, np.int32, 2 * (n - 1))
    recs, _ = c.debug_read(crt.DBG_RECS, np.uint32, 16 * n)
    lbox, _ = c.debug_read(crt.DBG_LEAF_BOX, np.float32, 8 * n)
nodes = nodes.reshape(-1, 32)
rng_ = rng_.reshape(-1, 2)
ids = recs.reshape(-1, 16)[:, 12]
lbox = lbox.reshape(-1, 8)


def half(h):
    return np.asarray([h], np.uint16).view(np.float16).astype(np.float32)[0]


for ri, ray in enumerate(rays):
    target = int(br[1][ri])
    pos = int(np.nonzero(ids == target)[0][0])
    o = ray[:3].astype(np.float64)
    d = ray[3:].astype(np.float64)
    d = (d / np.linalg.norm(d))
    inv = 1.0 / d
    print("\nray %d: brute prim %d at sorted position %d, t %.6f; its leaf box lo %s hi %s" % (ri, target, pos, br[0][ri], lbox[pos, 0:3], lbox[pos, 4:7]))
    tl = (lbox[pos, 0:3].astype(np.float64) - o) * inv
    th = (lbox[pos, 4:7].astype(np.float64) - o) * inv
    print("   leaf box slab interval [%.6f, %.6f]" % (np.minimum(tl, th).max(), np.maximum(tl, th).min()))
    ref = root
    level = 0
    while ref >= 0:
        w = nodes[ref]
        found = None
        for k in range(4):
            r = int(w[20 + k])
            if r == 0x80000000:
                continue
            r = r - (1 << 32) if r >= (1 << 31) else r
            if r < 0:
                lr = (~r) & 0xFFFFFFFF
                a, b = lr >> 3, (lr >> 3) + (lr & 7)
            else:
                a, b = rng_[r]
            if a <= pos <= b:
                found = (k, r, a, b)
        if found is None:
            print("   level %d node %d: NO slot covers position %d; slots: %s" % (level, ref, pos, [hex(int(x)) for x in w[20:24]]))
            break
        k, r, a, b = found
        lo = w[5 * k:5 * k + 3].view(np.float32).astype(np.float64)
        e_xy, e_z = int(w[5 * k + 3]), int(w[5 * k + 4])
        ext = np.asarray([half(e_xy & 0xFFFF), half(e_xy >> 16), half(e_z & 0xFFFF)], np.float64)
        hi = (lo.astype(np.float32) + ext.astype(np.float32)).astype(np.float64)
        t0, t1 = (lo - o) * inv, (hi - o) * inv
        tn, tf = np.minimum(t0, t1).max(), np.maximum(t0, t1).min()
        cone = [half(int(w[4]) >> 16), half(int(w[9]) >> 16), half(int(w[14]) >> 16)]
        inside = (lbox[pos, 0:3] >= lo.astype(np.float32)).all() and (lbox[pos, 4:7] <= hi.astype(np.float32)).all()
        print("   level %2d node %8d slot %d -> %9d range [%d, %d]: box lo %s hi %s  interval [%.4f, %.4f] %s  leaf box inside: %s; node cone dot %.4f" % (
            level, ref, k, r, a, b, np.round(lo, 3), np.round(hi, 3), tn, tf, "ENTERED" if tn <= tf else "*** MISSED ***", inside,
            float(np.dot(d, cone))))
        ref = r
        level += 1
