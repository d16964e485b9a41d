"""Generates the golden vectors under tests/golden/ from the UNMODIFIED
reference headers (oracle/_ref, built from /root/reference). Run in the
authoring container only:  python tests/golden/make_golden.py

  small_bunny_parsed.npz      parseObjByName("objects/small_bunny.obj")  (obj_loader.h:20-111)
  bunny_primary_1024x512.npz  prim id + t of every primary ray of C1 through the reference list scan
  quirk_rays.npz              hand-picked rays for the semantics of SURVEY §9 (Q1-Q3, Q6, Q8)
  scatter_vectors.npz         Material::scatter outputs for fixed random sequences
  c1_lambert_256x128.npz      C1 image, LambertShade (reference default), float32
  c1_path_256x128.npz         C1 image, shade(), float32
  c1_lambert_1024x512_rgba8.npz  C1 at the reference's resolution after WritePng's conversion
  c2_path_150x100.npz         C2 (spheres, 3 materials) at reduced size
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

import oracle  # noqa: E402
import cudaraytracer_b200 as crt  # noqa: E402
from cudaraytracer_b200 import scenes  # noqa: E402
from util import oracle_scene  # noqa: E402


def quirk_scene_and_rays():
    """Two triangles and two spheres; rays that pin the reference's quirks."""
    tri = np.asarray([
        # z=0 plane, stored normal +z: hit only by rays travelling +z (Q1)
        [-1, -1, 0, 1, -1, 0, 0, 1, 0, 0, 0, 1],
        # same place at z=2, normal -z
        [-1, -1, 2, 1, -1, 2, 0, 1, 2, 0, 0, -1],
        # two triangles sharing the edge x=0 at z=5 (Q3 ties), normal +z
        [-1, -1, 5, 0, -1, 5, 0, 1, 5, 0, 0, 1],
        [0, -1, 5, 1, -1, 5, 0, 1, 5, 0, 0, 1],
    ], np.float32)
    sph = np.asarray([[0, 0, -10, 2], [5, 0, 0, -1.0]], np.float32)   # negative radius flips the normal (Q8)
    rays = np.asarray([
        [0, 0, -1, 0, 0, 1],      # +z through tri 0 (hit), tri 1 culled (normal -z), tri 2/3 edge tie
        [0, 0, 3, 0, 0, -1],      # -z: tri 1 hit at t=1; tri 0 culled
        [0, -0.5, 1, 0, 0, 2],    # un-normalised direction: t is still Euclidean (Q5)
        [0, 0, 1, 0, 0, 1],       # tri 0 is BEHIND the origin: raw reference returns t=-1 (Q2), D1 rejects
        [0, -0.5, 4, 0, 0, 1],    # exactly on the shared edge x=0 (tie -> lower id)
        [0, 0, 0, 0, 0, -1],      # sphere 0, near root
        [0, 0, -10, 0, 0, -1],    # from inside sphere 0: far root
        [0, 0, -10, 1, 0, 0],     # from sphere 0 centre towards sphere 1 (negative radius)
        [0.25, -0.25, -1, 0, 0, 1],
        [100, 100, 100, 0, 1, 0], # miss
        [0, 0, 6, 1e-9, 0, -1],   # nearly axis-parallel, goes back through everything
    ], np.float32)
    sc = dict(tri=tri, tri_mat=np.zeros(4, np.uint32), sph=sph, sph_mat=np.zeros(2, np.uint32),
              mats=[(0, (0.5, 0.5, 0.5), 0.0)])
    return sc, rays


def main():
    ref = oracle.load("ref")
    out = {}

    pts, idx = ref.parse_obj("/root/reference/CudaTest/objects/small_bunny.obj")
    np.savez_compressed(os.path.join(HERE, "small_bunny_parsed.npz"), points=pts, idx=idx)

    # objects/cbox_smallbox.obj has "f a//n" faces, which the reference's parseObjByName mangles (it reads each face
    # token with operator>> as a float; its parseFaceIndex, obj_loader.h:113-118, is unused): the points - centred by
    # the reference's own computeMean / centering - come from the reference, the faces are read here field by field
    cpts, _ = ref.parse_obj("/root/reference/CudaTest/objects/cbox_smallbox.obj")
    cf = []
    for line in open("/root/reference/CudaTest/objects/cbox_smallbox.obj"):
        if line.startswith("f "):
            cf.append([float(t.split("/")[0]) for t in line.split()[1:4]])
    np.savez_compressed(os.path.join(HERE, "cbox_smallbox.npz"), points=cpts, idx=np.asarray(cf, np.float32))

    sc = scenes.c1_bunny()
    s = oracle_scene(ref, sc)
    cam = crt.camera_for(sc, 1024, 512)
    rays = scenes.primary_rays(cam, 1024, 512)
    t, pid = s.trace(rays)
    raw = oracle_scene(ref, sc, d1=False)
    _, pid_raw = raw.trace(rays[::16])
    np.savez_compressed(os.path.join(HERE, "bunny_primary_1024x512.npz"), t=t, id=pid,
                        d1_changed=np.int64((pid_raw != pid[::16]).sum()))
    print("bunny table:", int((pid != 0xFFFFFFFF).sum()), "hits; D1 changes", int((pid_raw != pid[::16]).sum()))

    qs, qr = quirk_scene_and_rays()
    for d1 in (1, 0):
        qsc = oracle_scene(ref, qs, d1=bool(d1))
        qt, qid, qp, qn = qsc.trace(qr, want_pn=True)
        out["quirk_t_d1_%d" % d1], out["quirk_id_d1_%d" % d1] = qt, qid
        out["quirk_p_d1_%d" % d1], out["quirk_n_d1_%d" % d1] = qp, qn
    np.savez_compressed(os.path.join(HERE, "quirk_rays.npz"), tri=qs["tri"], sph=qs["sph"], rays=qr, **out)
    print("quirks d1:", out["quirk_id_d1_1"], out["quirk_t_d1_1"])
    print("quirks raw:", out["quirk_id_d1_0"], out["quirk_t_d1_0"])

    rng = np.random.default_rng(42)
    n = 600
    recs = dict(type=np.zeros(n, np.uint32), albedo=np.zeros((n, 3), np.float32), param=np.zeros(n, np.float32),
                ray=np.zeros((n, 6), np.float32), p=np.zeros((n, 3), np.float32), normal=np.zeros((n, 3), np.float32),
                randoms=np.zeros((n, 48), np.float32), ok=np.zeros(n, np.uint8), att=np.zeros((n, 3), np.float32),
                org=np.zeros((n, 3), np.float32), dir=np.zeros((n, 3), np.float32), draws=np.zeros(n, np.int32))
    for i in range(n):
        mt = i % 3
        recs["type"][i] = mt
        recs["albedo"][i] = rng.random(3)
        recs["param"][i] = [0.0, rng.random() * 1.4, rng.uniform(1.05, 2.4)][mt]
        recs["ray"][i] = np.concatenate([rng.normal(size=3), rng.normal(size=3) * 3])
        nn = rng.normal(size=3)
        recs["normal"][i] = nn / np.linalg.norm(nn)
        recs["p"][i] = rng.normal(size=3) * 10
        recs["randoms"][i] = 1.0 - rng.random(48)
        ok, att, org, dr, nd = ref.scatter(mt, recs["albedo"][i], recs["param"][i], recs["ray"][i], recs["p"][i],
                                           recs["normal"][i], recs["randoms"][i])
        recs["ok"][i], recs["att"][i], recs["org"][i], recs["dir"][i], recs["draws"][i] = ok, att, org, dr, nd
    np.savez_compressed(os.path.join(HERE, "scatter_vectors.npz"), **recs)

    cam_s = crt.camera_for(sc, 256, 128)
    for name, mode in (("lambert", 1), ("path", 0)):
        img = s.render(cam_s, 256, 128, 4, 8, mode, 1)
        np.savez_compressed(os.path.join(HERE, "c1_%s_256x128.npz" % name), rgb=img, seed=1, spp=4, max_depth=8)
        print("c1", name, "rays", s.last_rays, "%.1fs" % s.last_seconds)
    img = s.render(cam, 1024, 512, 4, 8, 1, 1)
    np.savez_compressed(os.path.join(HERE, "c1_lambert_1024x512_rgba8.npz"), rgba=ref.pack_rgba8(img), seed=1, spp=4)
    print("c1 full lambert %.1fs" % s.last_seconds)

    sc2 = scenes.c2_spheres()
    s2 = oracle_scene(ref, sc2)
    cam2 = crt.camera_for(sc2, 150, 100)
    img2 = s2.render(cam2, 150, 100, 4, 8, 0, 2)
    np.savez_compressed(os.path.join(HERE, "c2_path_150x100.npz"), rgb=img2, seed=2, spp=4, max_depth=8)
    print("c2 rays", s2.last_rays, "%.1fs" % s2.last_seconds)


if __name__ == "__main__":
    main()
