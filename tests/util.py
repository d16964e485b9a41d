"""Helpers shared by the tests (TEST INFRASTRUCTURE)."""
import os

import numpy as np

import oracle

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FLT_MAX = 3.4028234663852886e38


def oracle_kinds():
    return [k for k in ("ref", "port") if oracle.available(k)]


def best_oracle():
    return oracle.load("ref" if oracle.available("ref") else "port")


def oracle_scene(lib, sc, d1=True):
    s = lib.scene(d1)
    for m in sc["mats"]:
        s.add_material(*m)
    nt, ns = len(sc["tri"]), len(sc["sph"])
    if nt:
        s.add_triangles(sc["tri"], sc["tri_mat"], np.arange(nt, dtype=np.uint32))
    if ns:
        s.add_spheres(sc["sph"], sc["sph_mat"], np.arange(nt, nt + ns, dtype=np.uint32))
    next_id = nt + ns
    for it in sc.get("chained", []):            # ids continue in list order, as scenes.flatten_chained numbers them
        mat = np.asarray(it["mat"], np.uint32).reshape(-1)
        ids = np.arange(next_id, next_id + len(mat), dtype=np.uint32)
        next_id += len(mat)
        s.add_chained(it["kind"], it.get("data"), mat, ids, it["levels"], it.get("flip"))
    return s.finalize()


def subset_scene(sc, n_tri=None, n_sph=None):
    out = dict(sc)
    if n_tri is not None:
        out["tri"] = np.ascontiguousarray(sc["tri"][:n_tri])
        out["tri_mat"] = np.ascontiguousarray(sc["tri_mat"][:n_tri])
    if n_sph is not None:
        out["sph"] = np.ascontiguousarray(sc["sph"][:n_sph])
        out["sph_mat"] = np.ascontiguousarray(sc["sph_mat"][:n_sph])
    return out
