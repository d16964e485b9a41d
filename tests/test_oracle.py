"""CPU tier: pins the oracle.

* the C restatement (oracle/crt_oracle.c) against the golden vectors that were
  generated from the UNMODIFIED reference headers (tests/golden/make_golden.py);
* when oracle/_ref is present (authoring container, or the prebuilt .so on the
  GPU box), the restatement against the reference itself on fresh inputs;
* the reference's quirks of SURVEY.md §9 as explicit known-answer tests.
The reference ships no tests or golden vectors of its own, so these
files are the pin.
"""
import os

import numpy as np
import pytest

import oracle
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes
from util import GOLDEN, oracle_kinds, oracle_scene

KINDS = oracle_kinds()


def _scene_from_arrays(lib, tri, sph, d1=True):
    sc = dict(tri=tri, tri_mat=np.zeros(len(tri), np.uint32), sph=sph, sph_mat=np.zeros(len(sph), np.uint32),
              mats=[(0, (0.5, 0.5, 0.5), 0.0)])
    return oracle_scene(lib, sc, d1)


@pytest.mark.parametrize("kind", KINDS)
def test_obj_loader_matches_golden(kind, tmp_path):
    """parseObjByName semantics (obj_loader.h:20-133) on a hand-written file, and
    the bundled bunny when the reference tree is mounted."""
    lib = oracle.load(kind)
    p = tmp_path / "t.obj"
    p.write_text("# c\nv 1 2 3\nv 3 2 1\nvn 0 1 0\nv -1 0 5\nvt 0 0\nf 1 2 3\nf 3 2 1\n")
    pts, idx = lib.parse_obj(str(p))
    assert pts.shape == (3, 3) and idx.shape == (2, 3)
    mean = (np.float32(1) + np.float32(3) + np.float32(-1)) / np.float32(3)
    assert pts[0, 0] == np.float32(1) - mean
    assert np.array_equal(idx, np.asarray([[1, 2, 3], [3, 2, 1]], np.float32))      # 1-based floats (Q14)
    bunny = "/root/reference/CudaTest/objects/small_bunny.obj"
    if os.path.exists(bunny):
        g = np.load(os.path.join(GOLDEN, "small_bunny_parsed.npz"))
        pts, idx = lib.parse_obj(bunny)
        assert np.array_equal(pts, g["points"]) and np.array_equal(idx, g["idx"])


def test_bunny_fixture_shape():
    pts, idx = scenes.load_bunny_arrays()
    assert pts.shape == (2503, 3) and idx.shape == (4968, 3)
    assert tuple(idx[0]) == (1069.0, 1647.0, 1578.0)


@pytest.mark.parametrize("kind", KINDS)
def test_bunny_hit_table_subset(kind):
    """Every 23rd primary ray of the 1024x512 golden table: ids and t bit-exact."""
    lib = oracle.load(kind)
    g = np.load(os.path.join(GOLDEN, "bunny_primary_1024x512.npz"))
    sc = scenes.c1_bunny()
    rays = scenes.primary_rays(crt.camera_for(sc, 1024, 512), 1024, 512)
    sel = np.arange(0, len(rays), 23)
    hit = np.nonzero(g["id"] != crt.MISS)[0][::7]
    sel = np.unique(np.concatenate([sel, hit]))
    t, pid = oracle_scene(lib, sc).trace(rays[sel])
    assert np.array_equal(pid, g["id"][sel])
    assert np.array_equal(t, g["t"][sel])
    assert int(g["d1_changed"]) == 0          # D1 changes no primary ray of a camera outside the mesh


@pytest.mark.parametrize("kind", KINDS)
def test_quirk_rays(kind):
    lib = oracle.load(kind)
    g = np.load(os.path.join(GOLDEN, "quirk_rays.npz"))
    for d1 in (1, 0):
        s = _scene_from_arrays(lib, g["tri"], g["sph"], d1=bool(d1))
        t, pid, p, n = s.trace(g["rays"], want_pn=True)
        assert np.array_equal(pid, g["quirk_id_d1_%d" % d1])
        assert np.array_equal(t, g["quirk_t_d1_%d" % d1])
        hit = pid != crt.MISS
        assert np.array_equal(p[hit], g["quirk_p_d1_%d" % d1][hit])
        assert np.array_equal(n[hit], g["quirk_n_d1_%d" % d1][hit])


def test_quirks_known_answers():
    """The behaviours themselves, spelled out."""
    g = np.load(os.path.join(GOLDEN, "quirk_rays.npz"))
    idd, t = g["quirk_id_d1_1"], g["quirk_t_d1_1"]
    raw_id, raw_t = g["quirk_id_d1_0"], g["quirk_t_d1_0"]
    # Q1: ray 0 travels +z: triangle 0 (normal +z) is hit, triangle 1 (normal -z) is culled
    assert idd[0] == 0 and t[0] == 1.0
    # Q1 mirrored: ray 1 travels -z: triangle 1 is hit, triangle 0 culled
    assert idd[1] == 1 and t[1] == 1.0
    # Q5: direction (0,0,2): t is Euclidean distance, not the ray parameter
    assert t[2] == 4.0
    # Q2: the raw reference returns the triangle BEHIND the origin (t = -1); D1 rejects it
    assert raw_id[3] == 0 and raw_t[3] == -1.0 and idd[3] == 2 and t[3] == 4.0
    # Q3: a ray exactly on the shared edge of triangles 2 and 3 hits both; the lower id wins
    assert idd[4] == 2
    # Q8/Q6: sphere near root; far root from inside; hit point reported in the sphere's frame
    assert idd[5] == 4 and t[5] == 8.0 and idd[6] == 4 and t[6] == 2.0
    assert np.array_equal(g["quirk_p_d1_1"][5], np.asarray([0, 0, 2], np.float32))
    assert idd[9] == crt.MISS


@pytest.mark.parametrize("kind", KINDS)
def test_scatter_vectors(kind):
    lib = oracle.load(kind)
    g = np.load(os.path.join(GOLDEN, "scatter_vectors.npz"))
    for i in range(len(g["type"])):
        ok, att, org, dr, nd = lib.scatter(int(g["type"][i]), g["albedo"][i], float(g["param"][i]), g["ray"][i],
                                           g["p"][i], g["normal"][i], g["randoms"][i])
        assert ok == bool(g["ok"][i]) and nd == int(g["draws"][i]), i
        assert np.array_equal(att, g["att"][i]) and np.array_equal(org, g["org"][i]), i
        assert np.array_equal(dr, g["dir"][i], equal_nan=True), i


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("name,mode", [("lambert", 1), ("path", 0)])
def test_c1_golden_image_rows(kind, name, mode):
    """Rows of the C1 golden images (reference default LambertShade and shade())."""
    lib = oracle.load(kind)
    g = np.load(os.path.join(GOLDEN, "c1_%s_256x128.npz" % name))
    sc = scenes.c1_bunny()
    s = oracle_scene(lib, sc)
    rows = (40, 64, 77)
    for y in rows:
        img = s.render(crt.camera_for(sc, 256, 128), 256, 128, int(g["spp"]), int(g["max_depth"]), mode, int(g["seed"]),
                       y0=y, y1=y + 1)
        assert np.array_equal(img[y], g["rgb"][y])


@pytest.mark.parametrize("kind", KINDS)
def test_c2_golden_image_rows(kind):
    lib = oracle.load(kind)
    g = np.load(os.path.join(GOLDEN, "c2_path_150x100.npz"))
    sc = scenes.c2_spheres()
    s = oracle_scene(lib, sc)
    img = s.render(crt.camera_for(sc, 150, 100), 150, 100, int(g["spp"]), int(g["max_depth"]), 0, int(g["seed"]),
                   y0=20, y1=60, y_step=8)
    for y in range(20, 60, 8):
        assert np.array_equal(img[y], g["rgb"][y])


def test_pack_rgba8_golden():
    lib = oracle.load(KINDS[-1])
    g = np.load(os.path.join(GOLDEN, "c1_lambert_1024x512_rgba8.npz"))
    rgba = g["rgba"]
    assert rgba.shape == (512, 1024, 4) and (rgba[..., 3] == 255).all()
    x = np.asarray([[[0.0, 1.0, 0.5]], [[0.999, 0.2, 0.00391]]], np.float32)
    out = lib.pack_rgba8(x)
    assert out[0, 0].tolist() == [255, 51, 1, 255]        # rows flipped: row 0 of the PNG is the top
    assert out[1, 0].tolist() == [0, 255, 127, 255]       # char(255.99 * 1.0) = 255


@pytest.mark.skipif(not oracle.available("ref"), reason="oracle/_ref not built (no /root/reference)")
def test_port_matches_reference_on_fresh_inputs():
    ref, port = oracle.load("ref"), oracle.load("port")
    rng = np.random.default_rng(123)
    # cameras
    for _ in range(200):
        a = (rng.normal(size=3) * 30, rng.normal(size=3) * 3, (0, 1, 0), rng.uniform(10, 100), rng.uniform(0.5, 2.5),
             rng.uniform(0, 0.5), rng.uniform(1, 30))
        c1, c2 = ref.camera(*a), port.camera(*a)
        assert np.array_equal(c1, c2)
        assert np.array_equal(c1, crt.camera22(*a))
        r1, r2 = ref.get_ray(c1, 0.3, 0.7, 5, 11, 2), port.get_ray(c1, 0.3, 0.7, 5, 11, 2)
        assert np.array_equal(r1[0], r2[0]) and r1[1] == r2[1]
    # mixed scene, random rays, with and without D1
    tri = scenes.c5_triangles(300, seed=11)["tri"]
    sph = np.concatenate([rng.uniform(0, 1, (40, 3)), rng.uniform(0.02, 0.15, (40, 1))], axis=1).astype(np.float32)
    rays = scenes.c5_incoherent_rays(3000, seed=12)
    for d1 in (True, False):
        a = _scene_from_arrays(ref, tri, sph, d1).trace(rays, want_pn=True)
        b = _scene_from_arrays(port, tri, sph, d1).trace(rays, want_pn=True)
        hit = a[1] != crt.MISS
        assert np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0])
        assert np.array_equal(a[2][hit], b[2][hit]) and np.array_equal(a[3][hit], b[3][hit])
    # renders, all three shaders
    sc = scenes.c2_spheres()
    cam = crt.camera_for(sc, 60, 40)
    sa, sb = oracle_scene(ref, sc), oracle_scene(port, sc)
    for mode in (0, 1, 2):
        assert np.array_equal(sa.render(cam, 60, 40, 2, 8, mode, 9), sb.render(cam, 60, 40, 2, 8, mode, 9), equal_nan=True)
    # CheckerTexture: which colour a point selects, then renders through checkered Lambertians and a checkered light
    pts = np.concatenate([rng.normal(size=(20000, 3)) * 3, rng.normal(size=(20000, 3)) * 1000, np.zeros((1, 3)),
                          np.asarray([[np.pi / 10, 1, 1], [1, -np.pi / 5, 2]])]).astype(np.float32)
    odd = ref.checker_odd(pts)
    assert np.array_equal(odd, port.checker_odd(pts)) and 0.4 < odd.mean() < 0.6 and not odd[40000]
    sc = scenes.checker_scene()
    cam = crt.camera_for(sc, 60, 40)
    sa, sb = oracle_scene(ref, sc), oracle_scene(port, sc)
    for mode in (0, 1):
        ia = sa.render(cam, 60, 40, 2, 8, mode, 9)
        assert np.array_equal(ia, sb.render(cam, 60, 40, 2, 8, mode, 9), equal_nan=True)
    plain = dict(sc)
    plain["mats"] = [m[:3] for m in sc["mats"]]
    assert not np.array_equal(oracle_scene(ref, plain).render(cam, 60, 40, 2, 8, 0, 9), sa.render(cam, 60, 40, 2, 8, 0, 9))
    # triangle normals of the scene generator == Triangle constructor
    t = scenes.c1_bunny()["tri"]
    for k in range(0, len(t), 211):
        assert np.array_equal(ref.triangle_normal(t[k, :9]), t[k, 9:12])


@pytest.mark.skipif(not oracle.available("ref"), reason="oracle/_ref not built")
def test_reference_bvh_agrees_with_its_list_scan():
    """The reference's own BVHNode (bvh.h) against its list scan under D1 on
    primary rays: same ids (ties aside), t within 1e-5."""
    ref = oracle.load("ref")
    sc = scenes.c1_bunny()
    sc["tri"], sc["tri_mat"] = sc["tri"][:800], sc["tri_mat"][:800]
    s = oracle_scene(ref, sc)
    s.bvh_build(seed=3)
    rays = scenes.primary_rays(crt.camera_for(sc, 96, 48), 96, 48)
    t0, i0 = s.trace(rays)
    t1, i1 = s.trace(rays, use_bvh=True)
    same = i0 == i1
    assert same.mean() > 0.995
    hit = same & (i0 != crt.MISS)
    assert np.allclose(t0[hit], t1[hit], rtol=1e-5)


@pytest.mark.skipif(not oracle.available("ref"), reason="oracle/_ref not built")
def test_reference_bvh_misses_flat_boxes():
    """Why the parity oracle is the reference's LIST SCAN and not its BVHNode: AABB::hit (core/aabb.h) rejects a box
    whose slab interval is empty, and an axis-aligned quad has a box of zero thickness - rays that the list scan
    reports on C3's ground quad are reported as misses by the reference's own BVH (bench.py's cpu_baseline therefore
    times the BVH but compares images against the list scan)."""
    ref = oracle.load("ref")
    g0, g1, g2, gn = scenes.ground_quad(-1.0, 50.0)
    tilted = np.asarray([[-1, 0, -1, 1, 0.5, -1, 0, 0.2, 1]], np.float32)       # a triangle with a box of real thickness
    tn = ref.triangle_normal(tilted[0])
    quad = np.concatenate([g0, g1, g2, gn], axis=1).astype(np.float32)
    tri = np.concatenate([quad, np.concatenate([tilted, tn[None, :]], axis=1)]).astype(np.float32)
    sc = dict(tri=tri, tri_mat=np.zeros(len(tri), np.uint32), sph=np.zeros((0, 4), np.float32), sph_mat=np.zeros(0, np.uint32),
              mats=[(scenes.LAMBERTIAN, (0.5, 0.5, 0.5), 0.0)])
    s = oracle_scene(ref, sc)
    s.bvh_build(seed=1)
    rng = np.random.default_rng(4)
    o = np.concatenate([rng.uniform(-20, 20, (400, 1)), np.full((400, 1), 5.0), rng.uniform(-20, 20, (400, 1))], axis=1)
    d = np.concatenate([rng.uniform(-0.3, 0.3, (400, 1)), np.full((400, 1), -1.0), rng.uniform(-0.3, 0.3, (400, 1))], axis=1)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    t0, i0 = s.trace(rays)
    t1, i1 = s.trace(rays, use_bvh=True)
    on_quad = (i0 == 0) | (i0 == 1)
    assert on_quad.sum() > 300, "the list scan finds the ground"
    assert (i1[on_quad] == crt.MISS).all(), "the reference's BVH never reports a primitive whose box is flat"
    on_tilted = i0 == 2
    assert on_tilted.any() and np.array_equal(i1[on_tilted], i0[on_tilted])
