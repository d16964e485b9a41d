"""Synthetic scenes and ray sets of BASELINE.json's configs.

Everything is generated on the host from a 64-bit seed with the integer hash of
``include/crt_rng.h`` (re-stated in numpy below), so the same arrays come out on
every machine. A scene is a plain dict of numpy arrays in the layout of
``include/crt.h``:

    tri      (n, 12) float32  vertices[0..2] + stored normal   (shapes/triangle.h:49-50)
    tri_mat  (n,)    uint32
    sph      (m, 4)  float32  centre + radius                  (shapes/sphere.h:22)
    sph_mat  (m,)    uint32
    mats     list of (type, (r, g, b), param)                  (material/material.h:57-143)
    camera   dict(lookfrom, lookat, vup, vfov, aperture, focus_dist)
    nx, ny, spp, max_depth

Reference paths are relative to /root/reference/CudaTest/src/.
"""
import os

import numpy as np

LAMBERTIAN, METAL, DIELECTRIC, DIFFUSE_LIGHT = 0, 1, 2, 3

_GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


# ----------------------------------------------------------------- hashing
def mix32(x):
    """include/crt_rng.h crt_mix32 on a uint32 array."""
    x = np.asarray(x, dtype=np.uint32).copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint32(16)
        x *= np.uint32(0x7FEB352D)
        x ^= x >> np.uint32(15)
        x *= np.uint32(0x846CA68B)
        x ^= x >> np.uint32(16)
    return x


def rng_key(seed, pixel, sample):
    """include/crt_rng.h crt_rng_key."""
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    with np.errstate(over="ignore"):
        a = mix32(np.uint32((seed & 0xFFFFFFFF) ^ 0x9E3779B9))
        b = mix32(np.uint32(seed >> 32) + a)
        k0 = mix32(np.asarray(pixel, np.uint32) ^ a)
        k1 = mix32(np.asarray(sample, np.uint32) ^ b) + k0
    return (k1.astype(np.uint64) << np.uint64(32)) | k0.astype(np.uint64)


def rng_uniform(key, ctr):
    """include/crt_rng.h crt_rng_uniform: float32 in (0, 1]."""
    with np.errstate(over="ignore"):
        c = np.asarray(ctr, np.uint32) * np.uint32(0x9E3779B1) + np.uint32(0x7F4A7C15)
        key = np.asarray(key, np.uint64)
        k0, k1 = (key & np.uint64(0xFFFFFFFF)).astype(np.uint32), (key >> np.uint64(32)).astype(np.uint32)
        bits = mix32(k0 ^ mix32(k1 ^ c))
    return ((bits >> np.uint32(8)) + np.uint32(1)).astype(np.float32) * np.float32(1.0 / 16777216.0)


_libm = None


def _sincosf(x):
    """libm's sinf / cosf of a float32, as the host-compiled reference calls them (core/vec3.h:202-209)."""
    global _libm
    import ctypes as C
    if _libm is None:
        _libm = C.CDLL("libm.so.6")
        for f in (_libm.sinf, _libm.cosf):
            f.restype = C.c_float
            f.argtypes = [C.c_float]
    return np.float32(_libm.sinf(float(x))), np.float32(_libm.cosf(float(x)))


def rotation_rows(rot):
    """rotate()'s matrix (core/vec3.h:200-214) for Euler angles in degrees: rows rotate0, rotate1, rotate2 as nine
    float32 values, every product and sum rounded to float32 in the reference's order (the Z angle is negated)."""
    f = np.float32
    rx = f(np.float64(np.pi / 180.0) * np.float64(f(rot[0])))
    ry = f(np.float64(np.pi / 180.0) * np.float64(f(rot[1])))
    rz = f(np.float64(-(np.pi / 180.0)) * np.float64(f(rot[2])))
    (sx, cx), (sy, cy), (sz, cz) = _sincosf(rx), _sincosf(ry), _sincosf(rz)
    r0 = [cy * cz, (-cy) * sz, sy]
    r1 = [f(f(sx * sy) * cz) + f(cx * sz), f(f((-sx) * sy) * sz) + f(cx * cz), (-sx) * cy]
    r2 = [f(f((-cx) * sy) * cz) + f(sx * sz), f(f(cx * sy) * sz) + f(sx * cz), cx * cy]
    return np.asarray(r0 + r1 + r2, np.float32)


IDENTITY_ROWS = np.asarray([1, 0, 0, 0, 1, 0, 0, 0, 1], np.float32)


def rotate_point(v, rot):
    """rotate(v, rot) of core/vec3.h:200-217 in float32: component k = (v * row_k).sum(), summed left to right."""
    f = np.float32
    m = rotation_rows(rot)
    v = np.asarray(v, f)
    return np.asarray([f(f(f(v[0] * m[3 * k]) + f(v[1] * m[3 * k + 1])) + f(v[2] * m[3 * k + 2])) for k in range(3)], f)


def flatten_chained(scene):
    """The arrays crt_scene_set_frames / _triangle_chains / _sphere_chains / _rectangles take for scene["chained"]:
    a list of items dict(kind, data, mat, levels[, flip]) - `levels` are the (position, rotation) pairs of the
    Transforms between the world list and the item's primitives, outermost first, the last the primitives' own
    (oracle.Scene.add_chained takes the same). On the device a chain also holds the world list's own identity
    level, because every level re-normalises the direction (crt_device.cuh:chain_ray). Primitive ids: the plain
    triangles, the plain spheres, then the items in list order (the order of the reference's world list).
    Returns dict(frames (n, 13), tri, tri_mat, tri_id, tri_chain, sph, sph_mat, sph_id, sph_chain,
    rect_chain, rect_flip, rect_mat, rect_id)."""
    frames = []
    n_t, n_s = len(scene["tri"]), len(scene["sph"])
    tri, tri_mat, tri_chain = [scene["tri"]], [scene["tri_mat"]], [np.zeros(n_t, np.uint32)]
    sph, sph_mat, sph_chain = [scene["sph"]], [scene["sph_mat"]], [np.zeros(n_s, np.uint32)]
    next_id = n_t + n_s
    tri_id = [np.arange(n_t, dtype=np.uint32)]
    sph_id = [np.arange(n_t, next_id, dtype=np.uint32)]
    rect = dict(chain=[], flip=[], mat=[], id=[])
    for it in scene.get("chained", []):
        first = len(frames)
        frames.append(np.concatenate([IDENTITY_ROWS, np.zeros(3, np.float32), [0]]))
        for pos, rot in it["levels"]:
            has_rot = any(float(a) != 0.0 for a in rot)
            rows = rotation_rows(rot) if has_rot else IDENTITY_ROWS
            frames.append(np.concatenate([rows, np.asarray(pos, np.float32), [1 if has_rot else 0]]))
        chain = np.uint32((first << 8) | (len(it["levels"]) + 1))
        mat = np.asarray(it["mat"], np.uint32).reshape(-1)
        n = len(mat)
        ids = np.arange(next_id, next_id + n, dtype=np.uint32)
        next_id += n
        if it["kind"] == 0:
            tri.append(np.asarray(it["data"], np.float32).reshape(-1, 12)); tri_mat.append(mat); tri_id.append(ids)
            tri_chain.append(np.full(n, chain, np.uint32))
        elif it["kind"] == 1:
            r = np.asarray(it["data"], np.float32).reshape(-1, 1)
            sph.append(np.concatenate([np.zeros((n, 3), np.float32), r], axis=1)); sph_mat.append(mat); sph_id.append(ids)
            sph_chain.append(np.full(n, chain, np.uint32))
        else:
            rect["chain"].append(np.full(n, chain, np.uint32)); rect["mat"].append(mat); rect["id"].append(ids)
            rect["flip"].append(np.asarray(it.get("flip", np.zeros(n)), np.uint8).reshape(-1))

    def cat(parts, dt):
        return np.ascontiguousarray(np.concatenate(parts).astype(dt)) if parts else np.zeros((0,), dt)

    return dict(frames=np.ascontiguousarray(np.stack(frames).astype(np.float32)) if frames else np.zeros((0, 13), np.float32),
                tri=cat(tri, np.float32).reshape(-1, 12), tri_mat=cat(tri_mat, np.uint32), tri_id=cat(tri_id, np.uint32),
                tri_chain=cat(tri_chain, np.uint32), sph=cat(sph, np.float32).reshape(-1, 4), sph_mat=cat(sph_mat, np.uint32),
                sph_id=cat(sph_id, np.uint32), sph_chain=cat(sph_chain, np.uint32), rect_chain=cat(rect["chain"], np.uint32),
                rect_flip=cat(rect["flip"], np.uint8), rect_mat=cat(rect["mat"], np.uint32), rect_id=cat(rect["id"], np.uint32))


class Stream:
    """Sequential (0, 1] draws of one stream, for host-side scene generation."""

    def __init__(self, seed, stream=0):
        self.key = rng_key(seed, np.uint32(stream), np.uint32(0xC0FFEE))
        self.ctr = 0

    def draw(self, n=None):
        if n is None:
            v = rng_uniform(self.key, np.uint32(self.ctr))
            self.ctr += 1
            return float(v)
        v = rng_uniform(self.key, np.arange(self.ctr, self.ctr + n, dtype=np.uint32))
        self.ctr += n
        return v


# ------------------------------------------------------------ mesh helpers
def tri_normals(v0, v1, v2):
    """Triangle constructor normal, unit(cross(v1-v0, v2-v0)) in float32 with
    the reference's operation order (shapes/triangle.h:14-17, core/vec3.h:113-117,38-40,191-193)."""
    v0 = np.asarray(v0, np.float32)
    e1 = np.asarray(v1, np.float32) - v0
    e2 = np.asarray(v2, np.float32) - v0
    cx = e1[:, 1] * e2[:, 2] - e1[:, 2] * e2[:, 1]
    cy = -(e1[:, 0] * e2[:, 2] - e1[:, 2] * e2[:, 0])
    cz = e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]
    ln = np.sqrt(cx * cx + cy * cy + cz * cz)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.stack([cx / ln, cy / ln, cz / ln], axis=1).astype(np.float32)


def pack_tris(v0, v1, v2, normals=None, flip_normals=False):
    v0, v1, v2 = (np.asarray(a, np.float32) for a in (v0, v1, v2))
    n = tri_normals(v0, v1, v2) if normals is None else np.asarray(normals, np.float32)
    if flip_normals:
        n = -n
    return np.ascontiguousarray(np.concatenate([v0, v1, v2, n], axis=1), dtype=np.float32)


def mesh_from_indexed(points, idx_1based_float, scale=1.0):
    """add_mesh_withNormal's triangle assembly (createScene.h:175-190) for an
    OBJ mesh: indices are the loader's 1-based floats (obj_loader.h:62-64,
    SURVEY §9 Q14), winding reversed to idx[2], idx[1], idx[0]
    (createScene.h:183); normals from the Triangle constructor."""
    pts = np.asarray(points, np.float32) * np.float32(scale)
    idx = np.asarray(idx_1based_float).astype(np.int64) - 1
    v0, v1, v2 = pts[idx[:, 2]], pts[idx[:, 1]], pts[idx[:, 0]]
    return pack_tris(v0, v1, v2)


def load_bunny_arrays():
    """points (2503,3) mean-centred + 1-based float indices (4968,3) exactly as
    the reference's parseObjByName returns them for objects/small_bunny.obj;
    stored as a fixture because /root/reference does not travel."""
    z = np.load(os.path.join(_GOLDEN, "small_bunny_parsed.npz"))
    return z["points"], z["idx"]


def camera_dict(lookfrom, lookat, vfov, aperture=0.0, focus_dist=10.0, vup=(0, 1, 0)):
    return dict(lookfrom=tuple(map(float, lookfrom)), lookat=tuple(map(float, lookat)),
                vup=tuple(map(float, vup)), vfov=float(vfov), aperture=float(aperture),
                focus_dist=float(focus_dist))


def _empty_tris():
    return np.zeros((0, 12), np.float32), np.zeros((0,), np.uint32)


def _empty_sph():
    return np.zeros((0, 4), np.float32), np.zeros((0,), np.uint32)


# ------------------------------------------------------------------ configs
def c1_bunny(camera="framed", nx=1024, ny=512, spp=4):
    """C1: bundled small_bunny.obj, points x100 (obj_loader.h:164), one
    Lambertian(0.65, 0.05, 0.05) (createScene.h:180); 1024x512, 4 spp, depth 8
    (kernel.cu:44-49). camera="reference" uses the literals of
    createScene.h:160, which do not frame a +-10 unit mesh; "framed" looks at it."""
    pts, idx = load_bunny_arrays()
    tri = mesh_from_indexed(pts, idx, scale=100.0)
    sph, sph_mat = _empty_sph()
    cam = camera_dict((0, 100, 1000), (0, 150, 0), 40) if camera == "reference" else camera_dict((0, 0, 30), (0, 0, 0), 40)
    return dict(name="C1_bunny", tri=tri, tri_mat=np.zeros(len(tri), np.uint32), sph=sph, sph_mat=sph_mat,
                mats=[(LAMBERTIAN, (0.65, 0.05, 0.05), 0.0)], camera=cam, nx=nx, ny=ny, spp=spp, max_depth=8)


def c2_spheres(seed=1, nx=1200, ny=800, spp=64):
    """C2: the random-spheres layout: ground sphere, 22x22 grid of
    radius-0.2 spheres with materials drawn uniformly, three radius-1 spheres."""
    st = Stream(seed, 2)
    mats = [(LAMBERTIAN, (0.5, 0.5, 0.5), 0.0)]
    sph = [(0.0, -1000.0, 0.0, 1000.0)]
    sph_mat = [0]
    for a in range(-11, 11):
        for b in range(-11, 11):
            choose = st.draw()
            cx = a + 0.9 * st.draw()
            cz = b + 0.9 * st.draw()
            r = [st.draw() for _ in range(6)]
            if np.sqrt((cx - 4.0) ** 2 + (0.2 - 0.2) ** 2 + cz ** 2) <= 0.9:
                continue
            if choose < 0.8:
                mats.append((LAMBERTIAN, (r[0] * r[1], r[2] * r[3], r[4] * r[5]), 0.0))
            elif choose < 0.95:
                mats.append((METAL, (0.5 * (1 + r[0]), 0.5 * (1 + r[1]), 0.5 * (1 + r[2])), 0.5 * r[3]))
            else:
                mats.append((DIELECTRIC, (1.0, 1.0, 1.0), 1.5))
            sph.append((cx, 0.2, cz, 0.2))
            sph_mat.append(len(mats) - 1)
    for centre, m in (((0, 1, 0), (DIELECTRIC, (1, 1, 1), 1.5)),
                      ((-4, 1, 0), (LAMBERTIAN, (0.4, 0.2, 0.1), 0.0)),
                      ((4, 1, 0), (METAL, (0.7, 0.6, 0.5), 0.0))):
        mats.append(m)
        sph.append((centre[0], centre[1], centre[2], 1.0))
        sph_mat.append(len(mats) - 1)
    tri, tri_mat = _empty_tris()
    return dict(name="C2_spheres", tri=tri, tri_mat=tri_mat, sph=np.asarray(sph, np.float32),
                sph_mat=np.asarray(sph_mat, np.uint32), mats=mats,
                camera=camera_dict((13, 2, 3), (0, 0, 0), 20, aperture=0.1, focus_dist=10.0),
                nx=nx, ny=ny, spp=spp, max_depth=8)


def displaced_sphere(n_theta, n_phi, radius=100.0, amp=5.0, centre=(0.0, 0.0, 0.0)):
    """UV sphere with radial displacement amp*sin(8 theta)*sin(6 phi).
    n_theta segments around, n_phi from pole to pole:
    2*n_theta*(n_phi-1) triangles. Returns v0, v1, v2 (outward winding) and the
    latitude band of each triangle."""
    th = (np.arange(n_theta + 1, dtype=np.float64) % n_theta) * (2.0 * np.pi / n_theta)
    ph = np.arange(n_phi + 1, dtype=np.float64) * (np.pi / n_phi)
    T, P = np.meshgrid(th, ph, indexing="xy")          # (n_phi+1, n_theta+1)
    R = radius + amp * np.sin(8.0 * T) * np.sin(6.0 * P)
    X = R * np.sin(P) * np.cos(T) + centre[0]
    Y = R * np.cos(P) + centre[1]
    Z = R * np.sin(P) * np.sin(T) + centre[2]
    V = np.stack([X, Y, Z], axis=-1).astype(np.float32)
    a = V[:-1, :-1]   # (j, i)
    b = V[:-1, 1:]    # (j, i+1)
    c = V[1:, :-1]    # (j+1, i)
    d = V[1:, 1:]     # (j+1, i+1)
    band = np.broadcast_to(np.arange(n_phi)[:, None], (n_phi, n_theta))
    # upper triangles (a, b, d) exist for j >= 1 (a == b at the north pole);
    # lower triangles (a, d, c) exist for j <= n_phi-2 (c == d at the south pole)
    up = (slice(1, None), slice(None))
    lo = (slice(0, n_phi - 1), slice(None))
    v0 = np.concatenate([a[up].reshape(-1, 3), a[lo].reshape(-1, 3)])
    v1 = np.concatenate([b[up].reshape(-1, 3), d[lo].reshape(-1, 3)])
    v2 = np.concatenate([d[up].reshape(-1, 3), c[lo].reshape(-1, 3)])
    bands = np.concatenate([band[up].reshape(-1), band[lo].reshape(-1)])
    return v0, v1, v2, bands


def displaced_sphere_indexed(n_theta, n_phi, radius=100.0, amp=5.0, centre=(0.0, 0.0, 0.0)):
    """The same mesh as displaced_sphere() in indexed form, the shape the reference
    animates (a point array and a face index array, createScene.h:99-140): points
    ((n_phi+1)*(n_theta+1), 3) float32 and idx (n_tri, 3) uint32 such that
    points[idx[:, k]] equals v_k of displaced_sphere() triangle for triangle."""
    th = (np.arange(n_theta + 1, dtype=np.float64) % n_theta) * (2.0 * np.pi / n_theta)
    ph = np.arange(n_phi + 1, dtype=np.float64) * (np.pi / n_phi)
    T, P = np.meshgrid(th, ph, indexing="xy")
    R = radius + amp * np.sin(8.0 * T) * np.sin(6.0 * P)
    X = R * np.sin(P) * np.cos(T) + centre[0]
    Y = R * np.cos(P) + centre[1]
    Z = R * np.sin(P) * np.sin(T) + centre[2]
    points = np.stack([X, Y, Z], axis=-1).astype(np.float32).reshape(-1, 3)
    num = np.arange((n_phi + 1) * (n_theta + 1), dtype=np.uint32).reshape(n_phi + 1, n_theta + 1)
    a, b, c, d = num[:-1, :-1], num[:-1, 1:], num[1:, :-1], num[1:, 1:]
    up = (slice(1, None), slice(None))
    lo = (slice(0, n_phi - 1), slice(None))
    i0 = np.concatenate([a[up].reshape(-1), a[lo].reshape(-1)])
    i1 = np.concatenate([b[up].reshape(-1), d[lo].reshape(-1)])
    i2 = np.concatenate([d[up].reshape(-1), c[lo].reshape(-1)])
    return np.ascontiguousarray(points), np.ascontiguousarray(np.stack([i0, i1, i2], axis=1).astype(np.uint32))


def ground_quad(y, half, normal_down=True):
    p = np.asarray([[-half, y, -half], [half, y, -half], [half, y, half], [-half, y, half]], np.float32)
    v0 = p[[0, 0]]
    v1 = p[[1, 2]]
    v2 = p[[2, 3]]
    n = np.asarray([[0, -1.0 if normal_down else 1.0, 0]] * 2, np.float32)
    return v0, v1, v2, n


def c3_mesh(n_theta=1000, n_phi=500, nx=1920, ny=1080, spp=16, centre=(0.0, 0.0, 0.0)):
    """C3: displaced UV sphere, 1000x500 segments = 998,000 triangles, radius
    100, plus a 2-triangle ground quad; Lambertian on even latitude bands,
    Metal(0.8,0.8,0.8) fuzz 0.1 on odd bands.

    Stored normals point INTO the sphere and the ground's points DOWN: the
    reference triangle test only reports hits with dot(direction, normal) >= 0
    (shapes/triangle.h:61), so this is the orientation that shows
    the mesh as a solid from the outside."""
    v0, v1, v2, bands = displaced_sphere(n_theta, n_phi, centre=centre)
    n = -tri_normals(v0, v1, v2)
    # guard against degenerate pole slivers producing NaN normals
    bad = ~np.isfinite(n).all(axis=1)
    if bad.any():
        n[bad] = np.asarray([0, 1, 0], np.float32)
    g0, g1, g2, gn = ground_quad(-110.0, 2000.0)
    tri = np.concatenate([np.concatenate([v0, v1, v2, n], axis=1),
                          np.concatenate([g0, g1, g2, gn], axis=1)]).astype(np.float32)
    tri_mat = np.concatenate([(bands % 2).astype(np.uint32), np.asarray([2, 2], np.uint32)])
    sph, sph_mat = _empty_sph()
    mats = [(LAMBERTIAN, (0.7, 0.3, 0.2), 0.0), (METAL, (0.8, 0.8, 0.8), 0.1), (LAMBERTIAN, (0.5, 0.5, 0.5), 0.0)]
    # indexed form of the sphere (triangles [0, len(tri) - 2); the ground quad is not animated): the pose input of
    # crt_scene_update_points, as the reference's update_pose takes it (createScene.h:99-109)
    points, tri_idx = displaced_sphere_indexed(n_theta, n_phi, centre=centre)
    return dict(name="C3_mesh_%dk" % (len(tri) // 1000), tri=np.ascontiguousarray(tri), tri_mat=tri_mat, sph=sph,
                sph_mat=sph_mat, mats=mats, camera=camera_dict((0, 60, 320), (0, 0, 0), 40),
                nx=nx, ny=ny, spp=spp, max_depth=8, points=points, tri_idx=tri_idx)


def c4_instanced(grid=10, pitch=250.0, n_theta=500, n_phi=101, nx=3840, ny=2160, spp=64):
    """C4: grid x grid instances of a ~100k-triangle displaced sphere, flattened
    (the reference's only instancing device, HitableList(list, n, Transform*),
    hitable_list.h:11, shares children under a translated frame; flattening the
    translation into the vertices gives the same world-space triangles)."""
    base = c3_mesh(n_theta, n_phi)
    tri0 = base["tri"][:-2]
    mat0 = base["tri_mat"][:-2]
    tris, mats_idx, parts = [], [], []
    off0 = -(grid - 1) * pitch / 2.0
    for a in range(grid):
        for b in range(grid):
            parts.append(np.full(len(tri0), a * grid + b, np.uint32))
            t = tri0.copy()
            shift = np.asarray([off0 + a * pitch, 0.0, off0 + b * pitch], np.float32)
            for k in range(3):
                t[:, 3 * k:3 * k + 3] += shift
            tris.append(t)
            mats_idx.append(mat0)
    g0, g1, g2, gn = ground_quad(-110.0, 20000.0)
    tris.append(np.concatenate([g0, g1, g2, gn], axis=1))
    mats_idx.append(np.asarray([2, 2], np.uint32))
    parts.append(np.full(2, grid * grid, np.uint32))                    # the ground is a part of its own
    out = dict(base)
    for k in ("points", "tri_idx"):                                     # the base mesh's point form does not describe this scene
        out.pop(k, None)
    out.update(name="C4_instanced_%dM" % round(sum(len(t) for t in tris) / 1e6), tri=np.ascontiguousarray(np.concatenate(tris)),
               part_of_tri=np.concatenate(parts), n_parts=grid * grid + 1,
               tri_mat=np.concatenate(mats_idx).astype(np.uint32),
               camera=camera_dict((0, 900, 2600), (0, 0, 0), 40), nx=nx, ny=ny, spp=spp)
    return out


def c5_triangles(n, seed=1):
    """C5 geometry: n triangles with uniform random centres in the unit cube and
    edge length n^(-1/3); random orientation, Triangle-constructor normals."""
    key = rng_key(seed, np.uint32(5), np.uint32(n & 0xFFFFFFFF))
    i = np.arange(n, dtype=np.uint64)

    def u(k):
        return rng_uniform(key, (i * np.uint64(12) + np.uint64(k)).astype(np.uint32)).astype(np.float32)

    c = np.stack([u(0), u(1), u(2)], axis=1)
    edge = np.float32(float(n) ** (-1.0 / 3.0))
    offs = [np.stack([u(3 + 3 * j), u(4 + 3 * j), u(5 + 3 * j)], axis=1) - np.float32(0.5) for j in range(3)]
    v0, v1, v2 = (c + edge * o for o in offs)
    tri = pack_tris(v0, v1, v2)
    bad = ~np.isfinite(tri).all(axis=1)
    tri[bad, 9:12] = np.asarray([0, 1, 0], np.float32)
    sph, sph_mat = _empty_sph()
    return dict(name="C5_%d" % n, tri=tri, tri_mat=np.zeros(n, np.uint32), sph=sph, sph_mat=sph_mat,
                mats=[(LAMBERTIAN, (0.5, 0.5, 0.5), 0.0)], camera=camera_dict((0.5, 0.5, 3.0), (0.5, 0.5, 0.5), 40),
                nx=1920, ny=1080, spp=1, max_depth=8)


def c5_incoherent_rays(n=1 << 21, seed=1):
    """Uniform random origins in the unit cube, uniform directions on the sphere."""
    key = rng_key(seed, np.uint32(55), np.uint32(n & 0xFFFFFFFF))
    i = np.arange(n, dtype=np.uint64)

    def u(k):
        return rng_uniform(key, (i * np.uint64(8) + np.uint64(k)).astype(np.uint32)).astype(np.float32)

    o = np.stack([u(0), u(1), u(2)], axis=1)
    z = np.float32(2.0) * u(3) - np.float32(1.0)
    phi = np.float32(2.0 * np.pi) * u(4)
    r = np.sqrt(np.maximum(np.float32(0), np.float32(1.0) - z * z))
    d = np.stack([r * np.cos(phi), r * np.sin(phi), z], axis=1).astype(np.float32)
    return np.ascontiguousarray(np.concatenate([o, d], axis=1), dtype=np.float32)


def primary_rays(cam22, nx, ny, sx=0.5, sy=0.5):
    """Pinhole primary rays through pixel centres (no lens, no jitter):
    Camera::get_ray with offset 0 (core/camera.h:40-44), float32 op order kept.
    Row-major, row 0 at the bottom like the reference's colorBuffer."""
    c = np.asarray(cam22, np.float32)
    llc, hor, ver, org = c[0:3], c[3:6], c[6:9], c[9:12]
    x = np.arange(nx, dtype=np.float32)
    y = np.arange(ny, dtype=np.float32)
    u = ((x + np.float32(sx)) / np.float32(nx)).astype(np.float32)
    v = ((y + np.float32(sy)) / np.float32(ny)).astype(np.float32)
    U, Vv = np.meshgrid(u, v, indexing="xy")
    U = U.reshape(-1, 1)
    Vv = Vv.reshape(-1, 1)
    zero = np.zeros(3, np.float32)
    d = (((llc + U * hor) + Vv * ver) - org) - zero
    o = np.broadcast_to(org + zero, d.shape)
    return np.ascontiguousarray(np.concatenate([o, d], axis=1), dtype=np.float32)


def scene_bounds(scene):
    lo = np.full(3, np.inf)
    hi = np.full(3, -np.inf)
    if len(scene["tri"]):
        v = scene["tri"][:, :9].reshape(-1, 3)
        lo = np.minimum(lo, v.min(0))
        hi = np.maximum(hi, v.max(0))
    if len(scene["sph"]):
        s = scene["sph"]
        lo = np.minimum(lo, (s[:, :3] - np.abs(s[:, 3:4])).min(0))
        hi = np.maximum(hi, (s[:, :3] + np.abs(s[:, 3:4])).max(0))
    return lo, hi


def light_panels(nx=160, ny=120, spp=8):
    """Triangles only, for the DiffuseLight path (material/material.h:146-159, core/render.h:56-61): a Lambertian
    wall and floor, a light panel in front of the wall (seen directly), a second one behind it, a metal and a glass
    block, through a lens with a finite aperture. A reference triangle answers only rays that travel ALONG its
    stored normal and scatters along that normal too, i.e. onwards through the surface: the paths that
    reach the second panel are the ones scattered by the wall."""
    def quad(a, b, c, d, n):
        a, b, c, d, n = (np.asarray(v, np.float32) for v in (a, b, c, d, n))
        return [np.concatenate([a, b, c, n]), np.concatenate([a, c, d, n])]

    def block(lo, hi):
        (x0, y0, z0), (x1, y1, z1) = lo, hi
        f = []
        f += quad((x0, y0, z0), (x0, y1, z0), (x1, y1, z0), (x1, y0, z0), (0, 0, 1))      # normals point into the block
        f += quad((x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1), (0, 0, -1))
        f += quad((x0, y0, z0), (x0, y0, z1), (x0, y1, z1), (x0, y1, z0), (1, 0, 0))
        f += quad((x1, y0, z0), (x1, y1, z0), (x1, y1, z1), (x1, y0, z1), (-1, 0, 0))
        f += quad((x0, y0, z0), (x1, y0, z0), (x1, y0, z1), (x0, y0, z1), (0, 1, 0))
        f += quad((x0, y1, z0), (x0, y1, z1), (x1, y1, z1), (x1, y1, z0), (0, -1, 0))
        return f

    mats = [(LAMBERTIAN, (0.73, 0.73, 0.73), 0.0), (LAMBERTIAN, (0.65, 0.05, 0.05), 0.0),
            (DIFFUSE_LIGHT, (4.0, 3.5, 3.0), 0.0), (METAL, (0.8, 0.85, 0.88), 0.05), (DIELECTRIC, (1.0, 1.0, 1.0), 1.5)]
    tris, tm = [], []

    def add(faces, m):
        tris.extend(faces)
        tm.extend([m] * len(faces))

    add(quad((-12, 0, 0), (-12, 20, 0), (12, 20, 0), (12, 0, 0), (0, 0, -1)), 0)           # wall, camera side is +z
    add(quad((-12, 0, 0), (12, 0, 0), (12, 0, 14), (-12, 0, 14), (0, -1, 0)), 1)           # floor
    add(quad((5, 13, 2), (5, 17, 2), (9, 17, 2), (9, 13, 2), (0, 0, -1)), 2)               # panel in front of the wall
    add(quad((-6, 4, -5), (-6, 16, -5), (6, 16, -5), (6, 4, -5), (0, 0, -1)), 2)           # panel behind the wall
    add(block((-7, 0, 3), (-3, 7, 7)), 3)
    add(block((1, 0, 4), (5, 4, 8)), 4)
    sph, sph_mat = _empty_sph()
    return dict(name="light_panels", tri=np.ascontiguousarray(np.stack(tris), np.float32), tri_mat=np.asarray(tm, np.uint32),
                sph=sph, sph_mat=sph_mat, mats=mats, camera=camera_dict((0, 9, 22), (0, 8, 0), 50, aperture=0.3, focus_dist=18.0),
                nx=nx, ny=ny, spp=spp, max_depth=8)


def rotated_scene(seed=1, n_items=24, nx=160, ny=120, spp=4):
    """Triangles, spheres and rectangles under the reference's Transform (hitable/transform.h): own rotations and
    translations, identity levels, and up to three nested HitableLists, beside plain primitives - the cases of
    VERDICT r1 row a8. Lengths ~1, spread over ~10 units around the origin."""
    rng = np.random.default_rng(seed)
    f = np.float32
    mats = [(LAMBERTIAN, (0.7, 0.3, 0.2), 0.0), (METAL, (0.8, 0.8, 0.9), 0.1), (DIELECTRIC, (1, 1, 1), 1.5),
            (DIFFUSE_LIGHT, (3.0, 2.5, 2.0), 0.0), (LAMBERTIAN, (0.2, 0.3, 0.1), 0.0, (0.9, 0.9, 0.9))]

    def tris(n, scale=2.0):
        v0 = rng.normal(size=(n, 3)) * 0.5
        v1 = v0 + rng.normal(size=(n, 3)) * scale
        v2 = v0 + rng.normal(size=(n, 3)) * scale
        nrm = np.cross(v1 - v0, v2 - v0)
        nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
        nrm *= rng.choice([-1.0, 1.0], size=(n, 1))
        return np.concatenate([v0, v1, v2, nrm], axis=1).astype(f)

    def level(kind):
        pos = (rng.normal(size=3) * 4.0).astype(f)
        rot = (rng.normal(size=3) * [60.0, 120.0, 45.0]).astype(f)
        if kind == "identity":
            return (np.zeros(3, f), np.zeros(3, f))
        if kind == "translate":
            return (pos, np.zeros(3, f))
        if kind == "rotate":
            return (np.zeros(3, f), rot)
        if kind == "axis":                       # multiples of 90 degrees: sin/cos land on (almost) 0 and 1
            return (pos, (rng.integers(-2, 3, size=3) * 90).astype(f))
        return (pos, rot)

    kinds = ["full", "full", "rotate", "translate", "identity", "axis"]
    chained = []
    for i in range(n_items):
        depth = 1 + (i % 3)
        levels = [level(kinds[int(rng.integers(0, len(kinds)))]) for _ in range(depth)]
        what = i % 4
        n = int(rng.integers(1, 4))
        mat = rng.integers(0, len(mats), size=n).astype(np.uint32)
        if what in (0, 3):
            chained.append(dict(kind=0, data=tris(n), mat=mat, levels=levels))
        elif what == 1:
            chained.append(dict(kind=1, data=(rng.uniform(0.5, 2.0, size=n) * rng.choice([1.0, 1.0, -1.0], size=n)).astype(f), mat=mat, levels=levels))
        else:
            chained.append(dict(kind=2, data=None, mat=mat, flip=rng.integers(0, 2, size=n).astype(np.uint8), levels=levels))
    plain = tris(6)
    plain[:, 0:9] += np.tile(rng.normal(size=(6, 3)) * 4.0, (1, 3)).astype(f)
    sph = np.concatenate([rng.normal(size=(5, 3)) * 4.0, rng.uniform(0.5, 1.5, size=(5, 1))], axis=1).astype(f)
    return dict(name="rotated_%d" % seed, tri=np.ascontiguousarray(plain), tri_mat=rng.integers(0, len(mats), size=6).astype(np.uint32),
                sph=np.ascontiguousarray(sph), sph_mat=rng.integers(0, len(mats), size=5).astype(np.uint32), mats=mats, chained=chained,
                camera=camera_dict((0, 3, 26), (0, 0, 0), 40, aperture=0.05, focus_dist=24.0), nx=nx, ny=ny, spp=spp, max_depth=8)


def cornell_box(nx=200, ny=200, spp=8, obj_path=None):
    """A Cornell box out of the reference's own parts (row f3 of SURVEY 8f): five walls and a ceiling light that are
    Rectangles - the unit square of shapes/rectangle.h, turned and placed by Transform rotations (scale stays 1,
    SURVEY 9 Q7, so the box is one unit wide) - the short block of objects/cbox_smallbox.obj (a//n faces, scaled by
    1/555 into the unit box) as triangles under two nested lists, a glass sphere under a rotated Transform, a
    checkered floor. A Rectangle answers rays that travel AGAINST its normal (rectangle.h:24), a Triangle rays that
    travel ALONG its stored normal (triangle.h:61)."""
    f = np.float32
    mats = [(LAMBERTIAN, (0.73, 0.73, 0.73), 0.0), (LAMBERTIAN, (0.65, 0.05, 0.05), 0.0), (LAMBERTIAN, (0.12, 0.45, 0.15), 0.0),
            (DIFFUSE_LIGHT, (1.2, 1.2, 1.1), 0.0), (LAMBERTIAN, (0.73, 0.73, 0.73), 0.0, (0.2, 0.2, 0.6)), (DIELECTRIC, (1, 1, 1), 1.5),
            (METAL, (0.8, 0.85, 0.88), 0.0)]

    def rect(pos, rot, mat, flip=0):
        return dict(kind=2, data=None, mat=np.asarray([mat], np.uint32), flip=np.asarray([flip], np.uint8),
                    levels=[(np.asarray(pos, f), np.asarray(rot, f))])

    # The box spans [-0.5, 0.5]^3 and the camera looks down -z from z = 2.4. A frame maps WORLD to LOCAL: local =
    # rotate(world) - position (transform.h:38-48). A wall whose inward normal is N sits at C = -0.5 N; its rotation
    # must turn N into the rectangle's +z, and then position = rotate(C) = (0, 0, -0.5) for every wall.
    wall = (0.0, 0.0, -0.5)
    chained = [
        rect(wall, (0, 0, 0), 0),                              # back wall, N = +z
        rect(wall, (90, 0, 0), 4),                             # floor, N = +y, checkered
        rect(wall, (-90, 0, 0), 0),                            # ceiling, N = -y
        rect(wall, (0, -90, 0), 1),                            # wall at x = -0.5, N = +x, red
        rect(wall, (0, 90, 0), 2),                             # wall at x = +0.5, N = -x, green
        rect((0.0, 0.0, -0.49), (-90, 0, 0), 3),               # a luminous sheet just below the ceiling
    ]
    tri = np.zeros((0, 12), f)
    if obj_path is None:
        obj_path = os.path.join(_GOLDEN, "cbox_smallbox.npz")
    if os.path.exists(obj_path):
        g = np.load(obj_path)
        pts = (g["points"].astype(f) / f(555.0)).astype(f)
        idx = g["idx"].astype(np.int64) - 1
        v0, v1, v2 = pts[idx[:, 2]], pts[idx[:, 1]], pts[idx[:, 0]]     # winding reversed as createScene.h:183
        tri = pack_tris(v0, v1, v2)
        tri[:, 9:12] = -tri_normals(v0, v1, v2)                           # stored normals point into the block
    if len(tri):
        # the block (centred on its own origin by the loader) sits under two nested lists - a shift onto the floor, a
        # turn about y - and every triangle carries the same small offset of its own
        chained.append(dict(kind=0, data=tri, mat=np.full(len(tri), 6, np.uint32),
                            levels=[(np.asarray((0.17, -0.351, 0.1), f), np.zeros(3, f)), (np.zeros(3, f), np.asarray((0, 20, 0), f)),
                                    (np.asarray((0.02, 0.0, -0.03), f), np.zeros(3, f))]))
    chained.append(dict(kind=1, data=np.asarray([0.17], f), mat=np.asarray([5], np.uint32),
                        levels=[(rotate_point((-0.22, -0.33, 0.18), (0, 30, 0)), np.asarray((0, 30, 0), f))]))
    e_tri, e_mat = _empty_tris()
    e_sph, e_smat = _empty_sph()
    return dict(name="cornell_unit", tri=e_tri, tri_mat=e_mat, sph=e_sph, sph_mat=e_smat, mats=mats, chained=chained,
                camera=camera_dict((0, 0, 2.4), (0, 0, 0), 40), nx=nx, ny=ny, spp=spp, max_depth=8)


def checker_scene(nx=160, ny=100, spp=4):
    """CheckerTexture (material/texture.h:25-42) under a Lambertian and under a DiffuseLight, on triangles and on
    spheres: a checkered ground sphere in the manner of the reference's random scene, a checkered wall of triangles
    (normals away from the camera), a checkered light, a metal and a glass sphere."""
    mats = [(LAMBERTIAN, (0.2, 0.3, 0.1), 0.0, (0.9, 0.9, 0.9)), (LAMBERTIAN, (0.8, 0.1, 0.1), 0.0, (0.1, 0.1, 0.8)),
            (DIFFUSE_LIGHT, (4.0, 4.0, 4.0), 0.0, (0.5, 2.0, 0.5)), (METAL, (0.7, 0.6, 0.5), 0.0), (DIELECTRIC, (1.0, 1.0, 1.0), 1.5),
            (LAMBERTIAN, (0.4, 0.2, 0.1), 0.0)]
    sph = np.asarray([[0, -1000, 0, 1000], [0, 1, 0, 1.0], [-4, 1, 0, 1.0], [4, 1, 0, 1.0], [0, 3.2, -1.0, 0.7]], np.float32)
    sph_mat = np.asarray([0, 4, 5, 3, 2], np.uint32)

    def quad(a, b, c, d, n):
        a, b, c, d, n = (np.asarray(v, np.float32) for v in (a, b, c, d, n))
        return [np.concatenate([a, b, c, n]), np.concatenate([a, c, d, n])]

    tri = np.stack(quad((-8, 0, -4), (-8, 6, -4), (8, 6, -4), (8, 0, -4), (0.3, -0.1, -1.0)))
    return dict(name="checker", tri=np.ascontiguousarray(tri, np.float32), tri_mat=np.asarray([1, 1], np.uint32), sph=sph, sph_mat=sph_mat,
                mats=mats, camera=camera_dict((13, 2, 3), (0, 0, 0), 20, aperture=0.1, focus_dist=10.0),
                nx=nx, ny=ny, spp=spp, max_depth=8)
