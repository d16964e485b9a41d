"""cudaraytracer_b200 — B200-native render hot path for CudaRaytracer scenes.

The product is ``libcrt_b200.so`` (hand-written sm_100a CUDA behind the C ABI of
``include/crt.h``). This module is the thin ctypes binding used by the tests and
``bench.py``; there is no CPU path: importing works anywhere, but creating a
``Context`` without a CUDA device or without the built library raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CRT_LIB") or os.path.join(_HERE, "libcrt_b200.so")   # CRT_LIB: development aid (A/B builds)

MISS = 0xFFFFFFFF
SHADE_PATH, SHADE_LAMBERT, SHADE_NORMAL = 0, 1, 2
ACCEL_LBVH, ACCEL_BRUTE, ACCEL_LBVH_PACKET = 0, 1, 2
OPT_LEAF_MAX, OPT_COUNTERS, OPT_BATCH_PATHS, OPT_REFILL, OPT_LEAF_THRESH, OPT_PACKET_PRIMARY, OPT_LANES, OPT_BUILDER = 1, 2, 3, 4, 5, 6, 7, 8
OPT_WIDE_BELOW = 9
OPT_TIMING, OPT_SLICE_MIN, OPT_TAIL_HINT, OPT_PACKET_DEFER, OPT_PACKET_DEFER_BELOW, OPT_BATCH_SAMPLES = 10, 11, 12, 13, 14, 15
COMM_ID_BYTES = 128
DBG_NODES, DBG_RECS, DBG_KEYS, DBG_PERM, DBG_CHILD, DBG_RANGE, DBG_LEAF_BOX, DBG_NODE_BOX = range(8)
TRACE_STATIC_MAX_RAYS = 65536
BUILDER_LBVH, BUILDER_PLOC = 0, 1
TEX_CONSTANT, TEX_CHECKER = 0, 1
POSE_VERTICES, POSE_POINTS, POSE_PARTS = 0, 1, 2
TILE = 32

# every symbol include/crt.h declares
ABI_SYMBOLS = [
    "crt_abi_version", "crt_create", "crt_destroy", "crt_last_error", "crt_set_option", "crt_sync", "crt_stream",
    "crt_scene_set_materials", "crt_scene_set_triangles", "crt_scene_set_spheres", "crt_scene_update_vertices",
    "crt_scene_set_frames", "crt_scene_set_triangle_chains", "crt_scene_set_sphere_chains", "crt_scene_set_rectangles",
    "crt_scene_set_indices", "crt_scene_update_points", "crt_scene_set_parts", "crt_scene_move_parts",
    "crt_render_frame_pose", "crt_scene_pose_async",
    "crt_accel_build", "crt_accel_refit", "crt_trace", "crt_trace_device", "crt_render", "crt_render_frame",
    "crt_render_accum_device", "crt_resolve_device", "crt_pack_rgba8_device", "crt_get_counters",
    "crt_reset_counters", "crt_accel_info",
    "crt_comm_unique_id", "crt_comm_init", "crt_comm_destroy", "crt_comm_info", "crt_scene_broadcast",
    "crt_reduce_accum_device", "crt_gather_tiles_device", "crt_gather_device", "crt_render_dist", "crt_trace_static",
    "crt_debug_sort_pairs", "crt_debug_read",
]


class CrtError(RuntimeError):
    pass


class Material(C.Structure):
    _fields_ = [("type", C.c_uint32), ("albedo", C.c_float * 3), ("param", C.c_float),
                ("texture", C.c_uint32), ("albedo2", C.c_float * 3)]


class Camera(C.Structure):
    _fields_ = [("v", C.c_float * 22)]


class Counters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64),
                ("spheres_tested", C.c_uint64), ("scatters", C.c_uint64), ("paths", C.c_uint64),
                ("build_ms", C.c_float), ("refit_ms", C.c_float), ("render_ms", C.c_float), ("trace_ms", C.c_float),
                ("bounce_ms", C.c_float), ("trace_queue_ms", C.c_float),
                ("node_fetches", C.c_uint64), ("tri_fetches", C.c_uint64), ("sphere_fetches", C.c_uint64),
                ("launches", C.c_uint64), ("lanes", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None
_libm = C.CDLL("libm.so.6")
_libm.tanf.restype = C.c_float
_libm.tanf.argtypes = [C.c_float]
_tanf = _libm.tanf


def load_library():
    """dlopen libcrt_b200.so and declare the ABI. Fails loudly when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CrtError("%s is not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, sz = C.c_void_p, C.c_size_t
    lib.crt_abi_version.restype = C.c_int
    lib.crt_last_error.restype = C.c_char_p
    lib.crt_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.crt_destroy.argtypes = [vp]
    lib.crt_destroy.restype = None
    lib.crt_set_option.argtypes = [vp, C.c_int, C.c_int64]
    lib.crt_sync.argtypes = [vp]
    lib.crt_stream.argtypes = [vp]
    lib.crt_stream.restype = vp
    lib.crt_scene_set_materials.argtypes = [vp, vp, sz]
    lib.crt_scene_set_triangles.argtypes = [vp, vp, vp, vp, sz]
    lib.crt_scene_set_spheres.argtypes = [vp, vp, vp, vp, sz]
    lib.crt_scene_update_vertices.argtypes = [vp, vp, sz, sz]
    lib.crt_scene_set_frames.argtypes = [vp, vp, sz]
    lib.crt_scene_set_triangle_chains.argtypes = [vp, vp, sz]
    lib.crt_scene_set_sphere_chains.argtypes = [vp, vp, sz]
    lib.crt_scene_set_rectangles.argtypes = [vp, vp, vp, vp, vp, sz]
    lib.crt_scene_set_indices.argtypes = [vp, vp, sz, sz, sz]
    lib.crt_scene_update_points.argtypes = [vp, vp, sz]
    lib.crt_render_frame_pose.argtypes = [vp, C.c_int, vp, sz, sz, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, vp]
    lib.crt_scene_pose_async.argtypes = [vp, C.c_int, vp, sz, sz]
    lib.crt_scene_set_parts.argtypes = [vp, vp, sz, sz]
    lib.crt_scene_move_parts.argtypes = [vp, vp, sz]
    lib.crt_accel_build.argtypes = [vp]
    lib.crt_accel_refit.argtypes = [vp]
    lib.crt_trace.argtypes = [vp, vp, sz, C.c_float, C.c_float, C.c_int, vp]
    lib.crt_trace_static.argtypes = [vp, vp, sz, C.c_float, C.c_float, vp]
    lib.crt_trace_device.argtypes = [vp, vp, sz, C.c_float, C.c_float, C.c_int, vp]
    lib.crt_render.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, vp]
    lib.crt_render_frame.argtypes = [vp, vp, sz, sz, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, vp]
    lib.crt_render_accum_device.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.c_uint64, C.c_int, C.c_int, vp]
    lib.crt_resolve_device.argtypes = [vp, vp, sz, C.c_int, vp]
    lib.crt_pack_rgba8_device.argtypes = [vp, vp, C.c_int, C.c_int, vp]
    lib.crt_get_counters.argtypes = [vp, C.POINTER(Counters)]
    lib.crt_reset_counters.argtypes = [vp]
    lib.crt_accel_info.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                   C.POINTER(C.c_int)]
    lib.crt_comm_unique_id.argtypes = [vp]
    lib.crt_comm_init.argtypes = [vp, vp, C.c_int, C.c_int]
    lib.crt_comm_destroy.argtypes = [vp]
    lib.crt_comm_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.crt_scene_broadcast.argtypes = [vp, C.c_int]
    lib.crt_reduce_accum_device.argtypes = [vp, vp, sz, C.c_int]
    lib.crt_debug_sort_pairs.argtypes = [vp, vp, vp, sz, C.c_int, C.c_int]
    lib.crt_debug_read.argtypes = [vp, C.c_int, vp, sz, C.POINTER(C.c_int)]
    lib.crt_gather_tiles_device.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int]
    lib.crt_gather_device.argtypes = [vp, vp, sz, vp, C.c_int]
    lib.crt_render_dist.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, vp]
    for name in ABI_SYMBOLS:
        if name not in ("crt_destroy", "crt_last_error", "crt_stream"):
            getattr(lib, name).restype = C.c_int
    _lib = lib
    return lib


def comm_unique_id():
    """crt_comm_unique_id: the 128-byte id rank 0 hands to every rank before comm_init."""
    lib = load_library()
    buf = C.create_string_buffer(COMM_ID_BYTES)
    rc = lib.crt_comm_unique_id(C.cast(buf, C.c_void_p))
    if rc != 0:
        raise CrtError("crt error %d: %s" % (rc, lib.crt_last_error().decode()))
    return buf.raw


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def camera22(lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist):
    """Reference Camera constructor (core/camera.h:18-38) in float32 with its
    operation order; the host computes it once and passes it by value."""
    f = np.float32
    lookfrom, lookat, vup = (np.asarray(v, f) for v in (lookfrom, lookat, vup))

    def unit(v):
        return (v / np.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2], dtype=f)).astype(f)

    def cross(a, b):
        return np.asarray([a[1] * b[2] - a[2] * b[1], -(a[0] * b[2] - a[2] * b[0]), a[0] * b[1] - a[1] * b[0]], f)

    lens_radius = f(aperture) / f(2.0)
    theta = f(np.float64(f(vfov)) * np.pi / np.float64(f(180.0)))
    half_height = f(_tanf(float(f(theta / f(2.0)))))          # libm tanf, as the host-compiled reference calls
    half_width = f(half_height * f(aspect))
    fd = f(focus_dist)
    origin = lookfrom
    z = unit(lookfrom - lookat)
    x = unit(cross(vup, z))
    y = cross(z, x)
    llc = ((origin - f(half_width * fd) * x) - f(half_height * fd) * y) - fd * z
    hor = f(f(f(2.0) * half_width) * fd) * x
    ver = f(f(f(2.0) * half_height) * fd) * y
    return np.concatenate([llc, hor, ver, origin, x, y, z, [lens_radius]]).astype(f)


def camera_for(scene, nx=None, ny=None):
    c = scene["camera"]
    nx = scene["nx"] if nx is None else nx
    ny = scene["ny"] if ny is None else ny
    return camera22(c["lookfrom"], c["lookat"], c["vup"], c["vfov"], np.float32(nx) / np.float32(ny),
                    c["aperture"], c["focus_dist"])


class Context:
    """One GPU's render context (crt_ctx). Mirrors the reference's call order:
    set scene -> build accel -> (update vertices -> refit)* -> render / trace."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        self._h = None
        self._check(self.lib.crt_create(int(device), C.byref(h)))
        self._h = h

    def _check(self, rc):
        if rc != 0:
            raise CrtError("crt error %d: %s" % (rc, self.lib.crt_last_error().decode()))

    def close(self):
        if self._h is not None:
            self.lib.crt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- scene ----
    def set_option(self, opt, value):
        self._check(self.lib.crt_set_option(self._h, int(opt), int(value)))

    def set_materials(self, mats):
        arr = (Material * max(len(mats), 1))()
        for i, m in enumerate(mats):          # (type, albedo, param) or, with a CheckerTexture, (type, even, param, odd)
            t, alb, p = m[0], m[1], m[2]
            arr[i].type = int(t)
            arr[i].albedo[0], arr[i].albedo[1], arr[i].albedo[2] = (float(np.float32(v)) for v in alb)
            arr[i].param = float(np.float32(p))
            arr[i].texture = TEX_CONSTANT
            if len(m) > 3 and m[3] is not None:
                arr[i].texture = TEX_CHECKER
                arr[i].albedo2[0], arr[i].albedo2[1], arr[i].albedo2[2] = (float(np.float32(v)) for v in m[3])
        self._check(self.lib.crt_scene_set_materials(self._h, C.cast(arr, C.c_void_p), len(mats)))

    def set_triangles(self, tri, mat=None, ids=None):
        tri = np.ascontiguousarray(tri, np.float32).reshape(-1, 12)
        mat = None if mat is None else np.ascontiguousarray(mat, np.uint32)
        ids = None if ids is None else np.ascontiguousarray(ids, np.uint32)
        self._check(self.lib.crt_scene_set_triangles(self._h, _ptr(tri), _ptr(mat), _ptr(ids), tri.shape[0]))

    def set_spheres(self, sph, mat=None, ids=None):
        sph = np.ascontiguousarray(sph, np.float32).reshape(-1, 4)
        mat = None if mat is None else np.ascontiguousarray(mat, np.uint32)
        ids = None if ids is None else np.ascontiguousarray(ids, np.uint32)
        self._check(self.lib.crt_scene_set_spheres(self._h, _ptr(sph), _ptr(mat), _ptr(ids), sph.shape[0]))

    def set_frames(self, frames):
        """crt_scene_set_frames: (n, 13) float32 rows = 9 matrix entries, position, has_rotation."""
        fr = np.ascontiguousarray(frames, np.float32).reshape(-1, 13)
        raw = np.zeros((len(fr), 13), np.uint32)
        raw[:, :12] = fr[:, :12].view(np.uint32)
        raw[:, 12] = (fr[:, 12] != 0).astype(np.uint32)
        self._check(self.lib.crt_scene_set_frames(self._h, _ptr(raw), raw.shape[0]))

    def set_triangle_chains(self, chain):
        chain = np.ascontiguousarray(chain, np.uint32)
        self._check(self.lib.crt_scene_set_triangle_chains(self._h, _ptr(chain), chain.shape[0]))

    def set_sphere_chains(self, chain):
        chain = np.ascontiguousarray(chain, np.uint32)
        self._check(self.lib.crt_scene_set_sphere_chains(self._h, _ptr(chain), chain.shape[0]))

    def set_rectangles(self, chain, flip, mat, ids):
        chain, mat, ids = (np.ascontiguousarray(a, np.uint32) for a in (chain, mat, ids))
        flip = np.ascontiguousarray(flip, np.uint8)
        self._check(self.lib.crt_scene_set_rectangles(self._h, _ptr(chain), _ptr(flip), _ptr(mat), _ptr(ids), chain.shape[0]))

    def set_scene(self, scene):
        """Upload a scenes.py dict; triangles get ids 0..n_tri-1, spheres follow, then the primitives of
        scene["chained"] (rotated / nested Hitables and rectangles, scenes.flatten_chained) in list order."""
        self.set_materials(scene["mats"])
        if scene.get("chained"):
            from . import scenes as _scenes
            f = _scenes.flatten_chained(scene)
            self.set_frames(f["frames"])
            self.set_triangles(f["tri"], f["tri_mat"], f["tri_id"])
            self.set_triangle_chains(f["tri_chain"])
            self.set_spheres(f["sph"], f["sph_mat"], f["sph_id"])
            self.set_sphere_chains(f["sph_chain"])
            self.set_rectangles(f["rect_chain"], f["rect_flip"], f["rect_mat"], f["rect_id"])
            return
        self.set_frames(np.zeros((0, 13), np.float32))
        self.set_triangles(scene["tri"], scene["tri_mat"])
        self.set_spheres(scene["sph"], scene["sph_mat"])

    def update_vertices(self, verts, first=0):
        verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        self._check(self.lib.crt_scene_update_vertices(self._h, _ptr(verts), int(first), verts.shape[0]))

    def set_indices(self, idx, n_points, first=0):
        """Index buffer of an indexed mesh (createScene.h:99-109): triangle first + t stores
        points[idx[t, 0]], points[idx[t, 1]], points[idx[t, 2]]."""
        idx = np.ascontiguousarray(idx, np.uint32).reshape(-1, 3)
        self._check(self.lib.crt_scene_set_indices(self._h, _ptr(idx), int(first), idx.shape[0], int(n_points)))

    def update_points(self, points):
        pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
        self._check(self.lib.crt_scene_update_points(self._h, _ptr(pts), pts.shape[0]))

    def set_parts(self, part_of_tri, n_parts):
        """Bone partition (createScene.h:221-307): the part every triangle moves with; the
        current vertices become the rest pose."""
        part = np.ascontiguousarray(part_of_tri, np.uint32).reshape(-1)
        self._check(self.lib.crt_scene_set_parts(self._h, _ptr(part), part.shape[0], int(n_parts)))

    def move_parts(self, translation):
        """vertex = rest vertex + translation[part]; follow with refit()."""
        t = np.ascontiguousarray(translation, np.float32).reshape(-1, 3)
        self._check(self.lib.crt_scene_move_parts(self._h, _ptr(t), t.shape[0]))

    def pose_async(self, kind, pose, first=0):
        """crt_scene_pose_async: queue pose upload + refit; the next render_accum_device overlaps its ray generation
        with them. `pose` must stay alive and untouched until the next call on this context returns."""
        width = 9 if kind == POSE_VERTICES else 3
        pose = np.ascontiguousarray(pose, np.float32).reshape(-1, width)
        self._pose_keepalive = pose
        self._check(self.lib.crt_scene_pose_async(self._h, int(kind), _ptr(pose), int(first), pose.shape[0]))

    def build(self):
        self._check(self.lib.crt_accel_build(self._h))

    def refit(self):
        self._check(self.lib.crt_accel_refit(self._h))

    def accel_info(self):
        a, b, c, d = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_int()
        self._check(self.lib.crt_accel_info(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return dict(n_prims=a.value, n_nodes=b.value, bytes=c.value, max_depth=d.value)

    # ---- closest hit ----
    def trace(self, rays, t_min=0.001, t_max=3.4028234663852886e38, accel=ACCEL_LBVH):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        hits = np.zeros((n, 4), np.float32)
        self._check(self.lib.crt_trace(self._h, _ptr(rays), n, float(t_min), float(t_max), int(accel), _ptr(hits)))
        t = hits[:, 0].copy()
        pid = hits[:, 1].copy().view(np.uint32)
        return t, pid, hits[:, 2].copy(), hits[:, 3].copy()

    def trace_static(self, rays, t_min=0.001, t_max=3.4028234663852886e38):
        """crt_trace_static: crt_trace (LBVH) with its device scratch in a static array of the library - no allocation,
        at most TRACE_STATIC_MAX_RAYS rays per call."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        hits = np.zeros((n, 4), np.float32)
        self._check(self.lib.crt_trace_static(self._h, _ptr(rays), n, float(t_min), float(t_max), _ptr(hits)))
        return hits[:, 0].copy(), hits[:, 1].copy().view(np.uint32), hits[:, 2].copy(), hits[:, 3].copy()

    def trace_device(self, d_rays_ptr, n, d_hits_ptr, t_min=0.001, t_max=3.4028234663852886e38, accel=ACCEL_LBVH):
        self._check(self.lib.crt_trace_device(self._h, C.c_void_p(d_rays_ptr), int(n), float(t_min), float(t_max),
                                              int(accel), C.c_void_p(d_hits_ptr)))

    # ---- render ----
    def render(self, cam22, nx, ny, spp, max_depth=8, mode=SHADE_PATH, seed=1, out=None):
        cam = np.ascontiguousarray(cam22, np.float32)
        assert cam.size == 22
        rgb = np.empty((ny, nx, 3), np.float32) if out is None else out
        self._check(self.lib.crt_render(self._h, _ptr(cam), nx, ny, spp, max_depth, int(mode), int(seed), _ptr(rgb)))
        return rgb

    def render_frame(self, verts, cam22, nx, ny, spp, max_depth=8, mode=SHADE_PATH, seed=1, out=None, first=0):
        """One animation frame: new vertices for triangles [first, first+len(verts)), refit, render, read back
        (crt_render_frame: the three steps overlap). verts may be None for a frame without a geometry change."""
        cam = np.ascontiguousarray(cam22, np.float32)
        assert cam.size == 22
        n = 0
        if verts is not None:
            verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
            n = verts.shape[0]
        rgb = np.empty((ny, nx, 3), np.float32) if out is None else out
        self._check(self.lib.crt_render_frame(self._h, _ptr(verts), int(first), n, _ptr(cam), nx, ny, spp, max_depth,
                                              int(mode), int(seed), _ptr(rgb)))
        return rgb

    def render_frame_pose(self, kind, pose, cam22, nx, ny, spp, max_depth=8, mode=SHADE_PATH, seed=1, out=None):
        """crt_render_frame_pose: like render_frame with the pose given as points (POSE_POINTS, after set_indices)
        or as one translation per part (POSE_PARTS, after set_parts)."""
        cam = np.ascontiguousarray(cam22, np.float32)
        assert cam.size == 22
        width = 9 if kind == POSE_VERTICES else 3
        pose = np.ascontiguousarray(pose, np.float32).reshape(-1, width)
        rgb = np.empty((ny, nx, 3), np.float32) if out is None else out
        self._check(self.lib.crt_render_frame_pose(self._h, int(kind), _ptr(pose), 0, pose.shape[0], _ptr(cam), nx, ny, spp,
                                                   max_depth, int(mode), int(seed), _ptr(rgb)))
        return rgb

    def render_accum_device(self, cam22, nx, ny, s0, s1, max_depth, mode, seed, rank, world, d_accum_ptr):
        cam = np.ascontiguousarray(cam22, np.float32)
        self._check(self.lib.crt_render_accum_device(self._h, _ptr(cam), nx, ny, s0, s1, max_depth, int(mode),
                                                     int(seed), rank, world, C.c_void_p(d_accum_ptr)))

    def resolve_device(self, d_accum_ptr, n_pixels, spp, d_rgb_ptr):
        self._check(self.lib.crt_resolve_device(self._h, C.c_void_p(d_accum_ptr), int(n_pixels), int(spp),
                                                C.c_void_p(d_rgb_ptr)))

    def pack_rgba8_device(self, d_rgb_ptr, nx, ny, d_rgba_ptr):
        self._check(self.lib.crt_pack_rgba8_device(self._h, C.c_void_p(d_rgb_ptr), nx, ny, C.c_void_p(d_rgba_ptr)))

    # ---- several GPUs (comm.cu) ----
    def comm_init(self, comm_id, rank, world):
        """Collective: join the communicator `comm_id` (bytes from comm_unique_id() on rank 0) as `rank` of `world`."""
        buf = C.create_string_buffer(bytes(comm_id), COMM_ID_BYTES)
        self._check(self.lib.crt_comm_init(self._h, C.cast(buf, C.c_void_p), int(rank), int(world)))

    def comm_destroy(self):
        self._check(self.lib.crt_comm_destroy(self._h))

    def comm_info(self):
        r, w, v = C.c_int(), C.c_int(), C.c_int()
        self._check(self.lib.crt_comm_info(self._h, C.byref(r), C.byref(w), C.byref(v)))
        return dict(rank=r.value, world=w.value, nccl_version=v.value)

    def scene_broadcast(self, root=0):
        self._check(self.lib.crt_scene_broadcast(self._h, int(root)))

    def reduce_accum_device(self, d_accum_ptr, n_floats, root=0):
        self._check(self.lib.crt_reduce_accum_device(self._h, C.c_void_p(d_accum_ptr), int(n_floats), int(root)))

    def gather_tiles_device(self, d_accum_ptr, nx, ny, root=0):
        self._check(self.lib.crt_gather_tiles_device(self._h, C.c_void_p(d_accum_ptr), int(nx), int(ny), int(root)))

    def gather_device(self, d_send_ptr, nbytes, d_recv_ptr, root=0):
        self._check(self.lib.crt_gather_device(self._h, C.c_void_p(d_send_ptr), int(nbytes),
                                               C.c_void_p(d_recv_ptr) if d_recv_ptr else None, int(root)))

    def render_dist(self, cam22, nx, ny, spp, max_depth=8, mode=SHADE_PATH, seed=1, out=None):
        """crt_render_dist (collective): the frame split by tiles over the communicator's GPUs; rank 0 gets the image."""
        cam = np.ascontiguousarray(cam22, np.float32)
        rank = self.comm_info()["rank"]
        rgb = out
        if rank == 0 and rgb is None:
            rgb = np.empty((ny, nx, 3), np.float32)
        self._check(self.lib.crt_render_dist(self._h, _ptr(cam), nx, ny, spp, max_depth, int(mode), int(seed),
                                             _ptr(rgb) if rank == 0 else None))
        return rgb if rank == 0 else None

    # ---- test hooks ----
    def debug_sort_pairs(self, d_keys_ptr, d_vals_ptr, n, begin_bit=0, end_bit=64):
        self._check(self.lib.crt_debug_sort_pairs(self._h, C.c_void_p(d_keys_ptr), C.c_void_p(d_vals_ptr), int(n), begin_bit, end_bit))

    def debug_read(self, which, dtype, count):
        """(array, root): `count` items of `dtype` from the internal buffer `which` (DBG_*)."""
        out = np.zeros(count, dtype)
        root = C.c_int(0)
        self._check(self.lib.crt_debug_read(self._h, int(which), _ptr(out), out.nbytes, C.byref(root)))
        return out, root.value

    def sync(self):
        self._check(self.lib.crt_sync(self._h))

    def stream(self):
        return self.lib.crt_stream(self._h)

    def counters(self):
        c = Counters()
        self._check(self.lib.crt_get_counters(self._h, C.byref(c)))
        return c.as_dict()

    def reset_counters(self):
        self._check(self.lib.crt_reset_counters(self._h))
