"""oracle — TEST INFRASTRUCTURE (not part of the product).

Python bindings for the two CPU checkers of the render hot path:

* ``load("port")``: ``oracle/libcrt_oracle.so``, the plain-C restatement
  (``oracle/crt_oracle.c``);
* ``load("ref")``: ``oracle/_ref/libcrt_ref.so``, the UNMODIFIED reference
  headers from ``/root/reference/CudaTest/src`` compiled host-side through
  ``oracle/shim`` (``oracle/ref_harness.cpp``). Built in the authoring
  container only; the prebuilt ``.so`` travels to the GPU box.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this package.
``cudaraytracer_b200`` never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_SRC = "/root/reference/CudaTest/src"

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def build(ref=True, port=True, verbose=False):
    """Compile the checkers (building the checker is not using it)."""
    targets = []
    if port:
        targets.append("port")
    if ref and os.path.isdir(REFERENCE_SRC):
        targets.append("ref")
    if targets:
        out = subprocess.run(["make", "-C", _HERE] + targets, capture_output=True, text=True)
        if verbose or out.returncode != 0:
            print(out.stdout, out.stderr)
        out.check_returncode()


def lib_path(kind):
    return os.path.join(_HERE, "libcrt_oracle.so") if kind == "port" else os.path.join(_HERE, "_ref", "libcrt_ref.so")


def available(kind):
    return os.path.exists(lib_path(kind))


def _opt(arr, dtype):
    if arr is None:
        return None
    a = np.ascontiguousarray(arr, dtype=dtype)
    return a


class Scene:
    """A scene held by one of the checkers (list order = ascending prim id)."""

    def __init__(self, lib, d1=True):
        self.lib = lib
        self.h = lib._fn("scene_create")(1 if d1 else 0)
        self.n_prims = 0

    def add_material(self, mtype, albedo=(0, 0, 0), param=0.0, odd=None):
        """odd given: CheckerTexture(even = albedo, odd) under a Lambertian / DiffuseLight (material/texture.h:25-42)."""
        a = np.asarray(albedo, dtype=np.float32)
        if odd is not None:
            r = self.lib._fn("scene_add_checker_material")(self.h, int(mtype), a, np.asarray(odd, dtype=np.float32))
            assert r >= 0
            return r
        r = self.lib._fn("scene_add_material")(self.h, int(mtype), a, float(param))
        assert r >= 0
        return r

    def add_triangles(self, tri, mat=None, ids=None):
        tri = np.ascontiguousarray(tri, dtype=np.float32).reshape(-1, 12)
        n = tri.shape[0]
        mat = _opt(mat, np.uint32)
        ids = _opt(ids, np.uint32)
        self.lib._fn("scene_add_triangles")(
            self.h, tri, None if mat is None else mat.ctypes.data_as(C.c_void_p),
            None if ids is None else ids.ctypes.data_as(C.c_void_p), n)
        self.n_prims += n

    def add_spheres(self, sph, mat=None, ids=None):
        sph = np.ascontiguousarray(sph, dtype=np.float32).reshape(-1, 4)
        n = sph.shape[0]
        mat = _opt(mat, np.uint32)
        ids = _opt(ids, np.uint32)
        self.lib._fn("scene_add_spheres")(
            self.h, sph, None if mat is None else mat.ctypes.data_as(C.c_void_p),
            None if ids is None else ids.ctypes.data_as(C.c_void_p), n)
        self.n_prims += n

    def add_chained(self, kind, data, mat, ids, levels, flip=None):
        """n primitives under the chain of Transforms `levels` = [(position, rotation_degrees), ...], outermost first,
        the last being the primitives' own (ref_scene_add_chained). kind 0: data (n, 12) triangles in their own
        frame; 1: data (n,) radii; 2: rectangles (shapes/rectangle.h), flip (n,) = flipNormal."""
        ids = np.ascontiguousarray(ids, np.uint32)
        n = len(ids)
        mat = np.ascontiguousarray(mat, np.uint32)
        lv = np.ascontiguousarray(np.asarray([list(p) + list(r) for p, r in levels], np.float32).reshape(-1, 6))
        d = np.zeros(1, np.float32) if data is None else np.ascontiguousarray(data, np.float32).reshape(-1)
        fl = np.zeros(n, np.uint8) if flip is None else np.ascontiguousarray(flip, np.uint8)
        fn = self.lib._fn("scene_add_chained")
        fn.restype = C.c_int
        fn.argtypes = [C.c_void_p, C.c_int, _f32p, _u32p, _u32p, _u8p, C.c_size_t, _f32p, C.c_int]
        r = fn(self.h, int(kind), d, mat, ids, fl, n, lv, len(lv))
        assert r == 0
        self.n_prims += n

    def finalize(self):
        r = self.lib._fn("scene_finalize")(self.h)
        assert r == 0, "scene_finalize failed (bad material index?)"
        return self

    def bvh_build(self, seed=1):
        """Reference BVHNode build (ref only). Returns seconds."""
        return self.lib._fn("bvh_build")(self.h, int(seed))

    def use_bvh(self, flag=True):
        """ref only: ref_render goes through the reference's BVHNode (after bvh_build) instead of the list scan."""
        return bool(self.lib._fn("scene_use_bvh")(self.h, 1 if flag else 0))

    def bvh_stats(self, rays, t_min=0.001, t_max=3.4028234663852886e38):
        """ref only: (box tests per ray, primitives handed to leaf scans per ray) of the reference's BVHNode walk."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        a, b = C.c_uint64(0), C.c_uint64(0)
        r = self.lib._fn("bvh_stats")(self.h, rays, rays.shape[0], float(t_min), float(t_max), C.byref(a), C.byref(b))
        assert r == 0, "bvh_stats needs bvh_build first"
        n = max(1, rays.shape[0])
        return a.value / n, b.value / n

    def trace(self, rays, t_min=0.001, t_max=3.4028234663852886e38, use_bvh=False, want_pn=False):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = rays.shape[0]
        t = np.zeros(n, np.float32)
        pid = np.zeros(n, np.uint32)
        p = np.zeros((n, 3), np.float32) if want_pn else None
        nn = np.zeros((n, 3), np.float32) if want_pn else None
        secs = self.lib._fn("trace")(
            self.h, rays, n, float(t_min), float(t_max), 1 if use_bvh else 0, t, pid,
            None if p is None else p.ctypes.data_as(C.c_void_p),
            None if nn is None else nn.ctypes.data_as(C.c_void_p))
        assert secs >= 0
        self.last_seconds = secs
        if want_pn:
            return t, pid, p, nn
        return t, pid

    def render(self, cam22, nx, ny, spp, max_depth=8, mode=0, seed=1, y0=0, y1=None, y_step=1,
               want_accum=False):
        cam22 = np.ascontiguousarray(cam22, dtype=np.float32)
        assert cam22.size == 22
        y1 = ny if y1 is None else y1
        rgb = np.zeros((ny, nx, 3), np.float32)
        acc = np.zeros((ny, nx, 3), np.float32) if want_accum else None
        rays = C.c_uint64(0)
        secs = self.lib._fn("render")(
            self.h, cam22, nx, ny, spp, max_depth, mode, int(seed), y0, y1, y_step, rgb,
            None if acc is None else acc.ctypes.data_as(C.c_void_p), C.byref(rays))
        assert secs >= 0
        self.last_seconds = secs
        self.last_rays = rays.value
        if want_accum:
            return rgb, acc
        return rgb


class OracleLib:
    def __init__(self, kind):
        assert kind in ("port", "ref")
        self.kind = kind
        self.prefix = "oracle_" if kind == "port" else "ref_"
        self.dll = C.CDLL(lib_path(kind))
        self._declare()

    def _fn(self, name):
        return getattr(self.dll, self.prefix + name)

    def _declare(self):
        vp = C.c_void_p
        f = self._fn
        f("num_threads").restype = C.c_int
        f("scene_create").restype = vp
        f("scene_create").argtypes = [C.c_int]
        f("scene_add_material").restype = C.c_int
        f("scene_add_material").argtypes = [vp, C.c_uint32, _f32p, C.c_float]
        f("scene_add_checker_material").restype = C.c_int
        f("scene_add_checker_material").argtypes = [vp, C.c_uint32, _f32p, _f32p]
        f("scene_add_triangles").restype = None
        f("scene_add_triangles").argtypes = [vp, _f32p, vp, vp, C.c_size_t]
        f("scene_add_spheres").restype = None
        f("scene_add_spheres").argtypes = [vp, _f32p, vp, vp, C.c_size_t]
        f("scene_finalize").restype = C.c_int
        f("scene_finalize").argtypes = [vp]
        f("trace").restype = C.c_double
        f("trace").argtypes = [vp, _f32p, C.c_size_t, C.c_float, C.c_float, C.c_int, _f32p, _u32p, vp, vp]
        f("scatter").restype = C.c_int
        f("scatter").argtypes = [C.c_uint32, _f32p, C.c_float, _f32p, _f32p, _f32p, C.c_int, _f32p,
                                 C.POINTER(C.c_int)]
        f("camera").restype = None
        f("camera").argtypes = [_f32p, _f32p, _f32p, C.c_float, C.c_float, C.c_float, C.c_float, _f32p]
        f("get_ray").restype = C.c_uint32
        f("get_ray").argtypes = [_f32p, C.c_float, C.c_float, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, _f32p]
        f("render").restype = C.c_double
        f("render").argtypes = [vp, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int,
                                C.c_int, C.c_int, _f32p, vp, C.POINTER(C.c_uint64)]
        f("pack_rgba8").restype = None
        f("pack_rgba8").argtypes = [_f32p, C.c_int, C.c_int, _u8p]
        f("parse_obj").restype = C.c_int
        f("parse_obj").argtypes = [C.c_char_p, _f32p, _f32p, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        f("triangle_normal").restype = None
        f("triangle_normal").argtypes = [_f32p, _f32p]
        if self.kind == "ref":
            f("bvh_build").restype = C.c_double
            f("bvh_build").argtypes = [vp, C.c_uint64]
            f("bvh_refit").restype = C.c_double
            f("bvh_refit").argtypes = [vp]
            f("scene_use_bvh").restype = C.c_int
            f("scene_use_bvh").argtypes = [vp, C.c_int]
            f("bvh_stats").restype = C.c_int
            f("bvh_stats").argtypes = [vp, _f32p, C.c_size_t, C.c_float, C.c_float, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]

    # ---- functions ----
    def num_threads(self):
        return self._fn("num_threads")()

    def scene(self, d1=True):
        return Scene(self, d1)

    def checker_odd(self, points):
        """CheckerTexture::value at n points: True where the odd texture is returned (material/texture.h:30-38)."""
        pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
        out = np.zeros(len(pts), np.uint8)
        fn = self._fn("checker_odd")
        fn.restype = None
        fn.argtypes = [_f32p, C.c_size_t, np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")]
        fn(pts, len(pts), out)
        return out.astype(bool)

    def rotation_rows(self, rot):
        """The three rows of the matrix rotate() builds for Euler angles `rot` in degrees (core/vec3.h:200-214)."""
        out = np.zeros(9, np.float32)
        fn = self._fn("rotation_rows")
        fn.restype = None
        fn.argtypes = [_f32p, _f32p]
        fn(np.ascontiguousarray(rot, np.float32), out)
        return out

    def scatter(self, mtype, albedo, param, ray, p, normal, randoms):
        """One Material::scatter call with an injected (0,1] sequence.
        Returns (ok, attenuation[3], origin[3], direction[3], n_draws)."""
        albedo = np.asarray(albedo, np.float32)
        ray = np.ascontiguousarray(ray, np.float32).reshape(6)
        recpn = np.concatenate([np.asarray(p, np.float32), np.asarray(normal, np.float32)]).astype(np.float32)
        randoms = np.ascontiguousarray(randoms, np.float32)
        out = np.zeros(9, np.float32)
        nd = C.c_int(0)
        ok = self._fn("scatter")(int(mtype), albedo, float(param), ray, recpn, randoms, randoms.size, out, C.byref(nd))
        assert ok >= 0
        return bool(ok), out[0:3].copy(), out[3:6].copy(), out[6:9].copy(), nd.value

    def camera(self, lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist):
        out = np.zeros(22, np.float32)
        self._fn("camera")(np.asarray(lookfrom, np.float32), np.asarray(lookat, np.float32),
                           np.asarray(vup, np.float32), float(vfov), float(aspect), float(aperture),
                           float(focus_dist), out)
        return out

    def get_ray(self, cam22, s, t, seed, pixel, sample, ctr=0):
        out = np.zeros(6, np.float32)
        c = self._fn("get_ray")(np.ascontiguousarray(cam22, np.float32), float(s), float(t), int(seed),
                                int(pixel), int(sample), int(ctr), out)
        return out, c

    def pack_rgba8(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.float32)
        ny, nx = rgb.shape[0], rgb.shape[1]
        out = np.zeros((ny, nx, 4), np.uint8)
        self._fn("pack_rgba8")(rgb, nx, ny, out)
        return out

    def parse_obj(self, path, max_points=1 << 22, max_tris=1 << 23):
        # size the buffers from the file so small files stay cheap
        nv = nf = 0
        with open(path, "r", errors="ignore") as fh:
            for line in fh:
                if line.startswith("v "):
                    nv += 1
                elif line.startswith("f "):
                    nf += 1
        pts = np.zeros((max(nv, 1), 3), np.float32)
        idx = np.zeros((max(nf, 1), 3), np.float32)
        np_ = C.c_int(0)
        nt_ = C.c_int(0)
        r = self._fn("parse_obj")(path.encode(), pts, idx, pts.shape[0], idx.shape[0], C.byref(np_), C.byref(nt_))
        if r != 0:
            raise IOError("parse_obj(%s) failed: %d" % (path, r))
        return pts[: np_.value].copy(), idx[: nt_.value].copy()

    def triangle_normal(self, v9):
        v9 = np.ascontiguousarray(v9, np.float32).reshape(9)
        out = np.zeros(3, np.float32)
        self._fn("triangle_normal")(v9, out)
        return out


_cache = {}


def load(kind):
    if kind not in _cache:
        if not available(kind):
            raise FileNotFoundError("%s is not built; run `make -C oracle %s`" % (lib_path(kind), kind))
        _cache[kind] = OracleLib(kind)
    return _cache[kind]
