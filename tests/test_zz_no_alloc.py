"""crt_trace_static: the closest-hit call whose device scratch is a static array of the library (include/crt.h; no
counterpart in the reference). Checked through the C ABI against the
existing entry point, crt_trace, on two calls in a row - the second reuses the array the first one left dirty.

This file sorts last on purpose: a failure here hides no other test. It passes alone
in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes

pytestmark = pytest.mark.gpu


def _same(a, b, what):
    # the tolerance of the parity tests: ids, t and barycentrics bit for bit (same kernel, same arithmetic)
    assert np.array_equal(a[1], b[1]), "%s: %d primitive ids differ" % (what, int((a[1] != b[1]).sum()))
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3]), what + ": t / u / v differ"


def test_trace_static_matches_trace_on_two_calls_in_a_row():
    sc = scenes.c1_bunny()
    cam = crt.camera_for(sc, 256, 128)
    first = scenes.primary_rays(cam, 256, 128)                    # 32768 rays, most of them hits
    rng = np.random.default_rng(3)
    second = np.concatenate([rng.normal(size=(50000, 3)) * 12, rng.normal(size=(50000, 3))], axis=1).astype(np.float32)
    with crt.Context(0) as c:
        c.set_scene(sc)
        c.build()
        want1, want2 = c.trace(first), c.trace(second)
        assert (want1[1] != crt.MISS).sum() > 5000 and (want2[1] != crt.MISS).sum() > 1000
        got1 = c.trace_static(first)
        got2 = c.trace_static(second)                             # the array still holds call 1's cursor, rays and hits
        _same(got1, want1, "call 1")
        _same(got2, want2, "call 2")
        _same(c.trace_static(first[:7]), [w[:7] for w in want1], "a short third call")
        # the limit is stated and enforced, with a message
        assert crt.TRACE_STATIC_MAX_RAYS == 65536
        full = np.tile(second, (2, 1))[:crt.TRACE_STATIC_MAX_RAYS]
        _same(c.trace_static(full), c.trace(full), "a call of exactly the maximum size")
        with pytest.raises(crt.CrtError, match="at most 65536"):
            c.trace_static(np.tile(second, (2, 1))[:crt.TRACE_STATIC_MAX_RAYS + 1])
        # empty batch: nothing to do
        assert c.lib.crt_trace_static(c._h, None, 0, C.c_float(0.001), C.c_float(1e30), None) == 0


def test_trace_static_general_scene_and_state_errors():
    """Spheres take the kernel's general instantiation; without an acceleration structure the call says so."""
    sc = scenes.c2_spheres()
    rays = scenes.primary_rays(crt.camera_for(sc, 200, 120), 200, 120)
    with crt.Context(0) as c:
        c.set_scene(sc)
        with pytest.raises(crt.CrtError, match="crt_accel_build"):
            c.trace_static(rays)
        c.build()
        _same(c.trace_static(rays), c.trace(rays), "spheres")
        _same(c.trace_static(rays), c.trace(rays), "spheres, again")
