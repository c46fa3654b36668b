"""rtc_trace_rays_static — the allocation-free variant of rtc_trace_rays (static __device__ staging, no counterpart in
the reference) — against the existing entry point, through the C-ABI, on two calls in a row. Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes as C
import importlib

import numpy as np
import pytest

scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")

STATIC_MAX = 4096  # RTC_TRACE_STATIC_MAX, include/rtc.h


def _batch(rng, n):
    o = rng.uniform(-3, 3, size=(n, 3)) + np.array([0, 2.0, -2.0])
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d


@pytest.mark.gpu
def test_trace_rays_static_matches_trace_rays_twice(pkg, prod):
    wp, _ = scenes.primitives_csg(prod, 16, 9)          # CSG kernel instantiation
    wg, _ = scenes.reflect_refract(prod, 16, 9)         # no-CSG instantiation, reflection + refraction
    rng = np.random.default_rng(20261105)
    for world in (wp, wg):
        ds = pkg.DeviceScene(pkg.FlatScene(world))
        # first call fills the whole array; the second is smaller, so stale results of the first sit right behind it
        for n in (STATIC_MAX, 1000, 1):
            o, d = _batch(rng, n)
            rgb0, t0 = ds.trace_rays(o, d)
            rgb1, t1 = ds.trace_rays(o, d, static=True)
            assert np.array_equal(t1, t0)
            assert np.array_equal(rgb1, rgb0)            # same kernel, same inputs: bit-identical
            assert np.any(t0 >= 0) or n == 1
        ds.close()


@pytest.mark.gpu
def test_trace_rays_static_rejects_oversize(pkg, prod):
    wp, _ = scenes.default_world(prod)
    ds = pkg.DeviceScene(pkg.FlatScene(wp))
    rng = np.random.default_rng(1)
    o, d = _batch(rng, STATIC_MAX + 1)
    with pytest.raises(pkg.RtcError) as e:
        ds.trace_rays(o, d, static=True)
    assert e.value.code == -1 and "RTC_TRACE_STATIC_MAX" in str(e.value)
    # and the array is still usable afterwards
    o, d = _batch(rng, 64)
    a = ds.trace_rays(o, d)
    b = ds.trace_rays(o, d, static=True)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    ds.close()


def test_static_max_matches_header():
    import os, re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    m = re.search(r"#define\s+RTC_TRACE_STATIC_MAX\s+(\d+)", open(os.path.join(root, "include", "rtc.h")).read())
    assert m and int(m.group(1)) == STATIC_MAX
