"""Shared fixtures. `-m "not gpu"` runs here (no GPU); `-m gpu` runs on a B200 box.

pkg      the product package (raytracer-challenge-rs_b200), imported by path
prod     product scene API (rth_* of host/librth_host.so)
orc      oracle scene API (orc_* of oracle/liboracle.so)  — the checker, tests only
emul     tests/cpu_emul/libemul.so — host compile of the kernel source, CPU-only test tier
"""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG_NAME = "raytracer-challenge-rs_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: longer CPU test")


def _ensure_built():
    """Build the test-side libraries if they are missing (the product libs are built by build())."""
    need = [os.path.join(ROOT, "oracle", "liboracle.so"), os.path.join(ROOT, "oracle", "kat"),
            os.path.join(ROOT, "tests", "cpu_emul", "libemul.so"),
            os.path.join(ROOT, PKG_NAME, "csrc", "librtc_cuda.so"), os.path.join(ROOT, PKG_NAME, "host", "librth_host.so")]
    if all(os.path.exists(p) for p in need):
        return
    subprocess.run([sys.executable, os.path.join(ROOT, PKG_NAME, "build.py")], check=True)


@pytest.fixture(scope="session")
def pkg():
    _ensure_built()
    return importlib.import_module(PKG_NAME)


@pytest.fixture(scope="session")
def prod(pkg):
    return pkg.scene_api()


class Oracle:
    """ctypes view of oracle/liboracle.so: scene API + render/probe entry points."""

    def __init__(self, pkg):
        self.lib = C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
        self.api = pkg.SceneApi(self.lib, "orc_")
        L = self.lib
        vp = C.c_void_p
        L.orc_counters_len.restype = C.c_size_t
        L.orc_counters_flops.restype = C.c_double
        L.orc_counters_flops.argtypes = [vp]
        L.orc_render.argtypes = [vp, vp, C.c_int, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp]
        L.orc_trace_rays.argtypes = [vp, C.c_size_t, vp, vp, C.c_int, vp, vp]
        L.orc_probe_hit.argtypes = [vp, vp, vp, vp]
        L.orc_is_shadowed.argtypes = [vp, vp, C.c_size_t]
        L.orc_canvas_to_ppm.argtypes = [vp, C.c_size_t, C.c_size_t, vp, C.c_size_t]
        L.orc_canvas_to_ppm.restype = C.c_size_t
        L.orc_quantize.argtypes = [vp, C.c_size_t, vp]
        self.ncount = L.orc_counters_len()

    def render(self, camera, world, aa=1, threads=None, rows=None, out=None):
        """threads=0 → Camera::render; threads>=1 → render_multithreaded semantics."""
        if threads is None:
            threads = os.cpu_count() or 1
        h, w = camera.vsize, camera.hsize
        if out is None:
            out = np.zeros((h, w, 3), dtype=np.float64)
        r0, r1 = rows if rows else (0, h)
        stats = np.zeros(self.ncount, dtype=np.uint64)
        rc = self.lib.orc_render(camera.h, world.h, aa, threads, r0, r1, out.ctypes.data, stats.ctypes.data)
        if rc != 0:
            raise RuntimeError("oracle: " + self.api.last_error())
        names = ["rays_primary", "rays_shadow", "rays_reflect", "rays_refract"]
        info = {n: int(stats[i]) for i, n in enumerate(names)}
        info["rays"] = sum(info[n] for n in names)
        info["flops"] = self.lib.orc_counters_flops(stats.ctypes.data)
        info["ms"] = float(stats[-1]) / 1000.0
        return out, info

    def trace_rays(self, world, origins, directions, remaining=5):
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        rgb = np.zeros_like(o)
        t = np.zeros(o.shape[0])
        rc = self.lib.orc_trace_rays(world.h, o.shape[0], o.ctypes.data, d.ctypes.data, remaining, rgb.ctypes.data, t.ctypes.data)
        if rc != 0:
            raise RuntimeError("oracle: " + self.api.last_error())
        return rgb, t

    def probe_hit(self, world, origin, direction):
        out = np.zeros(20)
        o, d = np.asarray(origin, dtype=np.float64), np.asarray(direction, dtype=np.float64)
        assert self.lib.orc_probe_hit(world.h, o.ctypes.data, d.ctypes.data, out.ctypes.data) == 0
        return out

    def is_shadowed(self, world, point, light=0):
        p = np.asarray(point, dtype=np.float64)
        return bool(self.lib.orc_is_shadowed(world.h, p.ctypes.data, light))

    def canvas_to_ppm(self, canvas):
        c = np.ascontiguousarray(canvas, dtype=np.float64)
        n = self.lib.orc_canvas_to_ppm(c.ctypes.data, c.shape[1], c.shape[0], None, 0)
        buf = C.create_string_buffer(n)
        self.lib.orc_canvas_to_ppm(c.ctypes.data, c.shape[1], c.shape[0], buf, n)
        return buf.raw

    def quantize(self, canvas):
        c = np.ascontiguousarray(canvas, dtype=np.float64)
        out = np.zeros(c.shape, dtype=np.uint8)
        self.lib.orc_quantize(c.ctypes.data, c.size, out.ctypes.data)
        return out


@pytest.fixture(scope="session")
def orc(pkg):
    return Oracle(pkg)


class Emul:
    """tests/cpu_emul/libemul.so — the kernel source compiled for the host (test-only)."""

    def __init__(self, pkg, path=None):
        self.pkg = pkg
        self.lib = C.CDLL(path or os.path.join(ROOT, "tests", "cpu_emul", "libemul.so"))
        vp = C.c_void_p
        self.lib.emul_render.argtypes = [vp, vp, vp, C.c_uint, vp, vp]
        self.lib.emul_trace_rays.argtypes = [vp, C.c_size_t, vp, vp, C.c_uint32, vp, vp]
        self.lib.emul_probe_hit.argtypes = [vp, vp, vp, vp]
        self.lib.emul_is_shadowed.argtypes = [vp, vp, C.c_uint32]

    def render(self, camera, world, flat=None, threads=None, **opts):
        pkg = self.pkg
        flat = flat or pkg.FlatScene(world)
        cam = pkg.camera_to_rtc(camera)
        o = pkg.make_opts(**opts)
        out = np.zeros((cam.vsize, cam.hsize, 3), dtype=np.float64)
        counters = np.zeros(6, dtype=np.uint64)
        rc = self.lib.emul_render(C.byref(flat.desc), C.byref(cam), C.byref(o), threads or (os.cpu_count() or 1),
                                  out.ctypes.data, counters.ctypes.data)
        assert rc == 0
        names = ["rays_primary", "rays_shadow", "rays_reflect", "rays_refract", "overflow", "nan"]
        info = {n: int(counters[i]) for i, n in enumerate(names)}
        info["rays"] = sum(info[n] for n in names[:4])
        return out, info

    def trace_rays(self, flat, origins, directions, remaining=5):
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        rgb = np.zeros_like(o)
        t = np.zeros(o.shape[0])
        err = self.lib.emul_trace_rays(C.byref(flat.desc), o.shape[0], o.ctypes.data, d.ctypes.data, remaining,
                                       rgb.ctypes.data, t.ctypes.data)
        assert err == 0, "device-logic error flags %d" % err
        return rgb, t

    def probe_hit(self, flat, origin, direction):
        out = np.zeros(20)
        o, d = np.asarray(origin, dtype=np.float64), np.asarray(direction, dtype=np.float64)
        assert self.lib.emul_probe_hit(C.byref(flat.desc), o.ctypes.data, d.ctypes.data, out.ctypes.data) == 0
        return out

    def is_shadowed(self, flat, point, light=0):
        p = np.asarray(point, dtype=np.float64)
        return bool(self.lib.emul_is_shadowed(C.byref(flat.desc), p.ctypes.data, light))


@pytest.fixture(scope="session")
def emul(pkg):
    return Emul(pkg)


def compare_images(got, want, quant):
    """The north-star parity rule: returns (fraction of identical 8-bit pixels, max 8-bit channel
    difference, max linear difference)."""
    g8, w8 = quant(got), quant(want)
    same = np.all(g8 == w8, axis=-1)
    max8 = int(np.max(np.abs(g8.astype(np.int32) - w8.astype(np.int32)))) if g8.size else 0
    lin = float(np.max(np.abs(got - want))) if got.size else 0.0
    return float(np.mean(same)) if same.size else 1.0, max8, lin


@pytest.fixture(scope="session")
def has_gpu(pkg):
    return pkg.cuda_lib().rtc_device_count() > 0
