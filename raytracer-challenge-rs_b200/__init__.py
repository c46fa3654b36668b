"""raytracer-challenge-rs_b200 — B200-native (sm_100a) render backend for the per-pixel
ray-cast/shade hot path of tlinford/raytracer-challenge-rs (`Camera::render(&World) -> Canvas`).

Layout
  csrc/   hand-written CUDA kernels + the C-ABI of include/rtc.h  → csrc/librtc_cuda.so
  host/   C++ mirror of the reference's scene API, flattener, OBJ front end → host/librth_host.so
  api.py / transform.py / scenes.py   ctypes binding and scene scripts (harness side)

The directory name contains '-', so import it with
    importlib.import_module("raytracer-challenge-rs_b200")
There is no CPU fallback: `render()` raises if the CUDA library is missing or no GPU is visible.
"""
import ctypes as C
import os

import numpy as np

from . import api as _api
from . import transform  # noqa: F401  (re-export)
from .api import (CSG, DIFFERENCE, INTERSECTION, UNION, Material, Pattern, SceneApi,  # noqa: F401
                  checkers_pattern, gradient_pattern, ring_pattern, stripe_pattern)

_HERE = os.path.dirname(os.path.abspath(__file__))
# RTC_CUDA_LIB: kernel-variant experiments only (same ABI, different build flags)
CUDA_LIB = os.environ.get("RTC_CUDA_LIB") or os.path.join(_HERE, "csrc", "librtc_cuda.so")
HOST_LIB = os.path.join(_HERE, "host", "librth_host.so")

RTC_OPT_FAITHFUL_TRAVERSAL = 1
RTC_OPT_AVERAGE = 2
RTC_OPT_WAVEFRONT = 4
RTC_OPT_WAVE_STREAM = 16
RTC_STAT_WAVEFRONT = 1
RTC_STAT_WAVE_FALLBACK = 2
RTC_STAT_WAVE_OVERFLOWED = 4
RTC_OPT_RECURSIVE = 8


class RtcNode(C.Structure):
    _fields_ = [("kind", C.c_int32), ("skip", C.c_int32), ("xform", C.c_int32), ("material", C.c_int32),
                ("aux", C.c_int32), ("flags", C.c_uint32), ("right", C.c_int32), ("eq_class", C.c_int32)]


class RtcSceneDesc(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("n_nodes", C.c_uint32), ("n_xforms", C.c_uint32),
                ("n_boxes", C.c_uint32), ("n_limits", C.c_uint32), ("n_triangles", C.c_uint32),
                ("n_materials", C.c_uint32), ("n_patterns", C.c_uint32), ("n_lights", C.c_uint32),
                ("nodes", C.c_void_p), ("xforms", C.c_void_p), ("boxes", C.c_void_p), ("limits", C.c_void_p),
                ("tri_p1e1e2", C.c_void_p), ("tri_normals", C.c_void_p), ("materials", C.c_void_p),
                ("patterns", C.c_void_p), ("lights", C.c_void_p)]


class RtcCamera(C.Structure):
    _fields_ = [("hsize", C.c_uint32), ("vsize", C.c_uint32), ("pixel_size", C.c_double),
                ("half_width", C.c_double), ("half_height", C.c_double), ("inverse", C.c_double * 16)]


class RtcOpts(C.Structure):
    _fields_ = [("aa_samples", C.c_uint32), ("max_depth", C.c_uint32), ("row_begin", C.c_uint32),
                ("row_end", C.c_uint32), ("tile_rows", C.c_uint32), ("rank", C.c_uint32),
                ("nranks", C.c_uint32), ("flags", C.c_uint32)]


class RtcStats(C.Structure):
    _fields_ = [("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_reflect", C.c_uint64),
                ("rays_refract", C.c_uint64), ("overflow_count", C.c_uint64), ("nan_count", C.c_uint64),
                ("kernel_ms", C.c_double), ("total_ms", C.c_double), ("launches", C.c_uint32), ("flags", C.c_uint32)]

    @property
    def rays(self):
        return self.rays_primary + self.rays_shadow + self.rays_reflect + self.rays_refract

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "_pad"}

    @property
    def wave_fallback(self):
        return bool(self.flags & RTC_STAT_WAVE_FALLBACK)

    @property
    def wave_overflowed(self):
        return bool(self.flags & RTC_STAT_WAVE_OVERFLOWED)


class RtcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("rtc error %d: %s" % (code, msg))
        self.code = code


_libs = {}


def cuda_lib():
    """librtc_cuda.so (the C-ABI boundary). Raises if it has not been built."""
    if "cuda" not in _libs:
        if not os.path.exists(CUDA_LIB):
            raise RuntimeError("CUDA extension missing: %s — run __graft_entry__.build() (nvcc, sm_100a). "
                               "There is no CPU fallback." % CUDA_LIB)
        lib = C.CDLL(CUDA_LIB, mode=C.RTLD_GLOBAL)
        vp, pd, u32, sz = C.c_void_p, C.c_void_p, C.c_uint32, C.c_size_t
        lib.rtc_last_error.restype = C.c_char_p
        lib.rtc_device_count.restype = C.c_int
        lib.rtc_scene_create.argtypes = [C.POINTER(RtcSceneDesc), C.c_int, C.POINTER(vp)]
        lib.rtc_scene_destroy.argtypes = [vp]
        lib.rtc_scene_destroy.restype = None
        lib.rtc_render.argtypes = [vp, C.POINTER(RtcCamera), C.POINTER(RtcOpts), pd, C.POINTER(RtcStats)]
        lib.rtc_render_device.argtypes = [vp, C.POINTER(RtcCamera), C.POINTER(RtcOpts), pd, vp, C.POINTER(RtcStats)]
        lib.rtc_scene_finish.argtypes = [vp, vp, C.POINTER(RtcStats)]
        lib.rtc_quantize_device.argtypes = [C.c_int, pd, sz, pd, vp]
        lib.rtc_render_rgb8.argtypes = [vp, C.POINTER(RtcCamera), C.POINTER(RtcOpts), pd, C.POINTER(RtcStats)]
        lib.rtc_trace_rays.argtypes = [vp, sz, pd, pd, u32, pd, pd]
        lib.rtc_trace_rays_static.argtypes = [vp, sz, pd, pd, u32, pd, pd]
        lib.rtc_probe_hit.argtypes = [vp, pd, pd, pd]
        lib.rtc_is_shadowed.argtypes = [vp, pd, u32]
        lib.rtc_canvas_create.argtypes = [C.c_int, sz, C.POINTER(vp)]
        lib.rtc_canvas_export.argtypes = [vp, C.c_char_p]
        lib.rtc_canvas_open.argtypes = [C.c_int, C.c_char_p, sz, C.POINTER(vp)]
        lib.rtc_canvas_ptr.argtypes = [vp]
        lib.rtc_canvas_ptr.restype = vp
        lib.rtc_canvas_download.argtypes = [vp, pd, sz, sz]
        lib.rtc_canvas_push_tiles.argtypes = [vp, vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, vp]
        lib.rtc_canvas_destroy.argtypes = [vp]
        lib.rtc_canvas_destroy.restype = None
        lib.rtc_host_canvas_create.argtypes = [C.c_char_p, sz, C.c_int, C.c_int, C.POINTER(vp)]
        lib.rtc_host_canvas_place.argtypes = [vp, u32, u32, u32, u32, u32]
        lib.rtc_host_canvas_ptr.argtypes = [vp]
        lib.rtc_host_canvas_ptr.restype = vp
        lib.rtc_host_canvas_arrive.argtypes = [vp, u32, u32]
        lib.rtc_host_canvas_wait.argtypes = [vp, u32, u32, C.c_double]
        lib.rtc_host_canvas_destroy.argtypes = [vp]
        lib.rtc_host_canvas_destroy.restype = None
        lib.rtc_measure_fp64.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double)]
        lib.rtc_debug_normalize.argtypes = [C.c_int, sz, pd, pd]
        _libs["cuda"] = lib
    return _libs["cuda"]


def host_lib():
    """librth_host.so (scene API mirror + flattener)."""
    if "host" not in _libs:
        cuda_lib()  # dependency, loaded RTLD_GLOBAL first
        if not os.path.exists(HOST_LIB):
            raise RuntimeError("host library missing: %s — run __graft_entry__.build()" % HOST_LIB)
        lib = C.CDLL(HOST_LIB)
        vp = C.c_void_p
        lib.rth_flatten.argtypes = [vp]
        lib.rth_flatten.restype = vp
        lib.rth_flat_free.argtypes = [vp]
        lib.rth_flat_free.restype = None
        lib.rth_flat_desc.argtypes = [vp, C.POINTER(RtcSceneDesc)]
        lib.rth_flat_desc.restype = None
        lib.rth_flat_info.argtypes = [vp, C.POINTER(C.c_uint64)]
        lib.rth_flat_info.restype = None
        lib.rth_camera_to_rtc.argtypes = [vp, C.POINTER(RtcCamera)]
        lib.rth_camera_to_rtc.restype = None
        lib.rth_render.argtypes = [vp, vp, C.POINTER(RtcOpts), C.c_int, vp, C.POINTER(RtcStats), vp]
        lib.rth_quantize.argtypes = [vp, C.c_size_t, vp]
        lib.rth_quantize.restype = None
        lib.rth_canvas_to_ppm.argtypes = [vp, C.c_size_t, C.c_size_t, vp, C.c_size_t]
        lib.rth_canvas_to_ppm.restype = C.c_size_t
        lib.rth_save_ppm.argtypes = [vp, C.c_size_t, C.c_size_t, C.c_char_p]
        lib.rth_save_p6.argtypes = [vp, C.c_size_t, C.c_size_t, C.c_char_p]
        lib.rth_save_png.argtypes = [vp, C.c_size_t, C.c_size_t, C.c_char_p]
        lib.rth_render_scene_file.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32, C.c_uint32, C.c_int, C.POINTER(RtcStats)]
        _libs["host"] = lib
    return _libs["host"]


_scene_api = None


def scene_api():
    """The product's scene-building API (`rth_*`)."""
    global _scene_api
    if _scene_api is None:
        _scene_api = SceneApi(host_lib(), "rth_")
    return _scene_api


def _check(rc):
    if rc < 0:
        raise RtcError(rc, (cuda_lib().rtc_last_error() or b"").decode())
    return rc


def make_opts(aa=1, max_depth=5, rows=None, tile_rows=8, rank=0, nranks=1, faithful=False, average=False, wavefront=None,
              wave_stream=False):
    o = RtcOpts()
    o.aa_samples, o.max_depth = aa, max_depth
    o.row_begin, o.row_end = (rows if rows else (0, 0))
    o.tile_rows, o.rank, o.nranks = tile_rows, rank, nranks
    o.flags = (RTC_OPT_FAITHFUL_TRAVERSAL if faithful else 0) | (RTC_OPT_AVERAGE if average else 0)
    if wavefront is not None:  # None: the library's default; True/False force one renderer (same bits either way)
        o.flags |= RTC_OPT_WAVEFRONT if wavefront else RTC_OPT_RECURSIVE
    if wave_stream:    # one barrier-free persistent launch instead of the level-synchronous passes (experimental)
        o.flags |= RTC_OPT_WAVE_STREAM
    return o


class FlatScene:
    """Flattened world (host SoA tables). `desc` is the rtc_scene_desc to hand to rtc_scene_create."""

    def __init__(self, world):
        self._world = world  # keeps the handles alive
        self.h = host_lib().rth_flatten(world.h)
        if not self.h:
            raise RuntimeError("flatten failed: " + scene_api().last_error())
        self.desc = RtcSceneDesc()
        host_lib().rth_flat_desc(self.h, C.byref(self.desc))
        info = (C.c_uint64 * 8)()
        host_lib().rth_flat_info(self.h, info)
        self.info = dict(zip(["nodes", "leaves", "groups", "csgs", "xforms", "materials", "triangles", "bytes"], list(info)))

    def nodes(self):
        arr = (RtcNode * self.desc.n_nodes).from_address(self.desc.nodes)
        return [{k: getattr(n, k) for k, _ in RtcNode._fields_} for n in arr]

    def __del__(self):
        if getattr(self, "h", None):
            try:
                host_lib().rth_flat_free(self.h)
            except TypeError:  # interpreter shutdown: module globals already cleared
                pass
            self.h = None


class DeviceScene:
    """A flattened world resident in HBM (rtc_scene)."""

    def __init__(self, flat, device=0):
        self.flat = flat
        self.device = device
        h = C.c_void_p()
        _check(cuda_lib().rtc_scene_create(C.byref(flat.desc), device, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            try:
                cuda_lib().rtc_scene_destroy(self.h)
            except TypeError:  # interpreter shutdown
                pass
            self.h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def render(self, camera, out=None, **opts):
        """rtc_render into a host canvas (numpy float64 [vsize, hsize, 3]); returns (canvas, stats)."""
        cam = camera_to_rtc(camera)
        if out is None:
            out = np.zeros((cam.vsize, cam.hsize, 3), dtype=np.float64)
        o = make_opts(**opts)
        st = RtcStats()
        _check(cuda_lib().rtc_render(self.h, C.byref(cam), C.byref(o), out.ctypes.data, C.byref(st)))
        return out, st

    def render_rgb8(self, camera, out=None, **opts):
        """rtc_render_rgb8: the frame as 8-bit pixels (uint8 [vsize, hsize, 3]), quantised on the device."""
        cam = camera_to_rtc(camera)
        if out is None:
            out = np.zeros((cam.vsize, cam.hsize, 3), dtype=np.uint8)
        o = make_opts(**opts)
        st = RtcStats()
        _check(cuda_lib().rtc_render_rgb8(self.h, C.byref(cam), C.byref(o), out.ctypes.data, C.byref(st)))
        return out, st

    def render_device(self, camera, dev_ptr, stream=None, stats=False, **opts):
        cam = camera_to_rtc(camera)
        o = make_opts(**opts)
        st = RtcStats() if stats else None
        _check(cuda_lib().rtc_render_device(self.h, C.byref(cam), C.byref(o), dev_ptr, stream,
                                            C.byref(st) if stats else None))
        return st

    def finish(self, stream=None):
        st = RtcStats()
        _check(cuda_lib().rtc_scene_finish(self.h, stream, C.byref(st)))
        return st

    def trace_rays(self, origins, directions, remaining=5, static=False):
        """World::color_at for a ray batch; static=True goes through rtc_trace_rays_static (no allocation)."""
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(directions, dtype=np.float64).reshape(-1, 3)
        rgb = np.zeros_like(o)
        t = np.zeros(o.shape[0], dtype=np.float64)
        fn = cuda_lib().rtc_trace_rays_static if static else cuda_lib().rtc_trace_rays
        _check(fn(self.h, o.shape[0], o.ctypes.data, d.ctypes.data, remaining, rgb.ctypes.data, t.ctypes.data))
        return rgb, t

    def probe_hit(self, origin, direction):
        out = np.zeros(20, dtype=np.float64)
        o = np.asarray(origin, dtype=np.float64)
        d = np.asarray(direction, dtype=np.float64)
        _check(cuda_lib().rtc_probe_hit(self.h, o.ctypes.data, d.ctypes.data, out.ctypes.data))
        return out

    def is_shadowed(self, point, light_index=0):
        p = np.asarray(point, dtype=np.float64)
        return bool(_check(cuda_lib().rtc_is_shadowed(self.h, p.ctypes.data, light_index)))


def camera_to_rtc(camera):
    if isinstance(camera, RtcCamera):
        return camera
    cam = RtcCamera()
    host_lib().rth_camera_to_rtc(camera.h, C.byref(cam))
    return cam


def render(camera, world, device=0, aa=None, **opts):
    """`Camera::render(&World) -> Canvas` on the GPU through the host library's drop-in call
    (flatten → rtc_scene_create → rtc_render → destroy). Returns (canvas[v,h,3] float64, stats, times_ms)."""
    lib = host_lib()
    o = make_opts(aa=aa or camera.aa_samples, **opts)
    out = np.zeros((camera.vsize, camera.hsize, 3), dtype=np.float64)
    st = RtcStats()
    times = (C.c_double * 4)()
    rc = lib.rth_render(camera.h, world.h, C.byref(o), device, out.ctypes.data, C.byref(st), times)
    if rc != 0:
        raise RtcError(rc, scene_api().last_error())
    return out, st, dict(zip(["flatten", "upload", "render", "total"], list(times)))


def quantize(canvas):
    """image/ppm.rs:73-75 on the host: (v*255).round() as u8."""
    c = np.ascontiguousarray(canvas, dtype=np.float64)
    out = np.zeros(c.shape, dtype=np.uint8)
    host_lib().rth_quantize(c.ctypes.data, c.size, out.ctypes.data)
    return out


def canvas_to_ppm(canvas):
    c = np.ascontiguousarray(canvas, dtype=np.float64)
    h, w = c.shape[0], c.shape[1]
    n = host_lib().rth_canvas_to_ppm(c.ctypes.data, w, h, None, 0)
    buf = C.create_string_buffer(n)
    host_lib().rth_canvas_to_ppm(c.ctypes.data, w, h, buf, n)
    return buf.raw


def save_image(rgb8, path):
    """Write 8-bit pixels (uint8 [h, w, 3]) as .png (image/png.rs:13-31) or binary P6."""
    img = np.ascontiguousarray(rgb8, dtype=np.uint8)
    fn = host_lib().rth_save_png if str(path).endswith(".png") else host_lib().rth_save_p6
    if fn(img.ctypes.data, img.shape[1], img.shape[0], os.fsencode(str(path))) != 0:
        raise OSError(scene_api().last_error())


def render_scene_file(scene_path, out_path, extended=False, aa=1, device=0):
    """scene-parser's `render_scene <scene-file> <output-file>` on the CUDA backend; returns RtcStats."""
    st = RtcStats()
    rc = host_lib().rth_render_scene_file(os.fsencode(str(scene_path)), os.fsencode(str(out_path)),
                                          1 if extended else 0, aa, device, C.byref(st))
    if rc != 0:
        raise RtcError(rc, scene_api().last_error()) if rc < 0 else OSError(scene_api().last_error())
    return st


def measure_fp64(device=0, fused=False):
    g = C.c_double()
    _check(cuda_lib().rtc_measure_fp64(device, 1 if fused else 0, C.byref(g)))
    return g.value
