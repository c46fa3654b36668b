"""The C-ABI library loads on a CPU-only machine and exports every symbol include/rtc.h declares;
compute entry points fail loudly (no CPU fallback) when there is no device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "rtc.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rtc_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(pkg):
    lib = pkg.cuda_lib()
    syms = declared_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(lib, s), "librtc_cuda.so does not export %s" % s


def test_abi_version(pkg):
    assert pkg.cuda_lib().rtc_abi_version() == 1


def test_struct_sizes_match_header(pkg):
    # layout pinned: rtc_node 32 B, rtc_material 88 B, rtc_pattern 152 B, rtc_light 48 B
    assert C.sizeof(pkg.RtcNode) == 32
    assert C.sizeof(pkg.RtcCamera) == 8 + 3 * 8 + 16 * 8
    assert C.sizeof(pkg.RtcOpts) == 32
    assert C.sizeof(pkg.RtcStats) == 6 * 8 + 2 * 8 + 8


def test_no_cpu_fallback_without_device(pkg, prod, has_gpu):
    if has_gpu:
        pytest.skip("a GPU is visible")
    scenes = __import__("importlib").import_module("raytracer-challenge-rs_b200.scenes")
    w, c = scenes.default_world(prod)
    with pytest.raises(pkg.RtcError) as e:
        pkg.render(c, w)
    assert e.value.code == -2  # RTC_ERR_CUDA
    flat = pkg.FlatScene(w)
    with pytest.raises(pkg.RtcError):
        pkg.DeviceScene(flat)


def test_scene_validation_rejects_bad_desc(pkg, prod):
    scenes = __import__("importlib").import_module("raytracer-challenge-rs_b200.scenes")
    w, c = scenes.default_world(prod)
    flat = pkg.FlatScene(w)
    nodes = (pkg.RtcNode * flat.desc.n_nodes).from_address(flat.desc.nodes)
    old = nodes[0].xform
    nodes[0].xform = 99
    h = C.c_void_p()
    rc = pkg.cuda_lib().rtc_scene_create(C.byref(flat.desc), 0, C.byref(h))
    nodes[0].xform = old
    assert rc == -1 and b"xform" in pkg.cuda_lib().rtc_last_error()


def test_scene_validation_rejects_null_tables(pkg, prod):
    """A non-empty table with a NULL pointer is RTC_ERR_INVALID before anything dereferences it (no device needed)."""
    scenes = __import__("importlib").import_module("raytracer-challenge-rs_b200.scenes")
    w, c = scenes.default_world(prod)
    flat = pkg.FlatScene(w)
    for field in ("materials", "xforms", "lights", "nodes"):
        d = pkg.RtcSceneDesc.from_buffer_copy(flat.desc)
        setattr(d, field, None)
        h = C.c_void_p()
        rc = pkg.cuda_lib().rtc_scene_create(C.byref(d), 0, C.byref(h))
        assert rc == -1, field
        assert b"NULL" in pkg.cuda_lib().rtc_last_error(), field
