#!/usr/bin/env python
"""eighth_call.py — ONE GPU plays rank 0 of 8 on the headline frame through rtc_render into a pinned host canvas: where the
call's time goes beyond its kernels (RTC_TIMING=1 prints the library's own phases)."""
import ctypes as C, importlib, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("raytracer-challenge-rs_b200"); scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
api = pkg.scene_api()
world, cam = scenes.reflect_refract(api, width=1920, height=1080)
ds = pkg.DeviceScene(pkg.FlatScene(world), device=0)
host = torch.empty(cam.hsize * cam.vsize * 3, dtype=torch.float64).pin_memory()
cam_rtc = pkg.camera_to_rtc(cam)
lib = pkg.cuda_lib()
for nranks in (8, 4, 1):
    opts = pkg.make_opts(tile_rows=8, rank=0, nranks=nranks)
    stats = pkg.RtcStats()
    ts, ks = [], []
    for i in range(30):
        t = time.perf_counter()
        pkg._check(lib.rtc_render(ds.h, C.byref(cam_rtc), C.byref(opts), C.c_void_p(host.data_ptr()), C.byref(stats)))
        ts.append((time.perf_counter() - t) * 1e3)
        ks.append(stats.kernel_ms)
    print("rank 0 of %d: call %.3f ms (min %.3f), kernels %.3f ms, launches %d, flags %d" % (nranks, float(np.median(ts[5:])), min(ts), float(np.median(ks[5:])), stats.launches, stats.flags), flush=True)
ds.close()
