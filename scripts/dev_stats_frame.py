#!/usr/bin/env python
"""dev_stats_frame.py <workload> — one frame through a -DRTK_DEVICE_STATS build (RTC_CUDA_LIB=build/var/librtc_cuda_dstats.so):
the library prints searches / BVH nodes visited / prim tests of the frame (stderr); this prints the frame's rays and ms."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "raytracer-challenge-rs_b200"
pkg = importlib.import_module(PKG); scenes = importlib.import_module(PKG + ".scenes")
import bench
wl = sys.argv[1]
fn, kw, desc = bench.WORKLOADS[wl]
world, cam = getattr(scenes, fn)(pkg.scene_api(), **kw)
ds = pkg.DeviceScene(pkg.FlatScene(world), device=0)
for rep in range(2):
    out, st = ds.render(cam, wavefront=False)
    print("%s rep %d: %.3f ms, rays %d" % (wl, rep, st.kernel_ms, st.rays), flush=True)
