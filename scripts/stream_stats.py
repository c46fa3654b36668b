#!/usr/bin/env python
"""stream_stats.py [workload] [nranks] — one streaming-wavefront frame share through a -DRTK_STREAM_STATS build
(RTC_CUDA_LIB=build/var/librtc_cuda_stats.so): prints the kernel's own counters (stderr of the library) and the time."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
PKG = "raytracer-challenge-rs_b200"
pkg = importlib.import_module(PKG); scenes = importlib.import_module(PKG + ".scenes")
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "reflect_refract_1080p"
nr = int(sys.argv[2]) if len(sys.argv) > 2 else 1
fn, kw, desc = bench.WORKLOADS[wl]
world, cam = getattr(scenes, fn)(pkg.scene_api(), **kw)
ds = pkg.DeviceScene(pkg.FlatScene(world), device=0)
canvas = torch.zeros(cam.hsize * cam.vsize * 3, dtype=torch.float64, device="cuda")
s = torch.cuda.current_stream(); sp = s.cuda_stream
for mode in (dict(wavefront=True, wave_stream=True), dict(wavefront=True)):
    for rep in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s)
        ds.render_device(cam, canvas.data_ptr(), stream=sp, tile_rows=8, rank=0, nranks=nr, **mode)
        b.record(s); torch.cuda.synchronize()
        st = ds.finish(sp)
        print("%s nranks=%d rep %d: %.3f ms, launches %d, flags %d" % (mode, nr, rep, a.elapsed_time(b), st.launches, st.flags), flush=True)
