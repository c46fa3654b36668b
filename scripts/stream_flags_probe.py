#!/usr/bin/env python
"""stream_flags_probe.py — the streaming wavefront's policy switches (RTC_STREAM_FLAGS), A/B in one process per value:
prints ms for the whole glass frame and for an eighth of it (slowest of ranks 0 and 3)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = r'''
import importlib, os, sys
sys.path.insert(0, %r)
import torch
PKG = "raytracer-challenge-rs_b200"
pkg = importlib.import_module(PKG); scenes = importlib.import_module(PKG + ".scenes")
import bench
fn, kw, desc = bench.WORKLOADS[sys.argv[1]]
world, cam = getattr(scenes, fn)(pkg.scene_api(), **kw)
ds = pkg.DeviceScene(pkg.FlatScene(world), device=0)
canvas = torch.zeros(cam.hsize * cam.vsize * 3, dtype=torch.float64, device="cuda")
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
s = torch.cuda.current_stream(); sp = s.cuda_stream
def t(nr, r):
    ms = []
    for i in range(9):
        flush.fill_(float(i))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s); ds.render_device(cam, canvas.data_ptr(), stream=sp, tile_rows=8, rank=r, nranks=nr, wavefront=True, wave_stream=True); b.record(s)
        torch.cuda.synchronize(); ms.append(a.elapsed_time(b))
    st = ds.finish(sp); assert not st.wave_overflowed
    ms.sort(); return ms[len(ms) // 2]
print("flags %%s  n=1 %%.3f  n=2 %%.3f  n=4 %%.3f  n=8 %%.3f" %% (os.environ.get("RTC_STREAM_FLAGS"), t(1, 0), max(t(2, 0), t(2, 1)), max(t(4, 0), t(4, 2)), max(t(8, 0), t(8, 3))), flush=True)
''' % ROOT
wl = sys.argv[1] if len(sys.argv) > 1 else "reflect_refract_1080p"
for flags in (sys.argv[2].split(",") if len(sys.argv) > 2 else ["0", "1", "2", "4", "5", "6", "7"]):
    env = dict(os.environ, RTC_STREAM_FLAGS=flags)
    subprocess.run([sys.executable, "-c", code, wl], env=env, check=False)
