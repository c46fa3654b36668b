#!/usr/bin/env python
"""shard_probe2.py [workload] — on ONE GPU, the share of the frame rank r of n would render (local device canvas, no
process group), for the three renderers: recursive kernel, streaming wavefront (rt_stream.cuh, one launch) and the
level-synchronous wavefront (rt_wave.cuh). Prints the slowest rank's ms per n and the efficiency against frame/n."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
PKG = "raytracer-challenge-rs_b200"
pkg = importlib.import_module(PKG); scenes = importlib.import_module(PKG + ".scenes")
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "reflect_refract_1080p"
ns = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 2, 4, 8]
fn, kw, desc = bench.WORKLOADS[wl]
world, cam = getattr(scenes, fn)(pkg.scene_api(), **kw)
flat = pkg.FlatScene(world); ds = pkg.DeviceScene(flat, device=0)
n = cam.hsize * cam.vsize * 3
canvas = torch.zeros(n, dtype=torch.float64, device="cuda")
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
s = torch.cuda.current_stream(); sp = s.cuda_stream
MODES = {"recursive": dict(wavefront=False), "stream": dict(wavefront=True, wave_stream=True), "levels": dict(wavefront=True)}
def t(nr, r, mode, reps=7):
    ms = []
    for i in range(reps):
        flush.fill_(float(i))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s)
        ds.render_device(cam, canvas.data_ptr(), stream=sp, tile_rows=8, rank=r, nranks=nr, **MODES[mode])
        b.record(s); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    st = ds.finish(sp)
    assert not st.wave_overflowed, (mode, nr, r)
    ms.sort(); return ms[len(ms) // 2]
best_full = None
for mode in MODES:
    row = []
    for nr in ns:
        per = [t(nr, r, mode) for r in range(nr)]
        row.append(max(per))
    if best_full is None or (ns[0] == 1 and row[0] < best_full):
        best_full = row[0]
    print("%-10s %s" % (mode, "  ".join("n=%d %.3f ms" % (nr, x) for nr, x in zip(ns, row))), flush=True)
print("workload %s; efficiency vs the best full frame (%.3f ms) / n is ms_ideal / ms" % (wl, best_full))
