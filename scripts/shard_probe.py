#!/usr/bin/env python
"""shard_probe.py [workload] — on ONE GPU, time the share of the frame that rank r of n would render (local canvas, no
peer stores, no process group): separates the kernel's own strong-scaling loss (waves, tail, fixed costs) from what the
multi-process run adds (barrier skew, NVLink stores). Prints ms per (n, rank) and the ideal frame/n."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
PKG = "raytracer-challenge-rs_b200"
pkg = importlib.import_module(PKG); scenes = importlib.import_module(PKG + ".scenes")
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "reflect_refract_1080p"
fn, kw, desc = bench.WORKLOADS[wl]
world, cam = getattr(scenes, fn)(pkg.scene_api(), **kw)
flat = pkg.FlatScene(world); ds = pkg.DeviceScene(flat, device=0)
n = cam.hsize * cam.vsize * 3
canvas = torch.zeros(n, dtype=torch.float64, device="cuda")
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
s = torch.cuda.current_stream(); sp = s.cuda_stream
def t(nr, r, tile_rows=8, reps=7):
    ms = []
    for i in range(reps):
        flush.fill_(float(i))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s)
        ds.render_device(cam, canvas.data_ptr(), stream=sp, tile_rows=tile_rows, rank=r, nranks=nr)
        b.record(s); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    ms.sort(); return ms[len(ms) // 2]
full = t(1, 0)
print("%s full frame %.3f ms" % (wl, full))
for nr in (2, 4, 8):
    for tr in (8, 16):
        per = [t(nr, r, tr) for r in range(nr)]
        print("n=%d tile_rows=%d ideal %.3f  ranks: %s  max %.3f  eff %.2f" % (nr, tr, full / nr, " ".join("%.3f" % x for x in per), max(per), full / nr / max(per)))
