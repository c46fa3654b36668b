#!/usr/bin/env python
"""scene_create_timing.py [workload] — RTC_TIMING=1 phases of flatten + rtc_scene_create for one workload (stderr of the library)."""
import importlib, os, sys, time
os.environ["RTC_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("raytracer-challenge-rs_b200"); scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "mesh_and_spheres_4k"
fn, kw, desc = bench.WORKLOADS[wl]
world, cam = getattr(scenes, fn)(pkg.scene_api(), **kw)
for rep in range(3):
    t0 = time.perf_counter(); flat = pkg.FlatScene(world); t1 = time.perf_counter()
    ds = pkg.DeviceScene(flat, device=0); t2 = time.perf_counter()
    print("%s rep %d: flatten %.1f ms, scene create %.1f ms" % (wl, rep, (t1 - t0) * 1e3, (t2 - t1) * 1e3), flush=True)
    ds.close()
