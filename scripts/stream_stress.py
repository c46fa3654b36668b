#!/usr/bin/env python
"""stream_stress.py [n] — the streaming wavefront over and over on the small parity scenes (both register budgets), looking
for a frame that is abandoned (RTC_STAT_WAVE_FALLBACK) or differs from the recursive kernel's."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
PKG = "raytracer-challenge-rs_b200"
pkg = importlib.import_module(PKG); scenes = importlib.import_module(PKG + ".scenes")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
bad = 0
for big in ("0", "1"):
    os.environ["RTC_BIG_ACCEL"] = big
    for name, kw in (("primitives_csg", dict(width=96, height=54)), ("reflect_refract", dict(width=160, height=90)),
                     ("solids_field", dict(width=96, height=54)), ("mesh_and_spheres", dict(width=96, height=54, nx=40, nz=20, n_spheres=30))):
        w, c = scenes.CONFIGS[name](pkg.scene_api(), **kw)
        ds = pkg.DeviceScene(pkg.FlatScene(w))
        ref, sr = ds.render(c, wavefront=False)
        t0 = time.perf_counter()
        for k in range(n):
            if k % 5 == 4:
                ds.render(c, wavefront=True)          # a frame of the passes in between (dirties the child slots)
            got, st = ds.render(c, wavefront=True, wave_stream=True)
            if st.wave_fallback or not np.array_equal(got, ref) or st.rays != sr.rays:
                bad += 1
                print("BAD big=%s %s iter %d fallback=%s identical=%s" % (big, name, k, st.wave_fallback, np.array_equal(got, ref)), flush=True)
        print("big=%s %-18s %d frames %.2f s" % (big, name, n, time.perf_counter() - t0), flush=True)
        ds.close()
print("STRESS", "OK" if bad == 0 else "FAILED %d" % bad)
