#!/usr/bin/env python
"""verify_config5.py — config 5 (1 M triangles + 200 spheres, 3840x2160) at its BASELINE size on one GPU:
  1. accelerated search == faithful traversal (RTC_OPT_FAITHFUL_TRAVERSAL: what Group::intersect visits) on the FULL
     frame, bit for bit, per-kind ray counts equal — recursive kernel, streaming wavefront, level-synchronous wavefront;
  2. a block of rows through the mesh silhouette against the oracle (bounded: the oracle needs seconds per row).
Prints one JSON line; exit code 1 on any mismatch."""
import importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
PKG = "raytracer-challenge-rs_b200"
pkg = importlib.import_module(PKG); scenes = importlib.import_module(PKG + ".scenes")
import conftest
W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (3840, 2160)
orows = int(sys.argv[3]) if len(sys.argv) > 3 else 16
t0 = time.perf_counter()
world, cam = scenes.mesh_and_spheres(pkg.scene_api(), width=W, height=H)
flat = pkg.FlatScene(world)
ds = pkg.DeviceScene(flat, device=0)
out = {"width": W, "height": H, "scene_s": time.perf_counter() - t0}
ok = True
ref, sr = ds.render(cam, faithful=True)
out["faithful_kernel_ms"] = sr.kernel_ms
for name, kw in (("recursive", dict(wavefront=False)), ("stream", dict(wavefront=True, wave_stream=True)), ("levels", dict(wavefront=True))):
    got, st = ds.render(cam, **kw)
    same = bool(np.array_equal(got, ref))
    rays = (st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract) == (sr.rays_primary, sr.rays_shadow, sr.rays_reflect, sr.rays_refract)
    out[name] = {"identical_to_faithful": same, "rays_equal": bool(rays), "kernel_ms": st.kernel_ms, "launches": st.launches,
                 "fallback": st.wave_fallback}
    ok = ok and same and rays and not st.wave_fallback
out["rays"] = int(sr.rays)
# oracle rows through the silhouette (the mesh sits around the image centre)
orc = conftest.Oracle(pkg)
wo, co = scenes.mesh_and_spheres(orc.api, width=W, height=H)
r0 = H // 2 - orows // 2
t0 = time.perf_counter()
want, info = orc.render(co, wo, threads=os.cpu_count() or 1, rows=(r0, r0 + orows))
out["oracle_s"] = time.perf_counter() - t0
g, w = ref[r0:r0 + orows], want[r0:r0 + orows]
q = lambda a: np.clip(np.round(a * 255.0), 0, 255).astype(np.uint8)
out["oracle_rows"] = [r0, r0 + orows]
out["oracle_identical_8bit"] = float(np.mean(np.all(q(g) == q(w), axis=-1)))
out["oracle_max_linear_diff"] = float(np.max(np.abs(g - w)))
out["oracle_flops_per_ray"] = info["flops"] / max(1, info["rays"])
ok = ok and out["oracle_identical_8bit"] >= 0.999 and out["oracle_max_linear_diff"] <= 1e-6
out["ok"] = bool(ok)
print(json.dumps(out), flush=True)
sys.exit(0 if ok else 1)
