#!/usr/bin/env python
"""bvh_scene_times.py [tag] — the BVH scenes at their BASELINE sizes through the recursive kernel and the default
(wavefront where the library picks it): device ms per frame; the frames are saved as /tmp/il_<scene>_<tag>.npy so that
two library variants (RTC_CUDA_LIB) can be compared bit for bit."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("raytracer-challenge-rs_b200")
scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
api = pkg.scene_api()
SIZES = {"teapot_standin": (1920, 1080), "cover": (1920, 1080), "primitives_csg": (1920, 1080), "mesh_and_spheres": (3840, 2160), "hexagon": (1920, 1080)}
TAG = sys.argv[1] if len(sys.argv) > 1 else "0"
for name in sys.argv[2:] or ["teapot_standin", "cover", "primitives_csg", "mesh_and_spheres"]:
    w, h = SIZES[name]
    world, cam = scenes.CONFIGS[name](api, width=w, height=h)
    ds = pkg.DeviceScene(pkg.FlatScene(world), device=0)
    img, st = ds.render(cam, wavefront=False)
    ms = []
    for _ in range(12):
        _, st = ds.render(cam, wavefront=False)
        ms.append(st.kernel_ms)
    ms2 = []
    for _ in range(8):
        _, st2 = ds.render(cam)
        ms2.append(st2.kernel_ms)
    ms3 = []
    for _ in range(8):
        img3, st3 = ds.render(cam, wavefront=True)
        ms3.append(st3.kernel_ms)
    print("%s [%s] recursive %.3f ms (min %.3f)  default %.3f ms (launches %d)  levels %.3f ms (launches %d, flags %d, same %s)  sum %.9f" % (
        name, TAG, float(np.median(ms)), min(ms), float(np.median(ms2)), st2.launches, float(np.median(ms3)), st3.launches, st3.flags,
        np.array_equal(img3, img), float(img.sum())), flush=True)
    np.save("/tmp/il_%s_%s.npy" % (name, TAG), img)
    ds.close()
