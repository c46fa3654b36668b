#!/usr/bin/env python
"""make_golden.py — regenerate the golden frames in this directory with the ORACLE (oracle/liboracle.so,
the C++ restatement of the reference pinned by the reference's own known-answer tests; the Rust
reference itself cannot be built in this environment). Each fixture is the linear f64 canvas of one
small scene plus its per-kind ray counts:  <name>.npz = {canvas[h,w,3] float64, rays[4] int64}.
Run from the repo root:  python tests/golden/make_golden.py"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import conftest  # noqa: E402

CASES = [
    ("default_world", dict(width=11, height=11), 1),
    ("chapter7", dict(width=80, height=40), 1),
    ("reflect_refract", dict(width=96, height=54), 1),
    ("cover", dict(width=64, height=36), 1),
    ("primitives_csg", dict(width=96, height=54), 1),
    ("hexagon", dict(width=64, height=36), 1),
    ("glass_sphere2", dict(width=32, height=32, aa=4), 4),
    ("teapot_standin", dict(width=64, height=36, segments=12, rings=8, nu=12, nv=10), 1),
    ("mesh_and_spheres", dict(width=64, height=36, nx=24, nz=12, n_spheres=20), 1),
]


def fixture_name(name, kw, aa):
    return "%s_%dx%d_aa%d.npz" % (name, kw["width"], kw["height"], aa)


if __name__ == "__main__":
    pkg = importlib.import_module("raytracer-challenge-rs_b200")
    scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
    orc = conftest.Oracle(pkg)
    here = os.path.dirname(os.path.abspath(__file__))
    for name, kw, aa in CASES:
        w, c = scenes.CONFIGS[name](orc.api, **kw)
        canvas, info = orc.render(c, w, aa=aa, threads=0 if aa == 1 else 4)
        rays = np.array([info["rays_primary"], info["rays_shadow"], info["rays_reflect"], info["rays_refract"]], dtype=np.int64)
        np.savez_compressed(os.path.join(here, fixture_name(name, kw, aa)), canvas=canvas, rays=rays)
        print(fixture_name(name, kw, aa), canvas.shape, rays.tolist())
