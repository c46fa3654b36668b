"""Scene cases shared by the CPU-tier (kernel source on the host) and GPU-tier parity tests."""
SMALL_CASES = [
    # name, kwargs, aa
    ("default_world", dict(width=11, height=11), 1),
    ("chapter7", dict(width=160, height=80), 1),
    ("reflect_refract", dict(width=160, height=90), 1),
    ("cover", dict(width=128, height=72), 1),
    ("primitives_csg", dict(width=160, height=90), 1),
    ("glass_sphere2", dict(width=80, height=80), 1),
    ("hexagon", dict(width=128, height=72), 1),
    ("solids_field", dict(width=160, height=90), 1),
    ("reflect_refract", dict(width=64, height=36), 4),
    ("glass_sphere2", dict(width=32, height=32, aa=16), 16),
    # round 2: the pattern kinds and the aa tables (camera.rs:95 x2, :97-106 x8) that nothing exercised
    ("pattern_gallery", dict(width=160, height=90), 1),
    ("chapter7", dict(width=64, height=32), 2),
    ("primitives_csg", dict(width=64, height=36), 8),
    ("pattern_gallery", dict(width=48, height=27), 2),
    ("cover", dict(width=33, height=19), 1),   # odd width, cubes in the BVH (see MESH_CASES)
]
MESH_CASES = [
    ("teapot_standin", dict(width=96, height=54, segments=16, rings=12, nu=20, nv=16), 1),
    # odd width: the centre column's rays have a zero (or 1e-17) x component — the BVH's double-precision slab path
    # (the single-precision one needs |1/d| <= 1e15 on every axis, rt_accel.cuh)
    ("teapot_standin", dict(width=33, height=19, segments=16, rings=12, nu=20, nv=16), 1),
    ("mesh_and_spheres", dict(width=96, height=54, nx=40, nz=20, n_spheres=30), 1),
]


def case_id(c):
    return "%s-%dx%d-aa%d" % (c[0], c[1].get("width", 0), c[1].get("height", 0), c[2])
