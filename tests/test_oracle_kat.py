"""The reference's own known-answer tests (hot path), restated against the oracle: oracle/kat_*.cpp.
This is what pins the oracle; every KAT cites the reference test it ports."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def kat_output(pkg):
    env = dict(os.environ, KAT_OBJ_DIR=os.path.join(ROOT, "tests", "golden", "obj"))  # the reference's obj_parser/test_data
    p = subprocess.run([os.path.join(ROOT, "oracle", "kat")], capture_output=True, text=True, timeout=600, env=env)
    return p.returncode, p.stdout


def test_all_kats_pass(kat_output):
    rc, out = kat_output
    failed = [l for l in out.splitlines() if l.startswith("FAIL")]
    assert rc == 0 and not failed, "\n".join(failed)


def test_kat_coverage(kat_output):
    _, out = kat_output
    names = [l.split()[1] for l in out.splitlines() if l.startswith("ok ")]
    assert len(names) >= 115
    for prefix in ["world_", "camera_", "sphere_", "plane_", "cube_", "cylinder_", "cone_", "triangle_", "smooth_triangle_",
                   "intersection_", "group_", "csg_", "bbox_",
                   # round 2: material.rs, pattern/*.rs, image/ppm.rs, matrix.rs, transform.rs, vector.rs, point.rs, color.rs,
                   # ray.rs, light.rs, canvas.rs, obj_parser/mod.rs
                   "material_", "pattern_stripe", "pattern_gradient", "pattern_ring", "pattern_checkers", "pattern_test_pattern",
                   "ppm_", "matrix_", "transform_", "vector_", "point_", "color_", "ray_", "light_", "canvas_", "obj_"]:
        assert any(n.startswith(prefix) for n in names), prefix
