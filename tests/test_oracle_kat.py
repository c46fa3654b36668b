"""The reference's own known-answer tests (hot path), restated against the oracle: oracle/kat_*.cpp.
This is what pins the oracle; every KAT cites the reference test it ports."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def kat_output(pkg):
    p = subprocess.run([os.path.join(ROOT, "oracle", "kat")], capture_output=True, text=True, timeout=600)
    return p.returncode, p.stdout


def test_all_kats_pass(kat_output):
    rc, out = kat_output
    failed = [l for l in out.splitlines() if l.startswith("FAIL")]
    assert rc == 0 and not failed, "\n".join(failed)


def test_kat_coverage(kat_output):
    _, out = kat_output
    names = [l.split()[1] for l in out.splitlines() if l.startswith("ok ")]
    assert len(names) >= 57
    for prefix in ["world_", "camera_", "sphere_", "plane_", "cube_", "cylinder_", "cone_", "triangle_", "smooth_triangle_",
                   "intersection_", "group_", "csg_", "bbox_"]:
        assert any(n.startswith(prefix) for n in names), prefix
