"""CPU tier: the kernel SOURCE compiled for the host (tests/cpu_emul) + the product flattener vs the
oracle, bit for bit, on small frames. Proves traversal/CSG/containers/iterative-recursion logic and
the host mirror before GPU time is spent; the GPU tier repeats the same cases through the C-ABI."""
import importlib
import os

import numpy as np
import pytest

from conftest import ROOT, compare_images
from scene_cases import MESH_CASES, SMALL_CASES, case_id

scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")


@pytest.mark.parametrize("case", SMALL_CASES + MESH_CASES, ids=case_id)
def test_kernel_source_matches_oracle(case, pkg, prod, orc, emul):
    name, kw, aa = case
    wo, co = scenes.CONFIGS[name](orc.api, **kw)
    wp, cp = scenes.CONFIGS[name](prod, **kw)
    want, iw = orc.render(co, wo, aa=aa, threads=0 if aa == 1 else 4)
    got, ig = emul.render(cp, wp, aa=aa, faithful=True)
    assert ig["overflow"] == 0 and ig["nan"] == 0
    for k in ("rays_primary", "rays_shadow", "rays_reflect", "rays_refract"):
        assert ig[k] == iw[k], (k, ig[k], iw[k])
    frac, max8, lin = compare_images(got, want, orc.quantize)
    assert frac == 1.0 and max8 == 0, (frac, max8)
    assert lin <= 1e-12, lin      # bit-exact except libm pow vs itself: same libm here → expect 0
    # the conservative t-range culling must not change a single bit
    got2, ig2 = emul.render(cp, wp, aa=aa, faithful=False)
    assert np.array_equal(got, got2)
    assert ig2["rays"] == ig["rays"]


@pytest.mark.parametrize("case", MESH_CASES, ids=case_id)
def test_threaded_bvh_build_matches_oracle(case, pkg, prod, orc, emul, monkeypatch):
    """The accelerator builder bins large ranges on worker threads and builds independent subtrees
    in parallel (rt_accel_build.hpp); RTC_BVH_PAR_MIN pushes these small meshes through that path.
    The frame must still be the oracle's, bit for bit — the search result never depends on the tree."""
    name, kw, aa = case
    wo, co = scenes.CONFIGS[name](orc.api, **kw)
    wp, cp = scenes.CONFIGS[name](prod, **kw)
    want, iw = orc.render(co, wo, aa=aa, threads=0)
    monkeypatch.setenv("RTC_BVH_PAR_MIN", "16")
    got, ig = emul.render(cp, wp, aa=aa, faithful=False)
    assert ig["rays"] == iw["rays"]
    frac, max8, lin = compare_images(got, want, orc.quantize)
    assert frac == 1.0 and max8 == 0 and lin <= 1e-12


def _flat_tables(pkg, flat):
    """every table of a FlatScene as bytes"""
    import ctypes as C
    d = flat.desc
    sizes = [("nodes", d.n_nodes * 32), ("xforms", d.n_xforms * 96), ("boxes", d.n_boxes * 48), ("limits", d.n_limits * 16),
             ("tri_p1e1e2", d.n_triangles * 72), ("tri_normals", d.n_triangles * 72), ("materials", d.n_materials * 88),
             ("patterns", d.n_patterns * 152), ("lights", d.n_lights * 48)]
    out = {}
    for name, nbytes in sizes:
        ptr = getattr(d, name)
        out[name] = C.string_at(ptr, nbytes) if nbytes else b""
    return out


@pytest.mark.parametrize("name,kw", [
    ("teapot_standin", dict(width=8, height=8, segments=16, rings=12, nu=20, nv=16)),
    ("mesh_and_spheres", dict(width=8, height=8, nx=40, nz=20, n_spheres=30)),
    ("primitives_csg", dict(width=8, height=8)),
    ("solids_field", dict(width=8, height=8)),
    ("hexagon", dict(width=8, height=8)),
    ("pattern_gallery", dict(width=8, height=8)),
    ("cover", dict(width=8, height=8)),
])
@pytest.mark.parametrize("grain", [1, 7, 150])
def test_parallel_flatten_equals_the_serial_walk(name, kw, grain, pkg, prod, monkeypatch):
    """host/rth_flatten.hpp: worlds of 65 536 nodes and more are flattened by worker threads, job by job, and merged in
    pre-order; every table (node records with skip / right / aux / ids / equality classes, transforms, boxes, limits,
    triangles, materials, patterns, lights) must be byte-identical to the serial walk's. Small grains force the spine
    pieces (groups and CSGs too big for one job) that a million-triangle world has at its top."""
    world, cam = scenes.CONFIGS[name](prod, **kw)
    monkeypatch.setenv("RTH_FLATTEN_THREADS", "1")
    serial = _flat_tables(pkg, pkg.FlatScene(world))
    monkeypatch.setenv("RTH_FLATTEN_THREADS", "4")
    monkeypatch.setenv("RTH_FLATTEN_MIN_NODES", "0")
    monkeypatch.setenv("RTH_FLATTEN_GRAIN", str(grain))
    par = pkg.FlatScene(world)
    parallel = _flat_tables(pkg, par)
    for table in serial:
        assert serial[table] == parallel[table], table


@pytest.mark.parametrize("flags", [["-DRTK_BVH_WIDE=1"], ["-DRTK_BVH_F32=0"]], ids=["four_children_per_node", "double_precision_nodes"])
def test_flagged_bvh_variants_still_match_the_oracle(flags, pkg, prod, orc, tmp_path):
    """The measured-and-shelved BVH layouts stay compilable and correct: the kernel source is compiled for the host with
    the variant's flag (four-child collapsed nodes; the 128-byte double-precision nodes of rounds 1-2) and must render
    the mesh cases like the oracle, bit for bit."""
    import subprocess
    from conftest import Emul
    so = str(tmp_path / "libemul_variant.so")
    subprocess.run(["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-pthread", "-shared"] + flags +
                   ["-o", so, os.path.join(ROOT, "tests", "cpu_emul", "emul.cpp")], check=True)
    variant = Emul(pkg, so)
    for name, kw, aa in MESH_CASES[:2] + [("cover", dict(width=64, height=36), 1)]:
        wo, co = scenes.CONFIGS[name](orc.api, **kw)
        wp, cp = scenes.CONFIGS[name](prod, **kw)
        want, iw = orc.render(co, wo, aa=aa, threads=0)
        got, ig = variant.render(cp, wp, aa=aa, faithful=False)
        assert ig["rays"] == iw["rays"], name
        frac, max8, lin = compare_images(got, want, orc.quantize)
        assert frac == 1.0 and max8 == 0 and lin <= 1e-12, name
