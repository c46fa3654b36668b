"""CPU tier: the kernel SOURCE compiled for the host (tests/cpu_emul) + the product flattener vs the
oracle, bit for bit, on small frames. Proves traversal/CSG/containers/iterative-recursion logic and
the host mirror before GPU time is spent; the GPU tier repeats the same cases through the C-ABI."""
import importlib

import numpy as np
import pytest

from conftest import compare_images
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
