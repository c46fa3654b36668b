"""Committed golden frames (tests/golden/*.npz, made by tests/golden/make_golden.py with the oracle).
CPU tier: the oracle still reproduces them bit for bit (guards the oracle against drift) and the
kernel source compiled for the host matches them. GPU tier: the CUDA path through the C-ABI matches
them (linear colour within the north-star tolerance; bit-exact except the libm-vs-CUDA pow ulps)."""
import importlib
import os

import numpy as np
import pytest

from conftest import compare_images

scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
HERE = os.path.dirname(os.path.abspath(__file__))
import sys
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden import CASES, fixture_name  # noqa: E402


def load(name, kw, aa):
    z = np.load(os.path.join(HERE, "golden", fixture_name(name, kw, aa)))
    return z["canvas"], [int(v) for v in z["rays"]]


@pytest.mark.parametrize("case", CASES, ids=lambda c: fixture_name(*c)[:-4])
def test_oracle_and_kernel_source_reproduce_golden(case, pkg, prod, orc, emul):
    name, kw, aa = case
    want, rays = load(name, kw, aa)
    wo, co = scenes.CONFIGS[name](orc.api, **kw)
    got_o, io = orc.render(co, wo, aa=aa, threads=0 if aa == 1 else 4)
    assert np.array_equal(got_o, want)
    assert [io["rays_primary"], io["rays_shadow"], io["rays_reflect"], io["rays_refract"]] == rays
    wp, cp = scenes.CONFIGS[name](prod, **kw)
    got_e, ie = emul.render(cp, wp, aa=aa)
    assert np.array_equal(got_e, want)
    assert [ie["rays_primary"], ie["rays_shadow"], ie["rays_reflect"], ie["rays_refract"]] == rays


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=lambda c: fixture_name(*c)[:-4])
def test_cuda_path_reproduces_golden(case, pkg, prod, orc):
    name, kw, aa = case
    want, rays = load(name, kw, aa)
    wp, cp = scenes.CONFIGS[name](prod, **kw)
    got, st, _ = pkg.render(cp, wp, device=0, aa=aa)
    assert [st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract] == rays
    frac, max8, lin = compare_images(got, want, orc.quantize)
    assert frac >= 0.999 and max8 <= 1 and lin <= 1e-6, (frac, max8, lin)
    assert lin <= 1e-12, lin
