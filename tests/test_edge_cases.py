"""Code paths no scene file reaches, CPU tier (kernel source on the host) and GPU tier (C-ABI) against the oracle:
the cone's |a| < EPSILON branch (cone.rs:102-110 — the reference has no test for it, so it is oracle-defined), the
32-warp `big_accel` kernel instantiations (RTC_BIG_ACCEL=1), and the two device error reports actually firing:
RTC_ERR_NAN (a NaN t reaches the sort: intersection.rs:113 panics) and RTC_ERR_OVERFLOW (bounded lists, cap 64)."""
import importlib

import numpy as np
import pytest

from conftest import compare_images
from scene_cases import MESH_CASES, SMALL_CASES, case_id

scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")

R2 = 2.0 ** 0.5 / 2.0


def cone_rays():
    """Directions with dx^2 - dy^2 + dz^2 == 0 exactly ((0,±1,±1)/sqrt2 and (±1,±1,0)/sqrt2 have equal squares), over
    origins that give b != 0 (one root: t = -c/2*b) and b == 0 (caps only); plus near-degenerate directions."""
    rng = np.random.default_rng(20261105)
    dirs = [(0, R2, R2), (0, -R2, R2), (0, R2, -R2), (0, -R2, -R2), (R2, R2, 0), (-R2, R2, 0), (R2, -R2, 0), (-R2, -R2, 0),
            (0, 1, 1), (0, -1, 1), (1, -1, 0), (0.5, -R2, 0.5)]
    o, d = [], []
    for dv in dirs:
        for _ in range(40):
            o.append(rng.uniform(-2.5, 2.5, 3))
            d.append(dv)
        for c in ((0, 0, 0), (4, 0, 0), (-4, 1, 1)):       # on an axis apex: b == 0 and c == 0
            o.append(c)
            d.append(dv)
        for c in ((0, 0.5, 0.5), (4, -0.25, 0.25)):         # oy*dy == oz*dz → b == 0
            o.append(c)
            d.append(dv)
    o, d = np.array(o, dtype=np.float64), np.array(d, dtype=np.float64)
    # near-degenerate: a = O(1e-6) straddling EPSILON = 1e-5
    eps = rng.uniform(-3e-5, 3e-5, size=(200, 1))
    o2 = rng.uniform(-2.5, 2.5, size=(200, 3))
    d2 = np.tile(np.array([[0.0, R2, R2]]), (200, 1)) + np.concatenate([np.zeros((200, 1)), eps, np.zeros((200, 1))], axis=1)
    return np.concatenate([o, o2]), np.concatenate([d, d2])


def test_cone_parallel_branch_kernel_source(pkg, prod, orc, emul):
    o, d = cone_rays()
    rgb0, t0 = orc.trace_rays(scenes.cone_field(orc.api), o, d)
    rgb1, t1 = emul.trace_rays(pkg.FlatScene(scenes.cone_field(prod)), o, d)
    assert np.sum(t0 >= 0) > 100                      # the branch really produces hits
    assert np.array_equal(t1, t0) and np.array_equal(rgb1, rgb0)


@pytest.mark.gpu
def test_cone_parallel_branch_gpu(pkg, prod, orc):
    o, d = cone_rays()
    rgb0, t0 = orc.trace_rays(scenes.cone_field(orc.api), o, d)
    ds = pkg.DeviceScene(pkg.FlatScene(scenes.cone_field(prod)))
    rgb1, t1 = ds.trace_rays(o, d)
    assert np.array_equal(t1, t0)
    assert np.max(np.abs(rgb1 - rgb0)) <= 1e-12
    ds.close()
    # and through the accelerated search (render path): a camera looking along (0,-1,1)/sqrt2 at the cones
    for api_, kind in ((orc.api, "o"), (prod, "p")):
        w = scenes.cone_field(api_)
        c = api_.camera(64, 48, 0.9)
        c.set_transform(pkg.transform.view_transform((0, 4, -4), (0, 0, 0), (0, 1, 0)))
        if kind == "o":
            want, iw = orc.render(c, w, threads=0)
        else:
            got, st, _ = pkg.render(c, w, device=0)
    assert st.rays == iw["rays"]
    frac, max8, lin = compare_images(got, want, orc.quantize)
    assert frac == 1.0 and lin <= 1e-12, (frac, max8, lin)


BIG_CASES = [MESH_CASES[0], MESH_CASES[1], ("primitives_csg", dict(width=96, height=54), 1), ("solids_field", dict(width=96, height=54), 1)]


@pytest.mark.gpu
@pytest.mark.parametrize("case", BIG_CASES, ids=case_id)
def test_big_accel_instantiations_match_oracle(case, pkg, prod, orc, monkeypatch):
    """render_kernel<..,8> / wave_forward_kernel<..,8,..> (64 registers, 32 warps per SM) are picked only when the BVH
    exceeds 64 MB — the 1 M-triangle scene; RTC_BIG_ACCEL=1 pushes the small cases through them, recursive and wavefront."""
    name, kw, aa = case
    wo, co = scenes.CONFIGS[name](orc.api, **kw)
    want, iw = orc.render(co, wo, aa=aa, threads=0)
    monkeypatch.setenv("RTC_BIG_ACCEL", "1")
    wp, cp = scenes.CONFIGS[name](prod, **kw)
    ds = pkg.DeviceScene(pkg.FlatScene(wp))
    for kw, launches in ((dict(wavefront=False), 1), (dict(wavefront=True), 11), (dict(wavefront=True, wave_stream=True), 1)):
        got, st = ds.render(cp, aa=aa, **kw)
        streaming = bool(kw.get("wave_stream"))
        assert st.launches == launches + (1 if st.wave_fallback else 0) and (streaming or not st.wave_fallback), kw
        assert (st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract) == \
            (iw["rays_primary"], iw["rays_shadow"], iw["rays_reflect"], iw["rays_refract"]), kw
        frac, max8, lin = compare_images(got, want, orc.quantize)
        assert frac >= 0.999 and max8 <= 1 and lin <= 1e-12, (kw, frac, max8, lin)
    ds.close()


def nested_glass(api, pkg, n):
    """n concentric glass spheres seen from inside the innermost: the containers walk of prepare_computations
    (intersection.rs:63-90) meets n negative-t entries before the hit."""
    w = api.world()
    w.add_light((0, 0, 0), (1, 1, 1))
    for k in range(n):
        r = 1.0 + 0.05 * k
        w.add_object(api.sphere().set_transform(pkg.transform.scaling(r, r, r))
                     .set_material(pkg.Material(color=(0.2, 0.2, 0.2), transparency=0.9, reflective=0.5, refractive_index=1.0 + 0.01 * (k + 1))))
    return w


@pytest.mark.gpu
def test_overflow_and_nan_are_reported(pkg, prod, orc):
    T = pkg.transform
    # 60 nested shells fit the bounded negative-t list (cap 64) and match the oracle ...
    o = np.array([[0.0, 0.0, 0.0]])
    d = np.array([[0.0, 0.0, 1.0]])
    ds = pkg.DeviceScene(pkg.FlatScene(nested_glass(prod, pkg, 60)))
    rgb1, t1 = ds.trace_rays(o, d, remaining=1)
    rgb0, t0 = orc.trace_rays(nested_glass(orc.api, pkg, 60), o, d, remaining=1)
    assert np.array_equal(t1, t0) and np.max(np.abs(rgb1 - rgb0)) <= 1e-12
    ds.close()
    # ... 70 do not: reported, never silently truncated — by the ray-batch entry and by a whole frame
    w = nested_glass(prod, pkg, 70)
    ds = pkg.DeviceScene(pkg.FlatScene(w))
    with pytest.raises(pkg.RtcError) as e:
        ds.trace_rays(o, d, remaining=1)
    assert e.value.code == -3                              # RTC_ERR_OVERFLOW
    c = prod.camera(16, 16, 0.5)
    c.set_transform(T.view_transform((0, 0, 0), (0, 0, 1), (0, 1, 0)))
    for kw in (dict(wavefront=False), dict(wavefront=True), dict(wavefront=True, wave_stream=True)):
        with pytest.raises(pkg.RtcError) as e:
            ds.render(c, **kw)
        assert e.value.code == -3, kw
    ds.close()
    # a zero direction makes the sphere's t = (-0 ± 0) / 0 = NaN: the reference panics in the sort
    # (intersection.rs:113), the oracle throws, the device reports RTC_ERR_NAN
    wd, _ = scenes.default_world(prod)
    ds = pkg.DeviceScene(pkg.FlatScene(wd))
    z = np.array([[0.0, 0.0, 0.0]])
    with pytest.raises(pkg.RtcError) as e:
        ds.trace_rays(np.array([[0.0, 0.0, -5.0]]), z)
    assert e.value.code == -4                              # RTC_ERR_NAN
    with pytest.raises(RuntimeError):
        orc.trace_rays(scenes.default_world(orc.api)[0], np.array([[0.0, 0.0, -5.0]]), z)
    # a degenerate camera (pixel == origin → normalize(0) = NaN direction) does the same to a whole frame
    c = prod.camera(8, 8, 0.5)
    cam = pkg.camera_to_rtc(c)
    for i in range(16):
        cam.inverse[i] = 0.0
    import ctypes as C
    out = np.zeros((8, 8, 3))
    st = pkg.RtcStats()
    opts = pkg.make_opts()
    rc = pkg.cuda_lib().rtc_render(ds.h, C.byref(cam), C.byref(opts), out.ctypes.data, C.byref(st))
    assert rc == -4 and st.nan_count > 0
    ds.close()


def test_overflow_is_reported_by_kernel_source(pkg, prod, emul):
    import ctypes as C
    flat = pkg.FlatScene(nested_glass(prod, pkg, 70))
    o = np.array([[0.0, 0.0, 0.0]])
    d = np.array([[0.0, 0.0, 1.0]])
    rgb, t = np.zeros_like(o), np.zeros(1)
    err = emul.lib.emul_trace_rays(C.byref(flat.desc), 1, o.ctypes.data, d.ctypes.data, 1, rgb.ctypes.data, t.ctypes.data)
    assert err & 1                                         # ERR_OVERFLOW


def cube_grazing_world(api, pkg, rotated=False):
    """A unit cube plus nine far cubes (enough boundable prims for a BVH). Cube::check_axis (cube.rs:30-48) calls an
    object-space direction component below EPSILON parallel and then decides by the ORIGIN alone, so a ray starting
    inside the x-slab with 0 < dx < 1e-5 'hits' the cube although its exact path has left the slab: a hit outside any
    box around the cube."""
    T = pkg.transform
    w = api.world()
    w.add_light((-10, 10, -10), (1, 1, 1))
    rot = T.rotation_y(0.3) if rotated else T.translation(0, 0, 0)
    w.add_object(api.cube().set_transform(rot).set_material(pkg.Material(color=(0.9, 0.3, 0.2), reflective=0.2)))
    for k in range(9):
        w.add_object(api.cube().set_transform(T.chain(T.scaling(0.5, 0.5, 0.5), rot, T.translation(40.0 + 3.0 * k, 5.0, 30.0)))
                     .set_material(pkg.Material(color=(0.2, 0.3 + 0.05 * k, 0.9))))
    return w


def grazing_camera(api, pkg, n, rotated=False):
    import math
    T = pkg.transform
    c = api.camera(n, n, 6e-5)   # directions span dx in about +-3e-5: both sides of EPSILON
    if rotated:
        ca, sa = math.cos(0.3), math.sin(0.3)
        # the unrotated set-up turned with the cube: origin just inside the x-slab in object space
        o = (0.99999 * ca + -5.0 * sa, 0.3, -0.99999 * sa + -5.0 * ca)
        t = (0.99999 * ca + 0.0 * sa, 0.3, -0.99999 * sa + 0.0 * ca)
    else:
        o, t = (0.99999, 0.3, -5.0), (0.99999, 0.3, 0.0)
    c.set_transform(T.view_transform(o, t, (0, 1, 0)))
    return c


@pytest.mark.parametrize("rotated", [False, True])
def test_cube_epsilon_parallel_hits_survive_the_bvh_kernel_source(pkg, prod, orc, emul, rotated):
    want, iw = orc.render(grazing_camera(orc.api, pkg, 24, rotated), cube_grazing_world(orc.api, pkg, rotated), threads=0)
    wp, cp = cube_grazing_world(prod, pkg, rotated), grazing_camera(prod, pkg, 24, rotated)
    flat = pkg.FlatScene(wp)
    a, ia = emul.render(cp, wp, flat=flat, faithful=True)
    b, ib = emul.render(cp, wp, flat=flat, faithful=False)
    assert np.array_equal(a, want) and ia["rays"] == iw["rays"]
    assert np.array_equal(b, want) and ib["rays"] == iw["rays"]
    lit = np.sum(np.any(want > 0, axis=-1))
    assert 0 < lit                                    # the quirk's hits are in the frame
    # the advisor's single ray: 1x1 camera at (0.999999, 0, -5) looking along (5e-6, 0, 1)
    if not rotated:
        for api_, rend in ((orc.api, None), (prod, emul)):
            c1 = api_.camera(1, 1, 0.1)
            c1.set_transform(pkg.transform.view_transform((0.999999, 0, -5), (0.999999 + 5e-6, 0, -4), (0, 1, 0)))
            w1 = cube_grazing_world(api_, pkg)
            if rend is None:
                w0, i0 = orc.render(c1, w1, threads=0)
            else:
                g0, j0 = emul.render(c1, w1, faithful=False)
        assert np.array_equal(g0, w0) and j0["rays"] == i0["rays"] and i0["rays"] >= 2


@pytest.mark.gpu
@pytest.mark.parametrize("rotated", [False, True])
def test_cube_epsilon_parallel_hits_survive_the_bvh_gpu(pkg, prod, orc, rotated):
    want, iw = orc.render(grazing_camera(orc.api, pkg, 48, rotated), cube_grazing_world(orc.api, pkg, rotated), threads=0)
    ds = pkg.DeviceScene(pkg.FlatScene(cube_grazing_world(prod, pkg, rotated)))
    cp = grazing_camera(prod, pkg, 48, rotated)
    for kw in (dict(faithful=True), dict(faithful=False), dict(wavefront=True), dict(wavefront=True, wave_stream=True)):
        got, st = ds.render(cp, **kw)
        assert st.rays == iw["rays"], kw
        assert np.max(np.abs(got - want)) <= 1e-12, kw
    ds.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", [MESH_CASES[0], MESH_CASES[1], ("solids_field", dict(width=96, height=54), 1)], ids=case_id)
def test_device_built_bvh_matches_oracle(case, pkg, prod, orc, monkeypatch):
    """RTC_DEVICE_BVH_MIN: the acceleration structure built on the GPU (rt_accel_device.cuh: Morton sort, SAH cuts over the
    Morton order, one warp per node) instead of the host's SAH builder. The search result never depends on the tree, so
    the frame must be the oracle's bit for bit — recursive kernel, both wavefronts, and the faithful traversal beside it.
    (solids_field has cubes in the tree: the device builder must step aside for the host's.)"""
    name, kw, aa = case
    wo, co = scenes.CONFIGS[name](orc.api, **kw)
    want, iw = orc.render(co, wo, aa=aa, threads=0)
    monkeypatch.setenv("RTC_DEVICE_BVH_MIN", "16")
    wp, cp = scenes.CONFIGS[name](prod, **kw)
    ds = pkg.DeviceScene(pkg.FlatScene(wp))
    for kwr in (dict(wavefront=False), dict(faithful=True), dict(wavefront=True), dict(wavefront=True, wave_stream=True)):
        got, st = ds.render(cp, aa=aa, **kwr)
        assert st.rays == iw["rays"], kwr
        frac, max8, lin = compare_images(got, want, orc.quantize)
        assert frac >= 0.999 and max8 <= 1 and lin <= 1e-12, (kwr, frac, max8, lin)
    ds.close()


def _far_cluster(api, pkg, centre, extent, n_tri=300, n_sph=40, n_cube=12, seed=7):
    """small triangles, spheres and cubes in a cluster of size `extent` around `centre` (far from the world origin when
    |centre| >> extent: the regime where a single-precision box test is coarsest), seen from just outside it"""
    T = pkg.transform
    rng = np.random.default_rng(seed)
    w = api.world()
    c = np.asarray(centre, dtype=np.float64)
    w.add_light(tuple(c + extent * np.array([-2.0, 3.0, -4.0])), (1.0, 1.0, 1.0))
    p0 = c + extent * rng.uniform(-1.0, 1.0, size=(n_tri, 1, 3))
    pts = p0 + 0.15 * extent * rng.uniform(-1.0, 1.0, size=(n_tri, 3, 3))
    nrm = np.tile(np.array([[[0.0, 1.0, 0.0]]]), (n_tri, 3, 1))
    mesh = api.group()
    mesh.add_triangles(np.ascontiguousarray(pts), np.ascontiguousarray(nrm))
    mesh.set_material(pkg.Material(color=(0.8, 0.5, 0.3), reflective=0.3))
    w.add_object(mesh)
    for k in range(n_sph):
        p = c + extent * rng.uniform(-1.0, 1.0, size=3)
        r = extent * rng.uniform(0.03, 0.12)
        mat = pkg.Material(color=(0.3, 0.6, 0.9), reflective=0.5) if k % 2 else pkg.Material(
            color=(0.1, 0.1, 0.1), transparency=0.9, reflective=0.9, refractive_index=1.5)
        w.add_object(api.sphere().set_transform(T.matmul(T.translation(*p), T.scaling(r, r, r))).set_material(mat))
    for k in range(n_cube):
        p = c + extent * rng.uniform(-1.0, 1.0, size=3)
        r = extent * rng.uniform(0.05, 0.1)
        w.add_object(api.cube().set_transform(T.chain(T.scaling(r, r, r), T.rotation_y(0.3 * k), T.translation(*p)))
                     .set_material(pkg.Material(color=(0.9, 0.9, 0.2))))
    cam = api.camera(48, 48, 0.9)
    cam.set_transform(T.view_transform(tuple(c + extent * np.array([0.3, 0.8, -3.5])), tuple(c), (0, 1, 0)))
    return w, cam


FAR_CLUSTERS = [((0.0, 0.0, 0.0), 1.0), ((5000.0, -3000.0, 8000.0), 5.0), ((3.0e6, 1.0e6, -2.0e6), 40.0),
                ((0.5, 0.5, 0.5), 0.8), ((-70000.0, 10.0, 0.5), 0.8)]


@pytest.mark.parametrize("centre,extent", FAR_CLUSTERS, ids=lambda v: str(v))
def test_single_precision_boxes_never_lose_a_hit_kernel_source(centre, extent, pkg, prod, emul):
    """rt_accel.cuh Slab32: the BVH's boxes and slab test are single precision, widened by a bound on their own rounding
    that grows with the tree's distance from the origin. Accelerated search == the reference's own traversal, bit for
    bit, on clusters at the origin and thousands and millions of units away from it (shapes no smaller than the
    reference's matrix inverse accepts: |det| >= 1e-5)."""
    w, cam = _far_cluster(prod, pkg, centre, extent)
    flat = pkg.FlatScene(w)
    want, iw = emul.render(cam, w, flat=flat, faithful=True)
    got, ig = emul.render(cam, w, flat=flat, faithful=False)
    assert ig["rays"] == iw["rays"] and ig["rays"] > 48 * 48
    assert np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("centre,extent", FAR_CLUSTERS, ids=lambda v: str(v))
def test_single_precision_boxes_never_lose_a_hit_gpu(centre, extent, pkg, prod):
    w, cam = _far_cluster(prod, pkg, centre, extent)
    ds = pkg.DeviceScene(pkg.FlatScene(w))
    want, sw = ds.render(cam, faithful=True)
    got, sg = ds.render(cam, wavefront=False)
    got2, sg2 = ds.render(cam, wavefront=True)
    ds.close()
    assert sg.rays == sw.rays and sg2.rays == sw.rays
    assert np.array_equal(got, want) and np.array_equal(got2, want)
