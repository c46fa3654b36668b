"""GPU tier (B200): the CUDA path through the C-ABI (librtc_cuda.so) vs the oracle on the same seeded
scenes. Bar (north_star): >= 99.9 % identical 8-bit pixels, no channel off by more than 1, linear
colour within 1e-6 — and identical per-kind ray counts, which catches control-flow divergence that a
colour tolerance could hide. Everything except libm-vs-CUDA pow() (specular term, <= 2 ulp) is
expected to be bit-exact."""
import ctypes as C
import importlib

import numpy as np
import pytest

from conftest import compare_images
from scene_cases import MESH_CASES, SMALL_CASES, case_id

scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
pytestmark = pytest.mark.gpu

LIN_TOL = 1e-6      # north_star: linear colour within 1e-6
EXPECT_LIN = 1e-12  # what the design actually delivers (pow ulps only)


@pytest.mark.parametrize("case", SMALL_CASES + MESH_CASES, ids=case_id)
def test_render_matches_oracle(case, pkg, prod, orc):
    name, kw, aa = case
    wo, co = scenes.CONFIGS[name](orc.api, **kw)
    wp, cp = scenes.CONFIGS[name](prod, **kw)
    want, iw = orc.render(co, wo, aa=aa, threads=0 if aa == 1 else 8)
    got, st, _ = pkg.render(cp, wp, device=0, aa=aa)
    assert st.overflow_count == 0 and st.nan_count == 0
    assert (st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract) == \
        (iw["rays_primary"], iw["rays_shadow"], iw["rays_reflect"], iw["rays_refract"])
    frac, max8, lin = compare_images(got, want, orc.quantize)
    assert frac >= 0.999 and max8 <= 1 and lin <= LIN_TOL, (frac, max8, lin)
    assert lin <= EXPECT_LIN, lin
    # device quantiser == reference quantiser on this frame
    assert np.array_equal(pkg.quantize(got), orc.quantize(got))
    # PPM text identical to the oracle's canvas_to_ppm of its own frame when the 8-bit images agree
    if frac == 1.0:
        assert pkg.canvas_to_ppm(got) == orc.canvas_to_ppm(want)


@pytest.mark.parametrize("case", [SMALL_CASES[2], SMALL_CASES[4], MESH_CASES[0], MESH_CASES[1]], ids=case_id)
def test_culling_is_bit_neutral(case, pkg, prod):
    """t-range box culling (not in the reference) vs the faithful traversal: same bits, same rays."""
    name, kw, aa = case
    wp, cp = scenes.CONFIGS[name](prod, **kw)
    ds = pkg.DeviceScene(pkg.FlatScene(wp))
    a, sa = ds.render(cp, aa=aa, faithful=True)
    b, sb = ds.render(cp, aa=aa, faithful=False)
    assert np.array_equal(a, b)
    assert sa.rays == sb.rays
    ds.close()


@pytest.mark.parametrize("case", [c for c in SMALL_CASES + MESH_CASES if c[2] == 1], ids=case_id)
def test_wavefront_is_bit_identical(case, pkg, prod):
    """The level-synchronous wavefront renderer (rt_wave.cuh: one work item per color_at call) vs the
    one-thread-per-pixel kernel: same bits, same per-kind ray counts; also sharded (cyclic tiles) and into
    a pinned host canvas, and with max_depth 0 (no children at all) and 2."""
    name, kw, aa = case
    wp, cp = scenes.CONFIGS[name](prod, **kw)
    ds = pkg.DeviceScene(pkg.FlatScene(wp))
    for depth in (5, 0, 2):
        a, sa = ds.render(cp, aa=1, max_depth=depth, wavefront=False)
        assert sa.launches == 1 and not (sa.flags & pkg.RTC_STAT_WAVEFRONT)
        # the level-synchronous passes (rt_wave.cuh) and the barrier-free persistent launch (rt_stream.cuh)
        for stream in (False, True):
            b, sb = ds.render(cp, aa=1, max_depth=depth, wavefront=True, wave_stream=stream)
            want_launches = (1 if stream else 2 * depth + 1) if depth > 0 else 1
            # (the experimental streaming launch may abandon a frame — a colour that did not arrive in time — and hand
            # it to the recursive kernel: one more launch, same frame; the passes never do on these scenes)
            assert sb.launches == want_launches + (1 if sb.wave_fallback else 0), (depth, stream, sb.launches)
            assert stream or not sb.wave_fallback, (depth, stream)
            assert bool(sb.flags & pkg.RTC_STAT_WAVEFRONT) == (depth > 0)
            assert np.array_equal(a, b), (depth, stream)
            assert (sa.rays_primary, sa.rays_shadow, sa.rays_reflect, sa.rays_refract) == \
                (sb.rays_primary, sb.rays_shadow, sb.rays_reflect, sb.rays_refract), (depth, stream)
    full, _ = ds.render(cp, aa=1, wavefront=False)
    parts = np.zeros_like(full)
    for stream in (False, True):
        parts[:] = 0.0
        for r in range(3):
            ds.render(cp, out=parts, aa=1, tile_rows=4, rank=r, nranks=3, wavefront=True, wave_stream=stream)
        assert np.array_equal(full, parts), stream
    # the same frame again and again through one scene handle, the two wavefronts alternating: the streaming queues'
    # ready stamps are per frame, and its child slots must be sentinels again after a frame of the passes
    for k in range(4):
        b, _ = ds.render(cp, aa=1, wavefront=True, wave_stream=(k != 2))
        assert np.array_equal(full, b), k
    ds.close()


def test_wavefront_queue_overflow_falls_back(pkg, prod):
    """Nested glass: every hit spawns a reflection AND a refraction ray, so a level needs more than the
    2 queue slots per pixel; the frame must still come back complete and identical (re-rendered by the
    recursive kernel inside rtc_scene_finish)."""
    w = prod.world()
    w.add_light((-4, 6, -8), (1, 1, 1))
    for k in range(4):
        s = prod.glass_sphere()
        s.set_transform(pkg.transform.scaling(3.0 - 0.5 * k, 3.0 - 0.5 * k, 3.0 - 0.5 * k))
        m = pkg.Material(color=(0.1, 0.1, 0.1), reflective=0.9, transparency=0.9, refractive_index=1.1 + 0.1 * k)
        s.set_material(m)
        w.add_object(s)
    c = prod.camera(96, 64, 0.9)
    c.set_transform(pkg.transform.view_transform((0, 0, -2.9), (0, 0, 0), (0, 1, 0)))
    ds = pkg.DeviceScene(pkg.FlatScene(w))
    a, sa = ds.render(c, wavefront=False)
    assert not sa.wave_fallback
    for stream in (False, True):
        b, sb = ds.render(c, wavefront=True, wave_stream=stream)
        assert sb.wave_fallback, stream                       # the overflow path really ran ...
        assert np.array_equal(a, b) and sa.rays == sb.rays    # ... and the frame is whole
    assert sa.rays_reflect + sa.rays_refract > 2 * 2 * 96 * 64  # some level really exceeded 2 per pixel
    # a frame that fits, right after one that did not, through the same handle
    c2 = prod.camera(96, 64, 0.9)
    c2.set_transform(pkg.transform.view_transform((0, 0, -40), (0, 0, 0), (0, 1, 0)))   # the glass covers 2 % of this frame
    a2, _ = ds.render(c2, wavefront=False)
    for stream in (True, False, True):
        b2, sb2 = ds.render(c2, wavefront=True, wave_stream=stream)
        assert np.array_equal(a2, b2) and (stream or not sb2.wave_fallback), stream
    ds.close()


def test_reference_known_answers_through_cabi(pkg, prod):
    """world.rs:186-234,249-274 and camera.rs:328-337 through rtc_trace_rays / rtc_is_shadowed."""
    w, c = scenes.default_world(prod)
    ds = pkg.DeviceScene(pkg.FlatScene(w))
    rgb, t = ds.trace_rays([[0, 0, -5], [0, 0, -5]], [[0, 0, 1], [0, 1, 0]])
    assert abs(t[0] - 4.0) < 1e-5 and t[1] == -1.0
    assert np.allclose(rgb[0], [0.38066, 0.47583, 0.2855], atol=1e-5) and np.all(rgb[1] == 0)
    assert ds.is_shadowed([0, 10, 0]) is False
    assert ds.is_shadowed([10, -10, 10]) is True
    assert ds.is_shadowed([-20, 20, -20]) is False
    assert ds.is_shadowed([-2, 2, -2]) is False
    canvas, _ = ds.render(c)
    assert np.allclose(canvas[5, 5], [0.38066, 0.47583, 0.2855], atol=1e-5)
    ds.close()


def test_n1_n2_table_through_cabi(pkg, prod, orc):
    """intersection.rs:298-335: three nested glass spheres; n1/n2 at each of the six crossings, by
    probing from just before each crossing."""
    def build(api):
        T = pkg.transform
        w = api.world()
        w.add_light((-10, 10, -10), (1, 1, 1))
        w.add_object(api.sphere().set_transform(T.scaling(2, 2, 2)).set_material(pkg.Material(transparency=1.0, refractive_index=1.5)))
        w.add_object(api.sphere().set_transform(T.translation(0, 0, -0.25)).set_material(pkg.Material(transparency=1.0, refractive_index=2.0)))
        w.add_object(api.sphere().set_transform(T.translation(0, 0, 0.25)).set_material(pkg.Material(transparency=1.0, refractive_index=2.5)))
        return w
    ds = pkg.DeviceScene(pkg.FlatScene(build(prod)))
    wo = build(orc.api)
    expected = [(1.0, 1.5), (1.5, 2.0), (2.0, 2.5), (2.5, 2.5), (2.5, 1.5), (1.5, 1.0)]
    crossings = [2.0, 2.75, 3.25, 4.75, 5.25, 6.0]
    for tc, (n1, n2) in zip(crossings, expected):
        o = [0, 0, -4 + tc - 0.1]
        got = ds.probe_hit(o, [0, 0, 1])
        want = orc.probe_hit(wo, o, [0, 0, 1])
        assert got[0] == 1.0 and abs(got[1] - 0.1) < 1e-9
        assert (got[17], got[18]) == (n1, n2), (tc, got[17], got[18])
        assert np.array_equal(got, want)
    ds.close()


def test_random_ray_batch_matches_oracle(pkg, prod, orc):
    """Seeded random rays through World::color_at on the CSG/primitive scene: colours and hit t."""
    wp, _ = scenes.primitives_csg(prod, 16, 9)
    wo, _ = scenes.primitives_csg(orc.api, 16, 9)
    rng = np.random.default_rng(20261104)
    n = 4000
    o = rng.uniform(-3, 3, size=(n, 3)) + np.array([0, 2.0, -2.0])
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    ds = pkg.DeviceScene(pkg.FlatScene(wp))
    rgb, t = ds.trace_rays(o, d)
    rgb0, t0 = orc.trace_rays(wo, o, d)
    assert np.array_equal(t, t0)
    assert np.max(np.abs(rgb - rgb0)) <= EXPECT_LIN
    ds.close()


def test_row_range_and_cyclic_tiles(pkg, prod):
    """rtc_opts row ranges and rank/nranks tile ownership partition the frame exactly."""
    wp, cp = scenes.reflect_refract(prod, 64, 40)
    ds = pkg.DeviceScene(pkg.FlatScene(wp))
    full, sf = ds.render(cp)
    part = np.full_like(full, -7.0)
    ds.render(cp, out=part, rows=(8, 24))
    assert np.array_equal(part[8:24], full[8:24]) and np.all(part[:8] == -7.0) and np.all(part[24:] == -7.0)
    acc = np.full_like(full, -7.0)
    rays = 0
    for r in range(3):
        _, s = ds.render(cp, out=acc, tile_rows=4, rank=r, nranks=3)
        rays += s.rays
    assert np.array_equal(acc, full) and rays == sf.rays
    ds.close()


def test_canvas_push_tiles(pkg):
    """rtc_canvas_push_tiles: exactly the tiles rank r of n owns (ragged last tile included) move from a local
    canvas into the target canvas; everything else is left alone."""
    import torch
    lib = pkg.cuda_lib()
    W, H, T = 40, 37, 8   # 4 whole tiles + a ragged one of 5 rows
    n = W * H * 3
    h = C.c_void_p()
    pkg._check(lib.rtc_canvas_create(0, n, C.byref(h)))
    host = torch.empty(n, dtype=torch.float64)
    for nranks in (1, 2, 3):
        for rank in range(nranks):
            local = torch.arange(n, dtype=torch.float64, device="cuda") + 1.0
            zero = torch.zeros(n, dtype=torch.float64, device="cuda")
            # clear the target through the same call (rank 0 of 1 owns everything)
            pkg._check(lib.rtc_canvas_push_tiles(h, zero.data_ptr(), W, H, T, 0, 1, None))
            pkg._check(lib.rtc_canvas_push_tiles(h, local.data_ptr(), W, H, T, rank, nranks, None))
            torch.cuda.synchronize()
            pkg._check(lib.rtc_canvas_download(h, host.data_ptr(), 0, n))
            got = host.numpy().reshape(H, W * 3)
            want = np.zeros_like(got)
            src = local.cpu().numpy().reshape(H, W * 3)
            for y in range(H):
                if (y // T) % nranks == rank:
                    want[y] = src[y]
            assert np.array_equal(got, want), (nranks, rank)
    lib.rtc_canvas_destroy(h)


def test_empty_world_and_no_lights(pkg, prod):
    w = prod.world()
    c = prod.camera(8, 4, 1.0)
    got, st, _ = pkg.render(c, w)
    assert np.all(got == 0.0) and st.rays == 32
    w.add_object(prod.sphere())
    c.set_transform(pkg.transform.view_transform((0, 0, -5), (0, 0, 0), (0, 1, 0)))
    got, st, _ = pkg.render(c, w)   # objects but no lights: surface = Σ over zero lights = black
    assert np.all(got == 0.0) and st.rays_shadow == 0


def test_device_quantize_kernel(pkg):
    import torch
    vals = np.array([-0.5, 0.0, 0.5, 1.0, 1.5, 0.0019607, 0.998, float("nan"), float("inf"), -float("inf"), 0.5 / 255, 1.5 / 255, 2.5 / 255])
    x = torch.tensor(vals, dtype=torch.float64, device="cuda")
    y = torch.zeros(len(vals), dtype=torch.uint8, device="cuda")
    assert pkg.cuda_lib().rtc_quantize_device(0, x.data_ptr(), len(vals), y.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(y.cpu().numpy(), pkg.quantize(vals))
    assert list(y.cpu().numpy()[:5]) == [0, 0, 128, 255, 255]   # ppm.rs:105-124


def test_pinned_canvas_is_written_zero_copy(pkg, prod):
    """rtc_render into a PINNED host canvas: the kernel stores its pixels straight into it over PCIe
    (no device canvas, no D2H copy) — must equal the staged path (pageable canvas) bit for bit."""
    import torch
    w, c = scenes.reflect_refract(prod, 320, 180)
    ds = pkg.DeviceScene(pkg.FlatScene(w))
    staged, st0 = ds.render(c)
    pinned = torch.full((180, 320, 3), -7.0, dtype=torch.float64).pin_memory()
    cam, opts, st = pkg.camera_to_rtc(c), pkg.make_opts(), pkg.RtcStats()
    pkg._check(pkg.cuda_lib().rtc_render(ds.h, C.byref(cam), C.byref(opts), pinned.data_ptr(), C.byref(st)))
    assert np.array_equal(pinned.numpy(), staged)
    assert st.rays == st0.rays
    # row range: only rows [40,100) are written
    pinned.fill_(-7.0)
    opts = pkg.make_opts(rows=(40, 100))
    pkg._check(pkg.cuda_lib().rtc_render(ds.h, C.byref(cam), C.byref(opts), pinned.data_ptr(), C.byref(st)))
    got = pinned.numpy()
    assert np.array_equal(got[40:100], staged[40:100]) and np.all(got[:40] == -7.0) and np.all(got[100:] == -7.0)
    ds.close()


@pytest.mark.gpu
def test_normalize_is_ieee_exact(pkg):
    """The device normalize() divides three components by one magnitude through a shared refined
    reciprocal (rt_math.cuh Divisor). Every quotient must be the correctly rounded IEEE one, i.e. what
    numpy's float64 `/` gives for the same magnitude — over ordinary, tiny, huge, zero, denormal,
    infinite and NaN inputs."""
    rng = np.random.default_rng(2024)
    parts = [rng.standard_normal((400000, 3)),
             rng.standard_normal((200000, 3)) * np.exp2(rng.integers(-40, 40, size=(200000, 3))),
             rng.standard_normal((50000, 3)) * np.exp2(rng.integers(-510, -480, size=(50000, 1))),   # squares underflow
             rng.standard_normal((50000, 3)) * np.exp2(rng.integers(480, 510, size=(50000, 1))),     # squares near overflow
             rng.integers(-3, 4, size=(50000, 3)).astype(np.float64),                                # exact zeros, small ints
             np.array([[0, 0, 0], [0.0, -0.0, 1.0], [np.inf, 1, 1], [np.nan, 1, 2], [5e-324, 0, 0], [1e308, 1e308, 1e308],
                       [1e-320, 1e-310, 0], [1, 2, 2], [3, 4, 0], [1e-160, 1e-160, 1e-160]], dtype=np.float64)]
    v = np.ascontiguousarray(np.concatenate(parts))
    out = np.empty_like(v)
    assert pkg.cuda_lib().rtc_debug_normalize(0, len(v), v.ctypes.data, out.ctypes.data) == 0
    with np.errstate(all="ignore"):
        m = np.sqrt(v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1] + v[:, 2] * v[:, 2])    # vector.rs:21, same order
        want = v / m[:, None]
    same = (out.view(np.uint64) == want.view(np.uint64)) | (np.isnan(out) & np.isnan(want))
    bad = np.argwhere(~same)
    assert len(bad) == 0, (len(bad), v[bad[:5, 0]], out[bad[:5, 0]], want[bad[:5, 0]])
