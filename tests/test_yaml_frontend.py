"""The YAML scene front end (host/rth_yaml.hpp) against scene-parser/src/lib.rs semantics.

CPU tier: (1) a scene file written for this test that exercises every construct the reference's
parser reads (define/extend, named and inline transforms, item*acc order, stripes/checkers/unknown
pattern, ignored pattern transform, unknown `add` kinds, comments) must flatten to the same tables,
byte for byte, as the same scene built through the constructor API, and render (kernel source on the
host) bit-identically to the oracle; (2) when /root/reference is present (this container only) its two
example files must flatten to the same tables as the transcriptions in scenes.py, which are what
bench.py and the parity tests use. GPU tier: the test scene through rtc_render vs the oracle."""
import ctypes as C
import importlib
import math
import os

import numpy as np
import pytest

from conftest import compare_images

scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
api_mod = importlib.import_module("raytracer-challenge-rs_b200.api")
T = importlib.import_module("raytracer-challenge-rs_b200.transform")

SCENE = """\
# a room corner with three balls and two boxes
- add: camera
  width: 96
  height: 54
  field-of-view: 0.9
  from: [-3, 2.5, -6]
  to: [0, 1, 0]
  up: [0, 1, 0]

- add: light
  at: [-6, 8, -7]
  intensity: [1, 0.9, 0.8]

- define: base
  value:
    color: [0.9, 0.9, 1]
    diffuse: 0.6
    specular: 0.3   # trailing comment
    reflective: 0.2

- define: glassy
  extend: base
  value:
    color: [0.1, 0.1, 0.1]
    transparency: 0.8
    refractive-index: 1.5
    shininess: 250

- define: lift
  value:
    - [translate, 0, 1, 0]

- define: small-lifted
  value:
    - [scale, 0.5, 0.5, 0.5]
    - lift
    - [rotate-y, 0.5]

- add: plane
  material:
    pattern:
      type: checkers
      colors:
        - [0.2, 0.2, 0.2]
        - [0.8, 0.8, 0.8]
      transform:
        - [scale, 0.25, 0.25, 0.25]
    specular: 0
    reflective: 0.3

- add: plane
  transform:
    - [rotate-x, 1.5708]
    - [translate, 0, 0, 4]
  material:
    pattern:
      type: stripes
      colors:
        - [0.5, 0.2, 0.2]
        - [0.2, 0.2, 0.5]
    ambient: 0.2

- add: sphere
  transform:
    - lift
  material: glassy

- add: sphere
  transform:
    - small-lifted
    - [translate, 2, 0, -1]
  material: base

- add: cube
  transform:
    - [scale, 0.6, 0.6, 0.6]
    - [rotate-z, 0.3]
    - [translate, -2, 0.6, 1]
  material:
    color: [1, 0.4, 0.1]
    pattern:
      type: gradient
      colors:
        - [1, 0, 0]
        - [0, 0, 1]

- add: torus
  transform:
    - lift

- add: sphere
  transform:
    - [scale, 0.3, 0.3, 0.3]
    - [translate, -0.8, 0.3, -2]
  material:
    color: [0.2, 1, 0.3]
    shininess: 20
"""


def build_by_hand(api):
    """SCENE through the constructor API, resolved the way lib.rs does."""
    M, w = api_mod.Material, api.world()
    w.add_light((-6, 8, -7), (1, 0.9, 0.8))
    base = M(color=(0.9, 0.9, 1), diffuse=0.6, specular=0.3, reflective=0.2)
    glassy = base.copy(color=(0.1, 0.1, 0.1), transparency=0.8, refractive_index=1.5, shininess=250)
    lift = T.chain(T.translation(0, 1, 0))
    small_lifted = T.chain(T.scaling(0.5, 0.5, 0.5), lift, T.rotation_y(0.5))
    w.add_object(api.plane().set_material(M(pattern=api_mod.checkers_pattern((0.2, 0.2, 0.2), (0.8, 0.8, 0.8)),
                                            specular=0, reflective=0.3)))
    w.add_object(api.plane().set_transform(T.chain(T.rotation_x(1.5708), T.translation(0, 0, 4)))
                 .set_material(M(pattern=api_mod.stripe_pattern((0.5, 0.2, 0.2), (0.2, 0.2, 0.5)), ambient=0.2)))
    w.add_object(api.sphere().set_transform(T.chain(lift)).set_material(glassy))
    w.add_object(api.sphere().set_transform(T.chain(small_lifted, T.translation(2, 0, -1))).set_material(base))
    # unknown pattern type → Pattern::default() = the test pattern (lib.rs:474)
    w.add_object(api.cube().set_transform(T.chain(T.scaling(0.6, 0.6, 0.6), T.rotation_z(0.3), T.translation(-2, 0.6, 1)))
                 .set_material(M(color=(1, 0.4, 0.1), pattern=api_mod.test_pattern())))
    w.add_object(api.sphere().set_transform(T.chain(T.scaling(0.3, 0.3, 0.3), T.translation(-0.8, 0.3, -2)))
                 .set_material(M(color=(0.2, 1, 0.3), shininess=20)))
    cam = api.camera(96, 54, 0.9)
    cam.set_transform(T.view_transform((-3, 2.5, -6), (0, 1, 0), (0, 1, 0)))
    return w, cam


def flat_tables(pkg, world):
    """Every table of the flattened scene as bytes (what rtc_scene_create is handed)."""
    f = pkg.FlatScene(world)
    d = f.desc
    node_sz = C.sizeof(pkg.RtcNode)
    mat_sz, pat_sz, light_sz = 8 * 10 + 8, 8 + 8 * 18, 8 * 6
    parts = {
        "nodes": (d.nodes, d.n_nodes * node_sz), "xforms": (d.xforms, d.n_xforms * 96),
        "boxes": (d.boxes, d.n_boxes * 48), "limits": (d.limits, d.n_limits * 16),
        "materials": (d.materials, d.n_materials * mat_sz), "patterns": (d.patterns, d.n_patterns * pat_sz),
        "lights": (d.lights, d.n_lights * light_sz),
    }
    out = {}
    for k, (ptr, nbytes) in parts.items():
        addr = ptr if isinstance(ptr, int) else C.cast(ptr, C.c_void_p).value
        out[k] = C.string_at(addr, nbytes) if nbytes else b""
    return out


def assert_same_scene(pkg, wa, ca, wb, cb):
    ta, tb = flat_tables(pkg, wa), flat_tables(pkg, wb)
    for k in ta:
        assert ta[k] == tb[k], k
    assert (ca.hsize, ca.vsize, ca.field_of_view) == (cb.hsize, cb.vsize, cb.field_of_view)
    assert ca.params() == cb.params()


def test_yaml_scene_equals_constructor_scene(pkg, prod):
    wy, cy, skipped = prod.yaml_scene(SCENE)
    assert skipped == 1                       # `add: torus` → "unhandled element" (lib.rs:132)
    assert (wy.n_objects, wy.n_lights) == (6, 1)
    wh, ch = build_by_hand(prod)
    assert_same_scene(pkg, wy, cy, wh, ch)


def test_yaml_scene_renders_like_the_oracle(pkg, prod, orc, emul):
    wy, cy, _ = prod.yaml_scene(SCENE)
    wo, co = build_by_hand(orc.api)
    want, iw = orc.render(co, wo)
    got, ig = emul.render(cy, wy)
    assert ig["rays"] == iw["rays"]
    frac, max8, lin = compare_images(got, want, orc.quantize)
    assert frac == 1.0 and max8 == 0 and lin <= 1e-12


@pytest.mark.parametrize("text,msg", [
    ("- add: camera\n  width: 10\n  height: 10\n  field-of-view: 1\n  from: [0,0,-5]\n  to: [0,0,0]\n  up: [0,1,0]\n",
     "field-of-view"),                                                         # integer fov (lib.rs:395-397)
    ("- add: camera\n  width: 10\n  height: 10\n  field-of-view: 1.0\n  from: [0,0,-5]\n  to: [0,0,0]\n", "up"),
    ("- add: sphere\n  transform:\n    - [shear, 1, 2, 3]\n", "shear"),     # unknown transform (lib.rs:362)
    ("- add: sphere\n  transform:\n    - nowhere\n", "nowhere"),              # undefined named transform
    ("- add: sphere\n  material: nothing\n", "nothing"),                      # undefined material
    ("- define: m2\n  extend: m1\n  value:\n    ambient: 1\n", "m1"),          # extend of an unknown material
    ("- add: light\n  at: [0, 1]\n  intensity: [1, 1, 1]\n", "at"),            # wrong arity (lib.rs:505-507)
    ("- add: sphere\n  material:\n    ambient: high\n", "ambient"),            # not a number (lib.rs:495)
    ("add: sphere\n", "list"),                                                 # top level must be an array
])
def test_yaml_errors(prod, text, msg):
    with pytest.raises(ValueError, match=msg):
        prod.yaml_scene(text)


def test_yaml_define_order_and_last_light(pkg, prod):
    """defines are resolved before adds wherever they stand in the file (lib.rs:97-114); a second
    camera replaces the first (lib.rs:126)."""
    text = ("- add: sphere\n  material: late\n  transform:\n    - up2\n"
            "- add: camera\n  width: 4\n  height: 4\n  field-of-view: 0.5\n  from: [0,0,-5]\n  to: [0,0,0]\n  up: [0,1,0]\n"
            "- add: camera\n  width: 8\n  height: 6\n  field-of-view: 0.75\n  from: [0,0,-5]\n  to: [0,0,0]\n  up: [0,1,0]\n"
            "- define: late\n  value:\n    ambient: 0.5\n"
            "- define: up2\n  value:\n    - [translate, 0, 2, 0]\n")
    w, cam, _ = prod.yaml_scene(text)
    assert (cam.hsize, cam.vsize, cam.field_of_view) == (8, 6, 0.75)
    wh = prod.world()
    wh.add_object(prod.sphere().set_transform(T.translation(0, 2, 0)).set_material(api_mod.Material(ambient=0.5)))
    ta, tb = flat_tables(pkg, w), flat_tables(pkg, wh)
    assert ta == tb


def test_yaml_extension_kinds(pkg, prod):
    text = "- add: cylinder\n  min: 0\n  max: 2\n  closed: true\n- add: cone\n  min: -1\n  max: 0\n"
    w, _, skipped = prod.yaml_scene(text)
    assert (w.n_objects, skipped) == (0, 2)   # reference behaviour: not handled
    w, _, skipped = prod.yaml_scene(text, extended=True)
    assert (w.n_objects, skipped) == (2, 0)
    wh = prod.world()
    wh.add_object(prod.cylinder(0, 2, True))
    wh.add_object(prod.cone(-1, 0, False))
    assert flat_tables(pkg, w) == flat_tables(pkg, wh)


EXAMPLES = "/root/reference/scene-parser/examples"


@pytest.mark.skipif(not os.path.isdir(EXAMPLES), reason="reference checkout not present (GPU box)")
@pytest.mark.parametrize("name,builder,size", [("reflect-refract", "reflect_refract", (2560, 1440)),
                                               ("cover", "cover", (4000, 4000))])
def test_reference_examples_equal_transcriptions(pkg, prod, name, builder, size):
    with open(os.path.join(EXAMPLES, name + ".yml")) as f:
        wy, cy, skipped = prod.yaml_scene(f.read())
    assert skipped == 0 and (cy.hsize, cy.vsize) == size
    wt, ct = getattr(scenes, builder)(prod, width=size[0], height=size[1])
    assert_same_scene(pkg, wy, cy, wt, ct)


@pytest.mark.gpu
def test_gpu_yaml_scene_matches_oracle(pkg, prod, orc):
    wy, cy, _ = prod.yaml_scene(SCENE)
    wo, co = build_by_hand(orc.api)
    want, iw = orc.render(co, wo)
    got, stats, _ = pkg.render(cy, wy)
    assert stats.rays == iw["rays"]
    frac, max8, lin = compare_images(got, want, orc.quantize)
    assert frac >= 0.999 and max8 <= 1 and lin <= 1e-6


EXTENDED = """\
- add: camera
  width: 64
  height: 48
  field-of-view: 1.0
  from: [0, 3, -7]
  to: [0, 1, 0]
  up: [0, 1, 0]
- add: light
  at: [-5, 8, -6]
  intensity: [1, 1, 1]
- define: red
  value:
    color: [1, 0.2, 0.2]
- add: plane
- add: group
  transform:
    - [rotate-y, 0.4]
    - [translate, -1.5, 0, 0]
  children:
    - add: cylinder
      min: 0
      max: 1.5
      closed: true
      transform:
        - [scale, 0.5, 1, 0.5]
    - add: cone
      min: -1
      max: 0
      transform:
        - [translate, 0, 2.5, 0]
      material: red
    - add: group
      children:
        - add: sphere
          shadow: false
          transform:
            - [scale, 0.3, 0.3, 0.3]
            - [translate, 1, 0.3, -1]
- add: csg
  operation: difference
  left:
    add: cube
    material: red
  right:
    add: sphere
    transform:
      - [scale, 1.3, 1.3, 1.3]
  transform:
    - [scale, 0.8, 0.8, 0.8]
    - [translate, 1.5, 0.8, 0]
- add: obj
  file: tri.obj
  divide: 1
  transform:
    - [translate, 0, 0.5, 2]
"""

TRI_OBJ = "v -1 0 0\nv 1 0 0\nv 0 2 0\nv 0 0 1\nf 1 2 3\nf 1 2 4\n"


def build_extended_by_hand(api):
    M = api_mod.Material
    red = M(color=(1, 0.2, 0.2))
    w = api.world()
    w.add_light((-5, 8, -6), (1, 1, 1))
    w.add_object(api.plane())
    g = api.group()
    g.add_child(api.cylinder(0, 1.5, True).set_transform(T.scaling(0.5, 1, 0.5)))
    g.add_child(api.cone(-1, 0, False).set_transform(T.translation(0, 2.5, 0)).set_material(red))
    inner = api.group()
    inner.add_child(api.sphere().set_transform(T.chain(T.scaling(0.3, 0.3, 0.3), T.translation(1, 0.3, -1))).no_shadow())
    g.add_child(inner)
    g.set_transform(T.chain(T.rotation_y(0.4), T.translation(-1.5, 0, 0)))
    w.add_object(g)
    c = api.csg(api_mod.DIFFERENCE, api.cube().set_material(red), api.sphere().set_transform(T.scaling(1.3, 1.3, 1.3)))
    c.set_transform(T.chain(T.scaling(0.8, 0.8, 0.8), T.translation(1.5, 0.8, 0)))
    w.add_object(c)
    o = api.obj_group(TRI_OBJ)
    o.set_transform(T.translation(0, 0.5, 2))
    o.divide(1)
    w.add_object(o)
    cam = api.camera(64, 48, 1.0)
    cam.set_transform(T.view_transform((0, 3, -7), (0, 1, 0), (0, 1, 0)))
    return w, cam


def test_yaml_extended_kinds_equal_constructor_scene(pkg, prod, orc, emul, tmp_path):
    (tmp_path / "tri.obj").write_text(TRI_OBJ)
    wy, cy, skipped = prod.yaml_scene(EXTENDED, extended=True, base_dir=str(tmp_path))
    assert skipped == 0
    wh, ch = build_extended_by_hand(prod)
    assert_same_scene(pkg, wy, cy, wh, ch)
    # and without the extension the reference's parser skips all five
    w0, _, skipped0 = prod.yaml_scene(EXTENDED)
    assert (w0.n_objects, skipped0) == (1, 3)
    wo, co = build_extended_by_hand(orc.api)
    want, iw = orc.render(co, wo)
    got, ig = emul.render(cy, wy)
    assert ig["rays"] == iw["rays"]
    frac, max8, lin = compare_images(got, want, orc.quantize)
    assert frac == 1.0 and max8 == 0 and lin <= 1e-12


def _read_png(path):
    """Minimal PNG reader (8-bit RGB, filter 0 rows) — checks chunk CRCs and the zlib stream."""
    import struct
    import zlib
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, size = 8, b"", None
    while pos < len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        assert zlib.crc32(typ + body) == crc
        if typ == b"IHDR":
            w, h, depth, ctype, comp, filt, inter = struct.unpack(">IIBBBBB", body)
            assert (depth, ctype, comp, filt, inter) == (8, 2, 0, 0, 0)
            size = (w, h)
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(size[1], size[0] * 3 + 1)
    assert not raw[:, 0].any()
    return raw[:, 1:].reshape(size[1], size[0], 3)


def test_image_writers_round_trip(pkg, tmp_path):
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, size=(301, 233, 3), dtype=np.uint8)   # > 64 KiB: several stored blocks
    pkg.save_image(img, tmp_path / "a.png")
    assert np.array_equal(_read_png(tmp_path / "a.png"), img)
    pkg.save_image(img, tmp_path / "a.pnm")
    blob = open(tmp_path / "a.pnm", "rb").read()
    head = b"P6\n233 301\n255\n"
    assert blob.startswith(head) and blob[len(head):] == img.tobytes()


@pytest.mark.gpu
def test_gpu_render_scene_file_and_rgb8(pkg, prod, orc, tmp_path):
    """render_scene <yml> <png> end to end on the device: YAML → flatten → rtc_render_rgb8 → PNG equals
    the oracle's frame quantised by the reference's rule (png.rs:29-31)."""
    (tmp_path / "tri.obj").write_text(TRI_OBJ)
    (tmp_path / "scene.yml").write_text(EXTENDED)
    st = pkg.render_scene_file(tmp_path / "scene.yml", tmp_path / "out.png", extended=True)
    wo, co = build_extended_by_hand(orc.api)
    want, iw = orc.render(co, wo)
    assert st.rays == iw["rays"]
    got8 = _read_png(tmp_path / "out.png")
    want8 = orc.quantize(want)
    diff = np.abs(got8.astype(np.int16) - want8.astype(np.int16))
    assert (diff.max(axis=2) == 0).mean() >= 0.999 and diff.max() <= 1
    # the rgb8 entry point agrees with rendering doubles and quantising on the host
    wy, cy, _ = prod.yaml_scene(EXTENDED, extended=True, base_dir=str(tmp_path))
    with pkg.DeviceScene(pkg.FlatScene(wy)) as ds:
        f64, _ = ds.render(cy)
        u8, st8 = ds.render_rgb8(cy)
    assert np.array_equal(u8, pkg.quantize(f64))
    assert np.array_equal(u8, got8)
    # the binary itself
    import subprocess
    exe = os.path.join(os.path.dirname(pkg.__file__), "host", "render_scene")
    out = subprocess.run([exe, str(tmp_path / "scene.yml"), str(tmp_path / "bin.ppm"), "--extended"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    assert open(tmp_path / "bin.ppm", "rb").read() == pkg.canvas_to_ppm(u8 / 255.0)
