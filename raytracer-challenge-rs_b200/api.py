"""ctypes binding of the handle-based scene API.

`SceneApi(lib, prefix)` wraps one shared library that exports the scene-building calls
(`<prefix>shape_new`, `<prefix>group_add_child`, …). The product's host library exports them as
`rth_*` (host/rth_capi.cpp); the parity tests bind the CPU oracle's `orc_*` with the same class so
a scene script builds the identical scene in both. This module never loads the oracle itself.

The objects mirror the reference's API names: World / Camera / shapes / Material / patterns
(raytracer/src/{world,camera,material}.rs, geometry/shape/*.rs, pattern/mod.rs).
"""
import ctypes as C
import os

import numpy as np

SPHERE, PLANE, CUBE, CYLINDER, CONE, TRIANGLE, SMOOTH_TRIANGLE, GROUP, CSG, TEST_SHAPE = range(10)
PAT_TEST, PAT_STRIPE, PAT_GRADIENT, PAT_RING, PAT_CHECKERS = range(5)
UNION, INTERSECTION, DIFFERENCE = range(3)

_IDENTITY = [1.0, 0, 0, 0, 0, 1.0, 0, 0, 0, 0, 1.0, 0, 0, 0, 0, 1.0]


class CMaterial(C.Structure):
    _fields_ = [("color", C.c_double * 3), ("ambient", C.c_double), ("diffuse", C.c_double),
                ("specular", C.c_double), ("shininess", C.c_double), ("reflective", C.c_double),
                ("transparency", C.c_double), ("refractive_index", C.c_double),
                ("has_pattern", C.c_int32), ("pattern_kind", C.c_int32),
                ("pattern_a", C.c_double * 3), ("pattern_b", C.c_double * 3),
                ("pattern_transform", C.c_double * 16)]


class Pattern:
    """pattern/mod.rs:17-91"""

    def __init__(self, kind=PAT_TEST, a=(0.0, 0.0, 0.0), b=(0.0, 0.0, 0.0), transform=None):
        self.kind, self.a, self.b = kind, tuple(map(float, a)), tuple(map(float, b))
        # None = set_transform never called: the reference keeps the exact identity then, while
        # set_transform(identity) stores identity.inverse() with its negative zeros (pattern/mod.rs:22-36)
        self.transform = list(transform) if transform is not None else None

    def set_transform(self, m):
        self.transform = list(m)
        return self


def test_pattern():
    return Pattern(PAT_TEST)


def stripe_pattern(a, b):
    return Pattern(PAT_STRIPE, a, b)


def gradient_pattern(a, b):
    return Pattern(PAT_GRADIENT, a, b)


def ring_pattern(a, b):
    return Pattern(PAT_RING, a, b)


def checkers_pattern(a, b):
    return Pattern(PAT_CHECKERS, a, b)


class Material:
    """material.rs:11-36 (defaults of Material::default)."""

    def __init__(self, color=(1.0, 1.0, 1.0), ambient=0.1, diffuse=0.9, specular=0.9, shininess=200.0,
                 reflective=0.0, transparency=0.0, refractive_index=1.0, pattern=None):
        self.color = tuple(map(float, color))
        self.ambient, self.diffuse, self.specular, self.shininess = float(ambient), float(diffuse), float(specular), float(shininess)
        self.reflective, self.transparency, self.refractive_index = float(reflective), float(transparency), float(refractive_index)
        self.pattern = pattern

    def copy(self, **kw):
        m = Material(self.color, self.ambient, self.diffuse, self.specular, self.shininess, self.reflective,
                     self.transparency, self.refractive_index, self.pattern)
        for k, v in kw.items():
            if not hasattr(m, k):
                raise AttributeError(k)
            setattr(m, k, tuple(map(float, v)) if k == "color" else v)
        return m

    def to_c(self):
        c = CMaterial()
        c.color[:] = self.color
        c.ambient, c.diffuse, c.specular, c.shininess = self.ambient, self.diffuse, self.specular, self.shininess
        c.reflective, c.transparency, c.refractive_index = self.reflective, self.transparency, self.refractive_index
        p = self.pattern or Pattern()
        c.has_pattern = (1 if self.pattern is not None else 0) | (2 if p.transform is not None else 0)
        c.pattern_kind = p.kind
        c.pattern_a[:] = p.a
        c.pattern_b[:] = p.b
        c.pattern_transform[:] = p.transform if p.transform is not None else _IDENTITY
        return c


def _d(seq, n):
    arr = (C.c_double * n)(*[float(v) for v in seq])
    return arr


class Shape:
    def __init__(self, api, handle, owned=True):
        self.api, self.h, self.owned = api, handle, owned

    def _give(self):
        if not self.owned:
            raise RuntimeError("shape already given to a parent/world")
        self.owned = False
        return self.h

    def __del__(self):
        if getattr(self, "owned", False) and self.h:
            self.api._f("shape_free")(self.h)
            self.h = None

    # Shape trait, geometry/mod.rs:38-108
    def set_transform(self, m):
        if self.api._f("shape_set_transform")(self.h, _d(m, 16)) != 0:
            raise ValueError(self.api.last_error())
        return self

    def set_material(self, material):
        c = material.to_c()
        if self.api._f("shape_set_material")(self.h, C.byref(c)) != 0:
            raise ValueError(self.api.last_error())
        return self

    def no_shadow(self):
        self.api._f("shape_no_shadow")(self.h)
        return self

    def divide(self, threshold):
        self.api._f("shape_divide")(self.h, int(threshold))
        return self

    def bounds(self):
        out = (C.c_double * 6)()
        self.api._f("shape_bounds")(self.h, out)
        return list(out)

    def transform(self):
        m, inv = (C.c_double * 16)(), (C.c_double * 16)()
        self.api._f("shape_get_transform")(self.h, m, inv)
        return list(m), list(inv)

    def kind(self):
        return self.api._f("shape_kind")(self.h)

    def child_count(self):
        return self.api._f("shape_child_count")(self.h)

    def child(self, i):
        return Shape(self.api, self.api._f("shape_child")(self.h, i), owned=False)

    # Group, group.rs:128-133
    def add_child(self, shape):
        if self.api._f("group_add_child")(self.h, shape._give()) != 0:
            raise ValueError(self.api.last_error())
        return self

    def add_triangles(self, points, normals=None):
        """points: (n,3,3) float64 p1,p2,p3; normals: (n,3,3) n1,n2,n3 or None (flat triangles)."""
        p = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 9)
        nptr = None
        if normals is not None:
            nn = np.ascontiguousarray(normals, dtype=np.float64).reshape(-1, 9)
            assert nn.shape == p.shape
            nptr = nn.ctypes.data_as(C.c_void_p)
        if self.api._f("group_add_triangles")(self.h, p.shape[0], p.ctypes.data_as(C.c_void_p), nptr) != 0:
            raise ValueError(self.api.last_error())
        return self


class World:
    """world.rs:18-29,83-93"""

    def __init__(self, api, handle=None):
        self.api = api
        self.h = handle if handle else api._f("world_new")()
        self.n_lights = 0

    def __del__(self):
        if getattr(self, "h", None):
            self.api._f("world_free")(self.h)
            self.h = None

    def add_object(self, shape):
        self.api._f("world_add_object")(self.h, shape._give())
        return self

    def add_light(self, position, intensity):
        self.api._f("world_add_light")(self.h, _d(position, 3), _d(intensity, 3))
        self.n_lights += 1
        return self


class Camera:
    """camera.rs:20-55,128-131"""

    def __init__(self, api, hsize, vsize, field_of_view, handle=None):
        self.api, self.hsize, self.vsize, self.field_of_view = api, int(hsize), int(vsize), float(field_of_view)
        self.h = handle if handle else api._f("camera_new")(self.hsize, self.vsize, self.field_of_view)
        self.aa_samples = 1

    def __del__(self):
        if getattr(self, "h", None):
            self.api._f("camera_free")(self.h)
            self.h = None

    def set_transform(self, m):
        if self.api._f("camera_set_transform")(self.h, _d(m, 16)) != 0:
            raise ValueError(self.api.last_error())
        return self

    def params(self):
        out = (C.c_double * 19)()
        self.api._f("camera_params")(self.h, out)
        return {"pixel_size": out[0], "half_width": out[1], "half_height": out[2], "inverse": list(out[3:19])}


class SceneApi:
    """Binds `<prefix>*` scene-building calls of one shared library."""

    def __init__(self, lib, prefix):
        self.lib, self.prefix = lib, prefix
        self._cache = {}
        vp, d, sz, i32 = C.c_void_p, C.c_double, C.c_size_t, C.c_int
        pd = C.c_void_p
        sigs = {
            "last_error": (C.c_char_p, []),
            "shape_new": (vp, [i32, d, d, i32]),
            "triangle_new": (vp, [pd]),
            "smooth_triangle_new": (vp, [pd, pd]),
            "csg_new": (vp, [i32, vp, vp]),
            "shape_free": (None, [vp]),
            "shape_kind": (i32, [vp]),
            "group_add_child": (i32, [vp, vp]),
            "group_add_triangles": (i32, [vp, sz, pd, pd]),
            "shape_set_transform": (i32, [vp, pd]),
            "shape_get_transform": (None, [vp, pd, pd]),
            "shape_set_material": (i32, [vp, vp]),
            "shape_no_shadow": (None, [vp]),
            "shape_divide": (None, [vp, sz]),
            "shape_bounds": (None, [vp, pd]),
            "shape_child_count": (sz, [vp]),
            "shape_child": (vp, [vp, sz]),
            "obj_parse": (vp, [C.c_char_p, sz]),
            "world_new": (vp, []),
            "world_free": (None, [vp]),
            "world_add_object": (None, [vp, vp]),
            "world_add_light": (None, [vp, pd, pd]),
            "camera_new": (vp, [sz, sz, d]),
            "camera_free": (None, [vp]),
            "camera_set_transform": (i32, [vp, pd]),
            "camera_params": (None, [vp, pd]),
        }
        # front ends only the host library has (the oracle is handed their output)
        optional = {
            "yaml_load": (i32, [C.c_char_p, sz, C.c_uint32, C.c_char_p, C.POINTER(vp), C.POINTER(vp), C.POINTER(i32)]),
            "camera_dims": (None, [vp, C.POINTER(sz), C.POINTER(sz), C.POINTER(d)]),
            "world_counts": (None, [vp, C.POINTER(sz), C.POINTER(sz)]),
            "world_object": (vp, [vp, sz]),
        }
        for name, sig in optional.items():
            if hasattr(lib, prefix + name):
                sigs[name] = sig
        for name, (res, args) in sigs.items():
            fn = getattr(lib, prefix + name)
            fn.restype, fn.argtypes = res, args
            self._cache[name] = fn

    def _f(self, name):
        return self._cache[name]

    def last_error(self):
        return (self._f("last_error")() or b"").decode()

    # constructors named after the reference's shapes
    def _shape(self, kind, mn=float("-inf"), mx=float("inf"), closed=False):
        h = self._f("shape_new")(kind, float(mn), float(mx), 1 if closed else 0)
        if not h:
            raise ValueError(self.last_error())
        return Shape(self, h)

    def sphere(self):
        return self._shape(SPHERE)

    def glass_sphere(self):
        return self.sphere().set_material(Material(transparency=1.0, refractive_index=1.5))

    def plane(self):
        return self._shape(PLANE)

    def cube(self):
        return self._shape(CUBE)

    def cylinder(self, minimum=float("-inf"), maximum=float("inf"), closed=False):
        return self._shape(CYLINDER, minimum, maximum, closed)

    def cone(self, minimum=float("-inf"), maximum=float("inf"), closed=False):
        return self._shape(CONE, minimum, maximum, closed)

    def group(self):
        return self._shape(GROUP)

    def triangle(self, p1, p2, p3):
        return Shape(self, self._f("triangle_new")(_d(list(p1) + list(p2) + list(p3), 9)))

    def smooth_triangle(self, p1, p2, p3, n1, n2, n3):
        return Shape(self, self._f("smooth_triangle_new")(_d(list(p1) + list(p2) + list(p3), 9),
                                                          _d(list(n1) + list(n2) + list(n3), 9)))

    def csg(self, op, left, right):
        return Shape(self, self._f("csg_new")(op, left._give(), right._give()))

    def obj_group(self, text):
        """obj_parser: parse_obj_file(...).as_group()"""
        data = text.encode() if isinstance(text, str) else text
        h = self._f("obj_parse")(data, len(data))
        if not h:
            raise ValueError(self.last_error())
        return Shape(self, h)

    def yaml_scene(self, text, extended=False, base_dir=None):
        """scene-parser SceneParser::load_file (lib.rs:90-119) → (World, Camera | None, skipped `add` kinds)"""
        data = text.encode() if isinstance(text, str) else text
        w, c, skipped = C.c_void_p(), C.c_void_p(), C.c_int(0)
        bd = os.fsencode(base_dir) if base_dir else None
        if self._f("yaml_load")(data, len(data), 1 if extended else 0, bd, C.byref(w), C.byref(c), C.byref(skipped)) != 0:
            raise ValueError(self.last_error())
        world = World(self, handle=w.value)
        no, nl = C.c_size_t(), C.c_size_t()
        self._f("world_counts")(world.h, C.byref(no), C.byref(nl))
        world.n_objects, world.n_lights = no.value, nl.value
        camera = None
        if c.value:
            hs, vs, fov = C.c_size_t(), C.c_size_t(), C.c_double()
            self._f("camera_dims")(c.value, C.byref(hs), C.byref(vs), C.byref(fov))
            camera = Camera(self, hs.value, vs.value, fov.value, handle=c.value)
        return world, camera, skipped.value

    def world(self):
        return World(self)

    def camera(self, hsize, vsize, fov):
        return Camera(self, hsize, vsize, fov)
