"""Scene scripts for the five BASELINE.json configs, written once against the
scene-building API (`api` = a SceneApi) so the product host library and the test oracle build the
identical scene. Each function returns (world, camera). Sources in the reference are cited per scene;
the YAML scenes are transcribed (object by object, in file order) because /root/reference does not
exist on the GPU box.
"""
import math

import numpy as np

from . import transform as T
from .api import (DIFFERENCE, Material, checkers_pattern, gradient_pattern, ring_pattern, stripe_pattern, test_pattern)

PI = math.pi


def default_world(api, width=11, height=11):
    """World::default (world.rs:137-151) with the camera of camera.rs:328-337."""
    w = api.world()
    w.add_light((-10, 10, -10), (1, 1, 1))
    w.add_object(api.sphere().set_material(Material(color=(0.8, 1.0, 0.6), diffuse=0.7, specular=0.2)))
    w.add_object(api.sphere().set_transform(T.scaling(0.5, 0.5, 0.5)))
    cam = api.camera(width, height, PI / 2.0)
    cam.set_transform(T.view_transform((0, 0, -5), (0, 0, 0), (0, 1, 0)))
    return w, cam


def chapter7(api, width=800, height=400):
    """Config 1 — the book's chapter-7 world: floor and walls are flattened spheres; camera and the
    three sphere placements as in raytracer/src/bin/first_scene.rs:57-78,110-115."""
    w = api.world()
    w.add_light((-10, 10, -10), (1, 1, 1))
    wall = Material(color=(1, 0.9, 0.9), specular=0.0)
    w.add_object(api.sphere().set_transform(T.scaling(10, 0.01, 10)).set_material(wall))
    w.add_object(api.sphere().set_transform(
        T.matmul(T.matmul(T.matmul(T.translation(0, 0, 5), T.rotation_y(-PI / 4.0)), T.rotation_x(PI / 2.0)),
                 T.scaling(10, 0.01, 10))).set_material(wall))
    w.add_object(api.sphere().set_transform(
        T.matmul(T.matmul(T.matmul(T.translation(0, 0, 5), T.rotation_y(PI / 4.0)), T.rotation_x(PI / 2.0)),
                 T.scaling(10, 0.01, 10))).set_material(wall))
    w.add_object(api.sphere().set_transform(T.translation(-0.5, 1, 0.5))
                 .set_material(Material(color=(0.1, 1, 0.5), diffuse=0.7, specular=0.3)))
    w.add_object(api.sphere().set_transform(T.matmul(T.translation(1.5, 0.5, -0.5), T.scaling(0.5, 0.5, 0.5)))
                 .set_material(Material(color=(0.5, 1, 0.1), diffuse=0.7, specular=0.3)))
    w.add_object(api.sphere().set_transform(T.matmul(T.translation(-1.5, 0.33, -0.75), T.scaling(0.33, 0.33, 0.33)))
                 .set_material(Material(color=(1, 0.8, 0.1), diffuse=0.7, specular=0.3)))
    cam = api.camera(width, height, PI / 3.0)
    cam.set_transform(T.view_transform((0, 1.5, -5), (0, 1, 0), (0, 1, 0)))
    return w, cam


def reflect_refract(api, width=1920, height=1080):
    """Config 2 — scene-parser/examples/reflect-refract.yml (objects in file order; width/height of
    :16-17 overridden). YAML transform lists apply first-to-last as item*acc
    (scene-parser/src/lib.rs:337-341); the pattern `transform:` key (:42-44) is ignored by the
    reference's parser (lib.rs:448-483), so the wall stripes carry the identity."""
    w = api.world()
    w.add_light((-4.9, 4.9, -1), (1, 1, 1))
    wall = Material(pattern=stripe_pattern((0.45, 0.45, 0.45), (0.55, 0.55, 0.55)), ambient=0, diffuse=0.4,
                    specular=0, reflective=0.3)
    # the checkered floor
    w.add_object(api.plane().set_transform(T.chain(T.rotation_y(0.31415)))
                 .set_material(Material(pattern=checkers_pattern((0.35, 0.35, 0.35), (0.65, 0.65, 0.65)),
                                        specular=0, reflective=0.4)))
    # the ceiling
    w.add_object(api.plane().set_transform(T.chain(T.translation(0, 5, 0)))
                 .set_material(Material(color=(0.8, 0.8, 0.8), ambient=0.3, specular=0)))
    # west, east, north, south walls
    w.add_object(api.plane().set_transform(T.chain(T.rotation_y(1.5708), T.rotation_z(1.5708), T.translation(-5, 0, 0))).set_material(wall))
    w.add_object(api.plane().set_transform(T.chain(T.rotation_y(1.5708), T.rotation_z(1.5708), T.translation(5, 0, 0))).set_material(wall))
    w.add_object(api.plane().set_transform(T.chain(T.rotation_x(1.5708), T.translation(0, 0, 5))).set_material(wall))
    w.add_object(api.plane().set_transform(T.chain(T.rotation_x(1.5708), T.translation(0, 0, -5))).set_material(wall))
    # background balls
    for s, tr, col in [(0.4, (4.6, 0.4, 1), (0.8, 0.5, 0.3)), (0.3, (4.7, 0.3, 0.4), (0.9, 0.4, 0.5)),
                       (0.5, (-1, 0.5, 4.5), (0.4, 0.9, 0.6)), (0.3, (-1.7, 0.3, 4.7), (0.4, 0.6, 0.9))]:
        w.add_object(api.sphere().set_transform(T.chain(T.scaling(s, s, s), T.translation(*tr)))
                     .set_material(Material(color=col, shininess=50)))
    # foreground balls: red sphere, blue glass sphere, green glass sphere
    w.add_object(api.sphere().set_transform(T.chain(T.translation(-0.6, 1, 0.6)))
                 .set_material(Material(color=(1, 0.3, 0.2), specular=0.4, shininess=5)))
    w.add_object(api.sphere().set_transform(T.chain(T.scaling(0.7, 0.7, 0.7), T.translation(0.6, 0.7, -0.6)))
                 .set_material(Material(color=(0, 0, 0.2), ambient=0, diffuse=0.4, specular=0.9, shininess=300,
                                        reflective=0.9, transparency=0.9, refractive_index=1.5)))
    w.add_object(api.sphere().set_transform(T.chain(T.scaling(0.5, 0.5, 0.5), T.translation(-0.7, 0.5, -0.8)))
                 .set_material(Material(color=(0, 0.2, 0), ambient=0, diffuse=0.4, specular=0.9, shininess=300,
                                        reflective=0.9, transparency=0.9, refractive_index=1.5)))
    cam = api.camera(width, height, 1.152)
    cam.set_transform(T.view_transform((-2.6, 1.5, -3.9), (-0.6, 1, -0.8), (0, 1, 0)))
    return w, cam


def cover(api, width=1920, height=1080):
    """Config 3a — scene-parser/examples/cover.yml (1 plane, 1 glass sphere, 17 cubes, 2 lights;
    `extend`ed materials and named transforms resolved as the parser does, lib.rs:140-187,333-373)."""
    w = api.world()
    w.add_light((50, 100, -50), (1, 1, 1))
    w.add_light((-400, 50, -10), (0.2, 0.2, 0.2))
    white = Material(color=(1, 1, 1), diffuse=0.7, ambient=0.1, specular=0.0, reflective=0.1)
    blue = white.copy(color=(0.537, 0.831, 0.914))
    red = white.copy(color=(0.941, 0.322, 0.388))
    purple = white.copy(color=(0.373, 0.404, 0.550))
    standard = T.chain(T.translation(1, -1, 1), T.scaling(0.5, 0.5, 0.5))
    large = T.chain(standard, T.scaling(3.5, 3.5, 3.5))
    medium = T.chain(standard, T.scaling(3, 3, 3))
    small = T.chain(standard, T.scaling(2, 2, 2))
    # white backdrop
    w.add_object(api.plane().set_material(Material(color=(1, 1, 1), ambient=1, diffuse=0, specular=0))
                 .set_transform(T.chain(T.rotation_x(1.5707963267948966), T.translation(0, 0, 500))))
    w.add_object(api.sphere().set_material(Material(color=(0.373, 0.404, 0.550), diffuse=0.2, ambient=0.0, specular=1.0,
                                                    shininess=200, reflective=0.7, transparency=0.7, refractive_index=1.5))
                 .set_transform(T.chain(large)))
    cubes = [(white, medium, (4, 0, 0)), (blue, large, (8.5, 1.5, -0.5)), (red, large, (0, 0, 4)),
             (white, small, (4, 0, 4)), (purple, medium, (7.5, 0.5, 4)), (white, medium, (-0.25, 0.25, 8)),
             (blue, large, (4, 1, 7.5)), (red, medium, (10, 2, 7.5)), (white, small, (8, 2, 12)),
             (white, small, (20, 1, 9)), (blue, large, (-0.5, -5, 0.25)), (red, large, (4, -4, 0)),
             (white, large, (8.5, -4, 0)), (white, large, (0, -4, 4)), (purple, large, (-0.5, -4.5, 8)),
             (white, large, (0, -8, 4)), (white, large, (-0.5, -8.5, 8))]
    for mat, size, tr in cubes:
        w.add_object(api.cube().set_material(mat).set_transform(T.chain(size, T.translation(*tr))))
    cam = api.camera(width, height, 0.785)
    cam.set_transform(T.view_transform((-6, 6, -10), (6, 0, 6), (-0.45, 1, 0)))
    return w, cam


def primitives_csg(api, width=1920, height=1080):
    """Config 3b — the objects of raytracer/src/bin/first_scene.rs:23-108 (3 patterned planes,
    mirror / checker / glass spheres, rotated cube, closed cylinder 0..1, closed cone -1..0, 2 lights)
    plus the CSG difference of two glass spheres under rotation_y(-pi/4) of glass_sphere2.rs:54-80
    (its two balls moved/scaled into view; the CSG node keeps the pure rotation)."""
    w = api.world()
    w.add_light((-10, 10, -10), (1.0, 1.0, 1.0))
    w.add_light((-5.0, 10.0, -6.0), (0.33, 0.33, 0.33))
    w.add_object(api.plane().set_material(Material(pattern=checkers_pattern((0.0, 0.5, 0.5), (0.5, 0.0, 0.5)))))
    w.add_object(api.plane().set_transform(T.chain(T.rotation_x(PI / 2.0), T.rotation_y(-PI / 4.0), T.translation(0, 0, 5)))
                 .set_material(Material(pattern=ring_pattern((0.0, 0.0, 1.0), (0.0, 1.0, 1.0)).set_transform(T.scaling(0.333, 0.333, 0.333)))))
    w.add_object(api.plane().set_transform(T.chain(T.rotation_x(PI / 2.0), T.rotation_y(PI / 4.0), T.translation(0, 0, 5)))
                 .set_material(Material(pattern=stripe_pattern((1, 1, 1), (0, 0, 0)))))
    w.add_object(api.sphere().set_transform(T.translation(-0.5, 1.0, 0.5))
                 .set_material(Material(color=(0.1, 1.0, 0.5), diffuse=0.7, specular=0.3, reflective=0.9)))
    w.add_object(api.sphere().set_transform(T.matmul(T.translation(1.5, 0.5, -0.5), T.scaling(0.5, 0.5, 0.5)))
                 .set_material(Material(pattern=checkers_pattern((1.0, 0.0, 0.0), (0.0, 1.0, 0.0)))))
    w.add_object(api.sphere().set_material(Material(color=(0.1, 0.0, 0.0), ambient=0.1, diffuse=0.05, reflective=0.3,
                                                    specular=1.0, shininess=300.0, transparency=1.0, refractive_index=1.5))
                 .set_transform(T.matmul(T.translation(-1.5, 0.33, -0.75), T.scaling(0.33, 0.33, 0.33))))
    w.add_object(api.cube().set_transform(T.chain(T.rotation_y(PI / 4.0), T.scaling(0.25, 0.25, 0.25), T.translation(0.0, 0.25, -1.0))))
    w.add_object(api.cylinder(0.0, 1.0, True).set_transform(T.matmul(T.translation(1.0, 0.0, -1.2), T.scaling(0.33, 0.33, 0.33))))
    w.add_object(api.cone(-1.0, 0.0, True).set_transform(T.matmul(T.translation(-1.0, 0.33, -1.2), T.scaling(0.33, 0.33, 0.33))))
    glass = Material(color=(1, 1, 1), ambient=0.0, diffuse=0.0, specular=0.9, shininess=300.0, reflective=0.9,
                     transparency=0.9, refractive_index=1.5)
    ball1 = api.sphere().set_material(glass).set_transform(T.matmul(T.translation(0.09, 1.9, -2.3), T.scaling(0.35, 0.35, 0.35)))
    ball2 = api.sphere().set_material(glass).set_transform(T.matmul(T.translation(-0.09, 1.9, -2.3), T.scaling(0.35, 0.35, 0.35)))
    w.add_object(api.csg(DIFFERENCE, ball1, ball2).set_transform(T.rotation_y(-PI / 4.0)))
    cam = api.camera(width, height, PI / 3.0)
    cam.set_transform(T.view_transform((0.0, 1.5, -5.0), (0, 1, 0), (0, 1, 0)))
    return w, cam


def glass_sphere2(api, width=960, height=960, aa=1):
    """raytracer/src/bin/glass_sphere2.rs verbatim (CSG difference of two glass balls in front of a
    checkered wall); the reference renders it at 3840x3840 AA x16."""
    w = api.world()
    w.add_light((2.0, 10.0, -5.0), (0.9, 0.9, 0.9))
    w.add_object(api.plane().set_transform(T.matmul(T.translation(0, 0, 10), T.rotation_x(PI / 2.0)))
                 .set_material(Material(pattern=checkers_pattern((0.15, 0.15, 0.15), (0.8, 0.8, 0.8)), ambient=0.8,
                                        diffuse=0.2, specular=0.0)))
    glass = Material(color=(1, 1, 1), ambient=0.0, diffuse=0.0, specular=0.9, shininess=300.0, reflective=0.9,
                     transparency=0.9, refractive_index=1.5)
    ball1 = api.sphere().set_material(glass).set_transform(T.translation(0.25, 0.0, 0.0))
    ball2 = api.sphere().set_material(glass).set_transform(T.translation(-0.25, 0.0, 0.0))
    w.add_object(api.csg(DIFFERENCE, ball1, ball2).set_transform(T.rotation_y(-PI / 4.0)))
    cam = api.camera(width, height, 0.45)
    cam.set_transform(T.view_transform((0, 0, -7), (0, 0, 0), (0, 1, 0)))
    cam.aa_samples = aa
    return w, cam


def hexagon(api, width=640, height=360):
    """raytracer/src/bin/hexagon.rs — nested groups with baked transforms and a group-wide material."""
    w = api.world()
    w.add_light((-10, 10, -10), (1.0, 1.0, 1.0))
    w.add_light((-5.0, 10.0, -6.0), (0.33, 0.33, 0.33))
    hexg = api.group()
    for n in range(6):
        side = api.group()
        side.add_child(api.sphere().set_transform(T.matmul(T.translation(0, 0, -1), T.scaling(0.25, 0.25, 0.25))))
        side.add_child(api.cylinder(0, 1, False).set_transform(
            T.chain(T.scaling(0.25, 1.0, 0.25), T.rotation_z(-PI / 2.0), T.rotation_y(-PI / 6.0), T.translation(0, 0, -1))))
        side.set_transform(T.rotation_y(float(n) * PI / 3.0))
        hexg.add_child(side)
    hexg.set_material(Material(pattern=checkers_pattern((1.0, 0.0, 0.0), (0.0, 1.0, 0.0))))
    hexg.set_transform(T.scaling(1.5, 1.5, 1.5))
    w.add_object(hexg)
    cam = api.camera(width, height, PI / 3.0)
    cam.set_transform(T.view_transform((0.0, 2.5, -5.0), (0, 0, 0), (0, 1, 0)))
    return w, cam


# ---- procedural stand-in meshes (the teapot OBJ files are git-ignored in the reference) ----------

def lathe_obj(segments=64, rings=48):
    """Deterministic teapot-like surface of revolution about +z (z up, ~8 units tall), as OBJ text
    with `vn` records and `f v//n` faces → SmoothTriangle path (obj_parser/mod.rs:70-83)."""
    lines = []

    def radius(t):  # t in [0,1] bottom→top
        return 0.4 + 3.2 * math.sin(PI * (0.08 + 0.84 * t)) ** 1.5 + (0.9 * math.exp(-((t - 0.93) / 0.05) ** 2))

    height = 8.0
    for j in range(rings + 1):
        t = j / rings
        r, z = radius(t), height * t
        dr = (radius(min(1.0, t + 1e-4)) - radius(max(0.0, t - 1e-4))) / (height * 2e-4 if 0.0 < t < 1.0 else height * 1e-4)
        for i in range(segments):
            a = 2.0 * PI * i / segments
            x, y = r * math.cos(a), r * math.sin(a)
            nx, ny, nz = math.cos(a), math.sin(a), -dr
            nl = math.sqrt(nx * nx + ny * ny + nz * nz)
            lines.append("v %.9f %.9f %.9f" % (x, y, z))
            lines.append("vn %.9f %.9f %.9f" % (nx / nl, ny / nl, nz / nl))
    for j in range(rings):
        for i in range(segments):
            a = j * segments + i + 1
            b = j * segments + (i + 1) % segments + 1
            c = a + segments
            d = b + segments
            lines.append("f %d//%d %d//%d %d//%d" % (a, a, b, b, d, d))
            lines.append("f %d//%d %d//%d %d//%d" % (a, a, d, d, c, c))
    return "\n".join(lines) + "\n"


def torus_obj(nu=100, nv=100, major=1.6, minor=0.7):
    """Deterministic torus (y up, resting on y=0), plain `f` faces → flat Triangle path."""
    lines = []
    for j in range(nv):
        v = 2.0 * PI * j / nv
        for i in range(nu):
            u = 2.0 * PI * i / nu
            x = (major + minor * math.cos(v)) * math.cos(u)
            z = (major + minor * math.cos(v)) * math.sin(u)
            y = minor + minor * math.sin(v)
            lines.append("v %.9f %.9f %.9f" % (x, y, z))
    for j in range(nv):
        for i in range(nu):
            a = j * nu + i + 1
            b = j * nu + (i + 1) % nu + 1
            c = ((j + 1) % nv) * nu + i + 1
            d = ((j + 1) % nv) * nu + (i + 1) % nu + 1
            lines.append("f %d %d %d %d" % (a, b, d, c))  # quad → fan triangulation (mod.rs:103-116)
    return "\n".join(lines) + "\n"


def teapot_standin(api, width=1920, height=1080, aa=1, segments=64, rings=48, nu=100, nv=100):
    """Config 4 — the scene of raytracer/src/bin/teapot.rs:25-92 with generated stand-in meshes:
    checker floor (reflective 0.4), 2 lights, a smooth mesh divide(5) under
    scale(.12)·rotate_x(-pi/2)·rotate_y(pi/7)·translate(1.5,0,0), a flat mesh divide(1000) under
    scale(.6)·rotate_y(pi/5)·translate(-1.9,0,0), shared material reflective 0.1."""
    w = api.world()
    w.add_light((-4, 4, -5), (1.0, 1.0, 1.0))
    w.add_light((-3.0, 5.0, -3.0), (0.33, 0.33, 0.33))
    w.add_object(api.plane().set_material(Material(pattern=checkers_pattern((0.8, 0.8, 0.8), (0.6, 0.6, 0.6)),
                                                   specular=0.8, reflective=0.4)))
    material = Material(color=(0.6, 0.4, 0.2), ambient=0.35, diffuse=0.3, specular=0.2, reflective=0.1)
    smooth = api.obj_group(lathe_obj(segments, rings))
    smooth.set_transform(T.chain(T.scaling(0.12, 0.12, 0.12), T.rotation_x(-PI / 2.0), T.rotation_y(PI / 7.0), T.translation(1.5, 0.0, 0.0)))
    smooth.divide(5)
    smooth.set_material(material)
    flat = api.obj_group(torus_obj(nu, nv))
    flat.set_material(material)
    flat.set_transform(T.chain(T.scaling(0.6, 0.6, 0.6), T.rotation_y(PI / 5.0), T.translation(-1.9, 0.0, 0.0)))
    flat.divide(1000)
    w.add_object(smooth)
    w.add_object(flat)
    cam = api.camera(width, height, PI / 3.0)
    cam.set_transform(T.view_transform((0.0, 2.5, -7.0), (0.0, 1.25, 0.0), (0, 1, 0)))
    cam.aa_samples = aa
    return w, cam


class SplitMix64:
    def __init__(self, seed):
        self.s = seed & 0xFFFFFFFFFFFFFFFF

    def next(self):
        self.s = (self.s + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        return z ^ (z >> 31)

    def uniform(self, lo=0.0, hi=1.0):
        return lo + (hi - lo) * ((self.next() >> 11) * (1.0 / 9007199254740992.0))


def heightfield_triangles(nx=1000, nz=500):
    """nx x nz quads → 2*nx*nz smooth triangles of a rolling height field over x∈[-10,10], z∈[-5,5]
    (object space), with analytic vertex normals. Returns (points[n,3,3], normals[n,3,3])."""
    xs = np.linspace(-10.0, 10.0, nx + 1)
    zs = np.linspace(-5.0, 5.0, nz + 1)
    X, Z = np.meshgrid(xs, zs, indexing="ij")
    Y = 0.35 * np.sin(1.3 * X) * np.cos(1.7 * Z) + 0.15 * np.sin(3.1 * X + 0.5) * np.sin(2.3 * Z)
    dYdx = 0.35 * 1.3 * np.cos(1.3 * X) * np.cos(1.7 * Z) + 0.15 * 3.1 * np.cos(3.1 * X + 0.5) * np.sin(2.3 * Z)
    dYdz = -0.35 * 1.7 * np.sin(1.3 * X) * np.sin(1.7 * Z) + 0.15 * 2.3 * np.sin(3.1 * X + 0.5) * np.cos(2.3 * Z)
    N = np.stack([-dYdx, np.ones_like(Y), -dYdz], axis=-1)
    N /= np.linalg.norm(N, axis=-1, keepdims=True)
    P = np.stack([X, Y, Z], axis=-1)
    a, b = P[:-1, :-1], P[1:, :-1]
    c, d = P[:-1, 1:], P[1:, 1:]
    na, nb = N[:-1, :-1], N[1:, :-1]
    nc, nd = N[:-1, 1:], N[1:, 1:]
    t1 = np.stack([a, c, d], axis=2)   # (nx,nz,3,3), counter-clockwise seen from +y
    t2 = np.stack([a, d, b], axis=2)
    n1 = np.stack([na, nc, nd], axis=2)
    n2 = np.stack([na, nd, nb], axis=2)
    pts = np.stack([t1, t2], axis=2).reshape(-1, 3, 3)
    nrm = np.stack([n1, n2], axis=2).reshape(-1, 3, 3)
    return np.ascontiguousarray(pts), np.ascontiguousarray(nrm)


def mesh_and_spheres(api, width=3840, height=2160, nx=1000, nz=500, n_spheres=200, divide=8, aa=1):
    """Config 5 — synthetic, seed-fixed (splitmix64 0x5EED0005): an nx x nz height field of
    2*nx*nz smooth triangles under one group transform (scale + translate: no rotation, so
    Group::divide's boxes do not inflate) with divide(8); n_spheres spheres above it cycling
    diffuse / mirror / glass; 2 lights; checker floor plane."""
    w = api.world()
    w.add_light((-12, 14, -12), (1.0, 1.0, 1.0))
    w.add_light((9.0, 11.0, -7.0), (0.35, 0.35, 0.35))
    w.add_object(api.plane().set_transform(T.translation(0, -1.0, 0))
                 .set_material(Material(pattern=checkers_pattern((0.75, 0.75, 0.75), (0.45, 0.45, 0.45)), specular=0.2, reflective=0.2)))
    pts, nrm = heightfield_triangles(nx, nz)
    mesh = api.group()
    mesh.add_triangles(pts, nrm)
    mesh.set_transform(T.chain(T.scaling(0.9, 1.2, 0.9), T.translation(0.0, 0.2, 1.0)))
    mesh.divide(divide)
    mesh.set_material(Material(color=(0.35, 0.55, 0.3), ambient=0.15, diffuse=0.7, specular=0.25, shininess=80.0, reflective=0.05))
    w.add_object(mesh)
    rng = SplitMix64(0x5EED0005)
    kinds = [Material(color=(0.9, 0.35, 0.25), diffuse=0.8, specular=0.4, shininess=60.0),
             Material(color=(0.9, 0.9, 0.95), ambient=0.05, diffuse=0.1, specular=0.9, shininess=300.0, reflective=0.9),
             Material(color=(0.05, 0.1, 0.15), ambient=0.0, diffuse=0.2, specular=0.9, shininess=300.0, reflective=0.9,
                      transparency=0.9, refractive_index=1.5)]
    for i in range(n_spheres):
        cx, cy, cz = rng.uniform(-8.0, 8.0), rng.uniform(1.2, 4.5), rng.uniform(-2.5, 5.0)
        s = rng.uniform(0.05, 0.4)
        col = (rng.uniform(0.2, 1.0), rng.uniform(0.2, 1.0), rng.uniform(0.2, 1.0))
        m = kinds[i % 3]
        if i % 3 == 0:
            m = m.copy(color=col)
        w.add_object(api.sphere().set_transform(T.matmul(T.translation(cx, cy, cz), T.scaling(s, s, s))).set_material(m))
    cam = api.camera(width, height, PI / 3.0)
    cam.set_transform(T.view_transform((0.0, 6.5, -13.0), (0.0, 1.0, 1.0), (0, 1, 0)))
    cam.aa_samples = aa
    return w, cam


def solids_field(api, width=640, height=360):
    """Parity case for bounded solids in the BVH: a 6 x 4 field of cubes, cylinders and cones (open and
    closed, rotated and scaled, some glass or mirrored, every third one inside a group) plus one
    unbounded cylinder and one unbounded cone, which must stay every-ray prims."""
    w = api.world()
    w.add_light((-8, 10, -8), (1, 1, 1))
    w.add_light((6, 4, -9), (0.3, 0.3, 0.4))
    w.add_object(api.plane().set_material(Material(pattern=checkers_pattern((0.3, 0.3, 0.3), (0.7, 0.7, 0.7)), reflective=0.2)))
    w.add_object(api.cylinder().set_transform(T.chain(T.scaling(0.2, 1, 0.2), T.translation(-7, 0, 6)))
                 .set_material(Material(color=(0.9, 0.8, 0.1))))
    w.add_object(api.cone().set_transform(T.chain(T.scaling(0.15, 0.6, 0.15), T.translation(7, 2, 6)))
                 .set_material(Material(color=(0.1, 0.8, 0.9))))
    glass = Material(color=(0.05, 0.05, 0.1), diffuse=0.2, specular=0.9, shininess=250, reflective=0.8,
                     transparency=0.85, refractive_index=1.5)
    k = 0
    for ix in range(6):
        for iz in range(4):
            x, z = -5.0 + 2.0 * ix, -1.0 + 2.2 * iz
            kind = k % 5
            if kind == 0:
                shape = api.cube()
            elif kind == 1:
                shape = api.cylinder(-0.5, 1.0, True)
            elif kind == 2:
                shape = api.cone(-1.0, 0.0, True)
            elif kind == 3:
                shape = api.cylinder(0.0, 1.5, False)
            else:
                shape = api.cone(-1.0, 0.5, False)
            s = 0.35 + 0.07 * (k % 4)
            shape.set_transform(T.chain(T.scaling(s, s * (1.0 + 0.2 * (k % 3)), s), T.rotation_z(0.11 * (k % 5)),
                                        T.rotation_y(0.37 * k), T.translation(x, 1.0 + 0.15 * (k % 3), z)))
            if k % 7 == 3:
                shape.set_material(glass)
            elif k % 4 == 1:
                shape.set_material(Material(color=(0.2 + 0.1 * (k % 5), 0.3, 0.8 - 0.1 * (k % 6)), reflective=0.4))
            else:
                shape.set_material(Material(color=(0.9 - 0.1 * (k % 6), 0.4 + 0.1 * (k % 4), 0.2), shininess=30 + 10 * k))
            if k % 3 == 2:
                g = api.group()
                g.add_child(shape)
                g.set_transform(T.chain(T.rotation_y(0.05 * k), T.translation(0, 0.1, 0)))
                w.add_object(g)
            else:
                w.add_object(shape)
            k += 1
    cam = api.camera(width, height, 1.0)
    cam.set_transform(T.view_transform((0, 6, -11), (0, 0.5, 3), (0, 1, 0)))
    return w, cam


def pattern_gallery(api, width=640, height=360):
    """Parity case for the pattern kinds no reference scene file uses: gradient (pattern/gradient.rs:14-18) on a
    plane, a sphere (with a pattern transform) and a cube, the test pattern (test_pattern.rs:7-9) on a cylinder, a
    ring-patterned cone and a stripe-patterned mirror sphere; two lights."""
    w = api.world()
    w.add_light((-6, 8, -8), (0.9, 0.9, 0.9))
    w.add_light((7, 5, -6), (0.3, 0.25, 0.35))
    w.add_object(api.plane().set_material(Material(
        pattern=gradient_pattern((0.9, 0.2, 0.1), (0.1, 0.3, 0.9)).set_transform(T.chain(T.scaling(3.0, 1.0, 1.0), T.rotation_y(0.4))),
        reflective=0.15, specular=0.2)))
    w.add_object(api.sphere().set_transform(T.translation(-2.2, 1.0, 0.5))
                 .set_material(Material(pattern=gradient_pattern((1.0, 1.0, 0.0), (0.0, 0.6, 0.2))
                                        .set_transform(T.chain(T.scaling(0.5, 0.5, 0.5), T.rotation_z(0.7), T.translation(0.25, 0, 0))),
                                        diffuse=0.8, specular=0.4, shininess=50.0)))
    w.add_object(api.cube().set_transform(T.chain(T.scaling(0.7, 0.7, 0.7), T.rotation_y(0.6), T.translation(0.0, 0.7, 1.5)))
                 .set_material(Material(pattern=gradient_pattern((0.2, 0.2, 0.2), (1.0, 1.0, 1.0)), transparency=0.3,
                                        refractive_index=1.2)))
    w.add_object(api.cylinder(0.0, 1.5, True).set_transform(T.chain(T.scaling(0.6, 1.0, 0.6), T.translation(2.2, 0.0, 0.2)))
                 .set_material(Material(pattern=test_pattern().set_transform(T.scaling(2.0, 1.5, 2.0)), ambient=0.3, diffuse=0.6)))
    w.add_object(api.cone(-1.0, 0.0, True).set_transform(T.chain(T.scaling(0.8, 1.4, 0.8), T.translation(0.8, 1.4, -1.4)))
                 .set_material(Material(pattern=ring_pattern((0.9, 0.9, 0.9), (0.1, 0.1, 0.5)).set_transform(T.scaling(0.2, 0.2, 0.2)))))
    w.add_object(api.sphere().set_transform(T.chain(T.scaling(0.6, 0.6, 0.6), T.translation(-0.8, 0.6, -1.6)))
                 .set_material(Material(pattern=stripe_pattern((0.8, 0.1, 0.1), (0.95, 0.95, 0.95)).set_transform(
                     T.chain(T.scaling(0.2, 0.2, 0.2), T.rotation_z(0.5))), reflective=0.5)))
    cam = api.camera(width, height, 1.0)
    cam.set_transform(T.view_transform((0.5, 3.0, -7.0), (0, 0.8, 0), (0, 1, 0)))
    return w, cam


def cone_field(api):
    """Cones for the `abs(a) < EPSILON` branch of Cone::local_intersect (cone.rs:102-110, with its `-c / 2.0 * b`):
    untransformed and translated only, so that a world direction proportional to (0, 1, 1) has a == 0 in object space."""
    w = api.world()
    w.add_light((-10, 10, -10), (1, 1, 1))
    w.add_object(api.cone().set_material(Material(color=(0.9, 0.4, 0.2))))
    w.add_object(api.cone(-1.0, 2.0, True).set_transform(T.translation(4.0, 0.0, 0.0)).set_material(Material(color=(0.2, 0.4, 0.9), reflective=0.3)))
    w.add_object(api.cone(-0.5, 0.5, False).set_transform(T.translation(-4.0, 1.0, 1.0)).set_material(Material(color=(0.2, 0.9, 0.3))))
    w.add_object(api.plane().set_transform(T.translation(0, -3.0, 0)).set_material(Material(pattern=checkers_pattern((0.2, 0.2, 0.2), (0.8, 0.8, 0.8)))))
    return w


CONFIGS = {
    "default_world": default_world,
    "chapter7": chapter7,                # config 1
    "reflect_refract": reflect_refract,  # config 2 (bench workload)
    "cover": cover,                      # config 3a
    "primitives_csg": primitives_csg,    # config 3b
    "glass_sphere2": glass_sphere2,
    "hexagon": hexagon,
    "solids_field": solids_field,
    "pattern_gallery": pattern_gallery,
    "teapot_standin": teapot_standin,    # config 4
    "mesh_and_spheres": mesh_and_spheres,  # config 5
}
