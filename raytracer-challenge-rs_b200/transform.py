"""4x4 transform builders for scene scripts, in plain Python floats (IEEE f64, never fused).

Same operation order as the reference (raytracer/src/matrix.rs:155-230, transform.rs:7-90), so a
scene script hands bit-identical matrices to whichever backend builds the scene. Matrices are
row-major lists of 16 floats. Only products and builders live here — inverses are computed by the
backend (host library) with the reference's cofactor expansion.
"""
import math

IDENTITY = [1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0]


def matmul(a, b):
    """matrix.rs:210-230: c = 0.0; c += a[row][i] * b[i][col]"""
    out = [0.0] * 16
    for row in range(4):
        for col in range(4):
            c = 0.0
            for i in range(4):
                c += a[4 * row + i] * b[4 * i + col]
            out[4 * row + col] = c
    return out


def identity():
    return list(IDENTITY)


def translation(x, y, z):
    m = identity()
    m[3], m[7], m[11] = float(x), float(y), float(z)
    return m


def scaling(x, y, z):
    m = identity()
    m[0], m[5], m[10] = float(x), float(y), float(z)
    return m


def rotation_x(r):
    m = identity()
    m[5], m[6], m[9], m[10] = math.cos(r), -math.sin(r), math.sin(r), math.cos(r)
    return m


def rotation_y(r):
    m = identity()
    m[0], m[2], m[8], m[10] = math.cos(r), math.sin(r), -math.sin(r), math.cos(r)
    return m


def rotation_z(r):
    m = identity()
    m[0], m[1], m[4], m[5] = math.cos(r), -math.sin(r), math.sin(r), math.cos(r)
    return m


def shearing(xy, xz, yx, yz, zx, zy):
    m = identity()
    m[1], m[2], m[4], m[6], m[8], m[9] = float(xy), float(xz), float(yx), float(yz), float(zx), float(zy)
    return m


def chain(*ms):
    """Fluent `identity.a().b().c()` of matrix.rs:155-183: each step is `step * acc`."""
    acc = identity()
    for m in ms:
        acc = matmul(m, acc)
    return acc


def view_transform(frm, to, up):
    """transform.rs:73-90"""
    def sub(a, b):
        return [a[0] - b[0], a[1] - b[1], a[2] - b[2]]

    def norm(v):
        m = math.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])
        return [v[0] / m, v[1] / m, v[2] / m]

    def cross(a, b):
        return [a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]]

    frm = [float(v) for v in frm]
    to = [float(v) for v in to]
    up = [float(v) for v in up]
    forward = norm(sub(to, frm))
    upn = norm(up)
    left = cross(forward, upn)
    true_up = cross(left, forward)
    orientation = [left[0], left[1], left[2], 0.0,
                   true_up[0], true_up[1], true_up[2], 0.0,
                   -forward[0], -forward[1], -forward[2], 0.0,
                   0.0, 0.0, 0.0, 1.0]
    return matmul(orientation, translation(-frm[0], -frm[1], -frm[2]))
