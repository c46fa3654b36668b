// rth_math.hpp — host-side f64 math of the B200 backend's scene builder (namespace rth).
//
// The flattener needs the SAME bits the reference would produce for every transform, inverse and
// bounding box, so these helpers restate the reference's arithmetic (fixed 4x4 storage, no heap):
//   raytracer/src/matrix.rs:85-153 (determinant by cofactor expansion along row 0, inverse),
//   matrix.rs:210-260 (products), transform.rs:7-90 (builders, view_transform),
//   bounding_box.rs:39-93,138-169 (grow, contain, transform by 8 corners, split), lib.rs:18-31.
// Build with -ffp-contract=off.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <utility>

namespace rth {

constexpr double EPSILON = 0.00001;
constexpr double INF = std::numeric_limits<double>::infinity();
inline bool equal(double a, double b) { return std::fabs(a - b) < EPSILON; }

struct Vec3 {
  double x = 0, y = 0, z = 0;
};
inline Vec3 vsub(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 vcross(Vec3 a, Vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline Vec3 vnormalize(Vec3 a) {
  double m = std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z);
  return {a.x / m, a.y / m, a.z / m};
}

// determinant of a row-major n x n (n = 2..4) exactly as Matrix::determinant recurses
inline double det_n(const double* a, int n);
inline double minor_n(const double* a, int n, int row, int col) {
  double sub[9];
  int m = n - 1;
  for (int i = 0; i < m; i++)
    for (int j = 0; j < m; j++) sub[i * m + j] = a[(i < row ? i : i + 1) * n + (j < col ? j : j + 1)];
  return det_n(sub, m);
}
inline double cofactor_n(const double* a, int n, int row, int col) {
  double mn = minor_n(a, n, row, col);
  return ((row + col) % 2 == 1) ? -mn : mn;
}
inline double det_n(const double* a, int n) {
  if (n == 2) return a[0] * a[3] - a[1] * a[2];
  double det = 0.0;
  for (int c = 0; c < n; c++) det += a[c] * cofactor_n(a, n, 0, c);
  return det;
}

struct Mat4 {
  double m[16];
  static Mat4 identity() {
    Mat4 r;
    for (int i = 0; i < 16; i++) r.m[i] = (i % 5 == 0) ? 1.0 : 0.0;
    return r;
  }
  static Mat4 from(const double* p) {
    Mat4 r;
    std::memcpy(r.m, p, sizeof r.m);
    return r;
  }
  double& at(int i, int j) { return m[4 * i + j]; }
  double at(int i, int j) const { return m[4 * i + j]; }
  bool bits_equal(const Mat4& o) const { return std::memcmp(m, o.m, sizeof m) == 0; }
  double determinant() const { return det_n(m, 4); }
  // matrix.rs:138-153; throws where the reference panics (|det| < EPSILON)
  Mat4 inverse() const {
    double det = determinant();
    if (equal(det, 0.0)) throw std::domain_error("matrix is not invertible (|det| < 1e-5, matrix.rs:134-139)");
    Mat4 inv;
    for (int i = 0; i < 4; i++)
      for (int j = 0; j < 4; j++) inv.at(j, i) = cofactor_n(m, 4, i, j) / det;
    return inv;
  }
  Vec3 mul_point(Vec3 p) const {  // matrix.rs:232-245
    return {m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
            m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]};
  }
};
inline Mat4 operator*(const Mat4& a, const Mat4& b) {  // matrix.rs:210-230
  Mat4 r;
  for (int row = 0; row < 4; row++)
    for (int col = 0; col < 4; col++) {
      double c = 0.0;
      for (int i = 0; i < 4; i++) c += a.at(row, i) * b.at(i, col);
      r.at(row, col) = c;
    }
  return r;
}
inline Mat4 translation(double x, double y, double z) {
  Mat4 t = Mat4::identity();
  t.at(0, 3) = x; t.at(1, 3) = y; t.at(2, 3) = z;
  return t;
}
inline Mat4 scaling(double x, double y, double z) {
  Mat4 s = Mat4::identity();
  s.at(0, 0) = x; s.at(1, 1) = y; s.at(2, 2) = z;
  return s;
}
inline Mat4 rotation_x(double r) {
  Mat4 m = Mat4::identity();
  m.at(1, 1) = std::cos(r); m.at(1, 2) = -std::sin(r); m.at(2, 1) = std::sin(r); m.at(2, 2) = std::cos(r);
  return m;
}
inline Mat4 rotation_y(double r) {
  Mat4 m = Mat4::identity();
  m.at(0, 0) = std::cos(r); m.at(0, 2) = std::sin(r); m.at(2, 0) = -std::sin(r); m.at(2, 2) = std::cos(r);
  return m;
}
inline Mat4 rotation_z(double r) {
  Mat4 m = Mat4::identity();
  m.at(0, 0) = std::cos(r); m.at(0, 1) = -std::sin(r); m.at(1, 0) = std::sin(r); m.at(1, 1) = std::cos(r);
  return m;
}
inline Mat4 shearing(double xy, double xz, double yx, double yz, double zx, double zy) {
  Mat4 s = Mat4::identity();
  s.at(0, 1) = xy; s.at(0, 2) = xz; s.at(1, 0) = yx; s.at(1, 2) = yz; s.at(2, 0) = zx; s.at(2, 1) = zy;
  return s;
}
inline Mat4 view_transform(Vec3 from, Vec3 to, Vec3 up) {  // transform.rs:73-90
  Vec3 forward = vnormalize(vsub(to, from));
  Vec3 upn = vnormalize(up);
  Vec3 left = vcross(forward, upn);
  Vec3 true_up = vcross(left, forward);
  Mat4 o = Mat4::identity();
  o.at(0, 0) = left.x; o.at(0, 1) = left.y; o.at(0, 2) = left.z;
  o.at(1, 0) = true_up.x; o.at(1, 1) = true_up.y; o.at(1, 2) = true_up.z;
  o.at(2, 0) = -forward.x; o.at(2, 1) = -forward.y; o.at(2, 2) = -forward.z;
  return o * translation(-from.x, -from.y, -from.z);
}

struct BBox {
  Vec3 min{INF, INF, INF}, max{-INF, -INF, -INF};
  void add_point(Vec3 p) {
    if (p.x > max.x) max.x = p.x;
    if (p.y > max.y) max.y = p.y;
    if (p.z > max.z) max.z = p.z;
    if (p.x < min.x) min.x = p.x;
    if (p.y < min.y) min.y = p.y;
    if (p.z < min.z) min.z = p.z;
  }
  void add_box(const BBox& b) { add_point(b.min); add_point(b.max); }
  bool contains_point(Vec3 p) const {
    return (min.x <= p.x && p.x <= max.x) && (min.y <= p.y && p.y <= max.y) && (min.z <= p.z && p.z <= max.z);
  }
  bool contains_box(const BBox& b) const { return contains_point(b.min) && contains_point(b.max); }
  BBox transform(const Mat4& m) const {  // bounding_box.rs:76-93 — corner order as in the reference
    const Vec3 ps[8] = {min, {min.x, min.y, max.z}, {min.x, max.y, min.z}, {min.x, max.y, max.z},
                        {max.x, min.y, min.z}, {max.x, min.y, max.z}, {max.x, max.y, min.z}, max};
    BBox nb;
    for (const Vec3& p : ps) nb.add_point(m.mul_point(p));
    return nb;
  }
  std::pair<BBox, BBox> split() const {  // bounding_box.rs:138-169
    double dx = std::fabs(max.x - min.x), dy = std::fabs(max.y - min.y), dz = std::fabs(max.z - min.z);
    double greatest = std::fmax(std::fmax(dx, dy), dz);
    double x0 = min.x, y0 = min.y, z0 = min.z, x1 = max.x, y1 = max.y, z1 = max.z;
    if (equal(greatest, dx)) { x0 = x1 = min.x + dx / 2.0; }
    else if (equal(greatest, dy)) { y0 = y1 = min.y + dy / 2.0; }
    else { z0 = z1 = min.z + dz / 2.0; }
    BBox l, r;
    l.min = min; l.max = {x1, y1, z1};
    r.min = {x0, y0, z0}; r.max = max;
    return {l, r};
  }
};

}  // namespace rth
