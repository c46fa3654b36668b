// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement (C++17) of the f64 math core of tlinford/raytracer-challenge-rs.
// Every function keeps the reference's floating-point operation ORDER so that
// results are bit-comparable; build with `g++ -O2 -ffp-contract=off` (Rust never
// contracts a*b+c into an FMA, so neither may this file).
//
// Reference files restated here (paths relative to /root/reference/raytracer/src):
//   lib.rs:18-31      EPSILON, equal, equal_ignore_inf
//   point.rs:5-60     Point and its operators
//   vector.rs:5-109   Vector, magnitude/normalize/reflect, dot, cross
//   color.rs:6-100    Color, average, Sum
//   matrix.rs:12-260  Matrix (row-major, cofactor-expansion determinant/inverse)
//   transform.rs:7-90 translation/scaling/rotation_*/shearing/view_transform
//   ray.rs:3-29       Ray, position, transform
#pragma once
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <vector>

namespace orc {

// lib.rs:18
constexpr double EPSILON = 0.00001;
constexpr double INF = std::numeric_limits<double>::infinity();

// lib.rs:20-22
inline bool equal(double a, double b) { return std::fabs(a - b) < EPSILON; }

// lib.rs:24-31 — note the reference's (sloppy) sign logic is kept verbatim:
// (a>0 && a>0) || (b<0 && b<0).
inline bool equal_ignore_inf(double a, double b) {
  if (std::isinf(a) && std::isinf(b)) {
    return (!std::signbit(a) && !std::signbit(a)) || (std::signbit(b) && std::signbit(b));
  }
  return equal(a, b);
}

// Rust f64::max / f64::min: NaN-ignoring, identical to C fmax/fmin.
inline double rmax(double a, double b) { return std::fmax(a, b); }
inline double rmin(double a, double b) { return std::fmin(a, b); }

struct Vector;

// point.rs:5-60
struct Point {
  double x = 0, y = 0, z = 0;
  Point() = default;
  Point(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
  static Point origin() { return Point(0, 0, 0); }
  bool eq_ignore_inf(const Point& o) const {
    return equal_ignore_inf(x, o.x) && equal_ignore_inf(y, o.y) && equal_ignore_inf(z, o.z);
  }
  bool operator==(const Point& o) const { return equal(x, o.x) && equal(y, o.y) && equal(z, o.z); }
};

// vector.rs:5-109
struct Vector {
  double x = 0, y = 0, z = 0;
  Vector() = default;
  Vector(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
  double magnitude() const { return std::sqrt(x * x + y * y + z * z); }  // vector.rs:21
  Vector normalize() const {                                            // vector.rs:25
    double m = magnitude();
    return Vector(x / m, y / m, z / m);
  }
  bool operator==(const Vector& o) const { return equal(x, o.x) && equal(y, o.y) && equal(z, o.z); }
};

inline Vector operator+(Vector a, Vector b) { return Vector(a.x + b.x, a.y + b.y, a.z + b.z); }
inline Vector operator-(Vector a, Vector b) { return Vector(a.x - b.x, a.y - b.y, a.z - b.z); }
inline Vector operator-(Vector a) { return Vector(-a.x, -a.y, -a.z); }
inline Vector operator*(Vector a, double s) { return Vector(a.x * s, a.y * s, a.z * s); }
inline Vector operator/(Vector a, double s) { return Vector(a.x / s, a.y / s, a.z / s); }
inline Point operator+(Point p, Vector v) { return Point(p.x + v.x, p.y + v.y, p.z + v.z); }
inline Vector operator-(Point a, Point b) { return Vector(a.x - b.x, a.y - b.y, a.z - b.z); }
inline Point operator-(Point p, Vector v) { return Point(p.x - v.x, p.y - v.y, p.z - v.z); }

// vector.rs:99-109
inline double dot(Vector a, Vector b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vector cross(Vector a, Vector b) {
  return Vector(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// vector.rs:30-32: *self - normal * 2.0 * dot(*self, normal)  ==  v - ((n*2.0)*d)
inline Vector reflect(Vector v, Vector n) { return v - n * 2.0 * dot(v, n); }

// color.rs:6-100
struct Color {
  double red = 0, green = 0, blue = 0;
  Color() = default;
  Color(double r, double g, double b) : red(r), green(g), blue(b) {}
  static Color black() { return Color(0, 0, 0); }
  static Color white() { return Color(1, 1, 1); }
  bool operator==(const Color& o) const {
    return equal(red, o.red) && equal(green, o.green) && equal(blue, o.blue);
  }
};
inline Color operator+(Color a, Color b) { return Color(a.red + b.red, a.green + b.green, a.blue + b.blue); }
inline Color operator-(Color a, Color b) { return Color(a.red - b.red, a.green - b.green, a.blue - b.blue); }
inline Color operator*(Color a, double s) { return Color(a.red * s, a.green * s, a.blue * s); }
inline Color operator*(Color a, Color b) { return Color(a.red * b.red, a.green * b.green, a.blue * b.blue); }
// color.rs:26-33: sequential sum from black, then * (1.0 / n)
inline Color color_average(const std::vector<Color>& cs) {
  Color avg = Color::black();
  for (const Color& c : cs) avg = avg + c;
  return avg * (1.0 / (double)cs.size());
}

// matrix.rs:12-260 — dense row-major, any small size; the 4x4 inverse goes through the
// same recursive cofactor expansion (submatrix → minor → cofactor → determinant).
struct Matrix {
  size_t rows = 0, columns = 0;
  std::vector<double> e;

  Matrix() = default;
  Matrix(size_t r, size_t c) : rows(r), columns(c), e(r * c, 0.0) {}
  static Matrix zero(size_t r, size_t c) { return Matrix(r, c); }
  static Matrix identity(size_t r, size_t c) {  // matrix.rs:27-35
    Matrix m(r, c);
    for (size_t i = 0; i < c; i++) m(i, i) = 1.0;
    return m;
  }
  static Matrix from_slice(size_t r, size_t c, const double* s, size_t n) {  // matrix.rs:37
    if (r * c != n) throw std::invalid_argument("Matrix::from_slice: rows*columns != len");
    Matrix m(r, c);
    for (size_t i = 0; i < n; i++) m.e[i] = s[i];
    return m;
  }
  static Matrix from_list(size_t r, size_t c, std::initializer_list<double> l) {
    if (r * c != l.size()) throw std::invalid_argument("Matrix::from_list: rows*columns != len");
    Matrix m(r, c);
    size_t i = 0;
    for (double v : l) m.e[i++] = v;
    return m;
  }
  double& operator()(size_t i, size_t j) { return e[i * columns + j]; }
  double operator()(size_t i, size_t j) const { return e[i * columns + j]; }

  Matrix transpose() const {  // matrix.rs:73-83
    Matrix t(columns, rows);
    for (size_t i = 0; i < rows; i++)
      for (size_t j = 0; j < columns; j++) t(j, i) = (*this)(i, j);
    return t;
  }
  double determinant() const {  // matrix.rs:85-96
    if (rows == 2 && columns == 2) return (*this)(0, 0) * (*this)(1, 1) - (*this)(0, 1) * (*this)(1, 0);
    double det = 0.0;
    for (size_t c = 0; c < columns; c++) det += (*this)(0, c) * cofactor(0, c);
    return det;
  }
  Matrix submatrix(size_t row, size_t column) const {  // matrix.rs:98-120
    assert(row < rows && rows > 1 && column < columns && columns > 1);
    Matrix sub(rows - 1, columns - 1);
    for (size_t i = 0; i < sub.rows; i++)
      for (size_t j = 0; j < sub.columns; j++) {
        size_t ii = i < row ? i : i + 1;
        size_t jj = j < column ? j : j + 1;
        sub(i, j) = (*this)(ii, jj);
      }
    return sub;
  }
  double minor(size_t r, size_t c) const { return submatrix(r, c).determinant(); }  // matrix.rs:122
  double cofactor(size_t r, size_t c) const {                                       // matrix.rs:126
    return ((r + c) % 2 == 1) ? -minor(r, c) : minor(r, c);
  }
  bool is_invertible() const { return !equal(determinant(), 0.0); }  // matrix.rs:134
  Matrix inverse() const {                                           // matrix.rs:138-153
    if (!is_invertible()) throw std::domain_error("Matrix::inverse: not invertible (|det| < EPSILON)");
    assert(rows == columns);
    Matrix inv(rows, columns);
    double det = determinant();
    for (size_t i = 0; i < rows; i++)
      for (size_t j = 0; j < columns; j++) {
        double c = cofactor(i, j);
        inv(j, i) = c / det;
      }
    return inv;
  }
  bool operator==(const Matrix& o) const {  // matrix.rs:201-208 (zip: shortest length)
    size_t n = e.size() < o.e.size() ? e.size() : o.e.size();
    for (size_t i = 0; i < n; i++)
      if (!equal(e[i], o.e[i])) return false;
    return true;
  }
  bool operator!=(const Matrix& o) const { return !(*this == o); }
  bool bits_equal(const Matrix& o) const {
    return rows == o.rows && columns == o.columns &&
           std::memcmp(e.data(), o.e.data(), e.size() * sizeof(double)) == 0;
  }
  // fluent builders, matrix.rs:155-183: each is `&t * self`
  Matrix translate(double x, double y, double z) const;
  Matrix scale(double x, double y, double z) const;
  Matrix rotate_x(double r) const;
  Matrix rotate_y(double r) const;
  Matrix rotate_z(double r) const;
  Matrix shear(double xy, double xz, double yx, double yz, double zx, double zy) const;
};

// matrix.rs:210-230: c = 0.0; c += a[row][i] * b[i][col]
inline Matrix operator*(const Matrix& a, const Matrix& b) {
  assert(a.columns == b.rows);
  Matrix m(a.rows, b.columns);
  for (size_t row = 0; row < a.rows; row++)
    for (size_t col = 0; col < b.columns; col++) {
      double c = 0.0;
      for (size_t i = 0; i < a.columns; i++) c += a(row, i) * b(i, col);
      m(row, col) = c;
    }
  return m;
}
// matrix.rs:232-245: affine, rows 0..2 only, `+ m[i][3]` last
inline Point operator*(const Matrix& m, Point p) {
  assert(m.rows == 4 && m.columns == 4);
  return Point(m(0, 0) * p.x + m(0, 1) * p.y + m(0, 2) * p.z + m(0, 3),
               m(1, 0) * p.x + m(1, 1) * p.y + m(1, 2) * p.z + m(1, 3),
               m(2, 0) * p.x + m(2, 1) * p.y + m(2, 2) * p.z + m(2, 3));
}
// matrix.rs:247-260
inline Vector operator*(const Matrix& m, Vector v) {
  assert(m.columns == 4);
  return Vector(m(0, 0) * v.x + m(0, 1) * v.y + m(0, 2) * v.z,
                m(1, 0) * v.x + m(1, 1) * v.y + m(1, 2) * v.z,
                m(2, 0) * v.x + m(2, 1) * v.y + m(2, 2) * v.z);
}

// transform.rs:7-71
inline Matrix translation(double x, double y, double z) {
  Matrix t = Matrix::identity(4, 4);
  t(0, 3) = x; t(1, 3) = y; t(2, 3) = z;
  return t;
}
inline Matrix scaling(double x, double y, double z) {
  Matrix s = Matrix::identity(4, 4);
  s(0, 0) = x; s(1, 1) = y; s(2, 2) = z;
  return s;
}
inline Matrix rotation_x(double r) {
  Matrix m = Matrix::identity(4, 4);
  m(1, 1) = std::cos(r); m(1, 2) = -std::sin(r); m(2, 1) = std::sin(r); m(2, 2) = std::cos(r);
  return m;
}
inline Matrix rotation_y(double r) {
  Matrix m = Matrix::identity(4, 4);
  m(0, 0) = std::cos(r); m(0, 2) = std::sin(r); m(2, 0) = -std::sin(r); m(2, 2) = std::cos(r);
  return m;
}
inline Matrix rotation_z(double r) {
  Matrix m = Matrix::identity(4, 4);
  m(0, 0) = std::cos(r); m(0, 1) = -std::sin(r); m(1, 0) = std::sin(r); m(1, 1) = std::cos(r);
  return m;
}
inline Matrix shearing(double xy, double xz, double yx, double yz, double zx, double zy) {
  Matrix s = Matrix::identity(4, 4);
  s(0, 1) = xy; s(0, 2) = xz; s(1, 0) = yx; s(1, 2) = yz; s(2, 0) = zx; s(2, 1) = zy;
  return s;
}
// transform.rs:73-90
inline Matrix view_transform(Point from, Point to, Vector up) {
  Vector forward = (to - from).normalize();
  Vector upn = up.normalize();
  Vector left = cross(forward, upn);
  Vector true_up = cross(left, forward);
  Matrix orientation = Matrix::from_list(4, 4,
      {left.x, left.y, left.z, 0.0,
       true_up.x, true_up.y, true_up.z, 0.0,
       -forward.x, -forward.y, -forward.z, 0.0,
       0.0, 0.0, 0.0, 1.0});
  return orientation * translation(-from.x, -from.y, -from.z);
}

inline Matrix Matrix::translate(double x, double y, double z) const { return translation(x, y, z) * *this; }
inline Matrix Matrix::scale(double x, double y, double z) const { return scaling(x, y, z) * *this; }
inline Matrix Matrix::rotate_x(double r) const { return rotation_x(r) * *this; }
inline Matrix Matrix::rotate_y(double r) const { return rotation_y(r) * *this; }
inline Matrix Matrix::rotate_z(double r) const { return rotation_z(r) * *this; }
inline Matrix Matrix::shear(double xy, double xz, double yx, double yz, double zx, double zy) const {
  return shearing(xy, xz, yx, yz, zx, zy) * *this;
}

// ray.rs:3-29
struct Ray {
  Point origin;
  Vector direction;
  Ray() = default;
  Ray(Point o, Vector d) : origin(o), direction(d) {}
  Point position(double t) const { return origin + direction * t; }          // ray.rs:22
  Ray transform(const Matrix& m) const { return Ray(m * origin, m * direction); }  // ray.rs:26
  bool operator==(const Ray& o) const { return origin == o.origin && direction == o.direction; }
};

}  // namespace orc
