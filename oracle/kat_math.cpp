// ORACLE KATs — matrix.rs:262-669, transform.rs:93-327, vector.rs:110-250, point.rs:62-97, color.rs:105-150,
// ray.rs:33-72, light.rs:26-38, canvas.rs:55-100 test modules restated against oracle/math.hpp + world.hpp.
#include "kat.hpp"
using namespace orc;

static const double PI = 3.14159265358979323846;
static const double S2 = std::sqrt(2.0);
using L = std::initializer_list<double>;
static Matrix M4(L l) { return Matrix::from_list(4, 4, l); }

// matrix.rs:268-282 (#[should_panic]: 16 values for a 5x5)
KAT(matrix_create_wrong_dimensions_panics) {
  CHECK_THROWS(Matrix::from_list(5, 5, {1, 2, 3, 4, 5.5, 6.5, 7.5, 8.5, 9, 10, 11, 12, 13.5, 14.5, 15.5, 16.5}));
}
// matrix.rs:284-308
KAT(matrix_create_4x4) {
  Matrix m = M4({1, 2, 3, 4, 5.5, 6.5, 7.5, 8.5, 9, 10, 11, 12, 13.5, 14.5, 15.5, 16.5});
  CHECK_EQ(m.rows, 4u); CHECK_EQ(m.columns, 4u);
  CHECK_NEAR(m(0, 0), 1.0); CHECK_NEAR(m(0, 3), 4.0); CHECK_NEAR(m(1, 0), 5.5); CHECK_NEAR(m(1, 2), 7.5);
  CHECK_NEAR(m(2, 2), 11.0); CHECK_NEAR(m(3, 0), 13.5); CHECK_NEAR(m(3, 2), 15.5);
}
// matrix.rs:310-333
KAT(matrix_create_2x2_3x3) {
  Matrix m = Matrix::from_list(2, 2, {-3, 5, 1, -2});
  CHECK_EQ(m.rows, 2u); CHECK_EQ(m.columns, 2u);
  CHECK_NEAR(m(0, 0), -3.0); CHECK_NEAR(m(0, 1), 5.0); CHECK_NEAR(m(1, 0), 1.0); CHECK_NEAR(m(1, 1), -2.0);
  Matrix n = Matrix::from_list(3, 3, {-3, 5, 0, 1, -2, -7, 0, 1, 1});
  CHECK_EQ(n.rows, 3u); CHECK_EQ(n.columns, 3u);
  CHECK_NEAR(n(0, 0), -3.0); CHECK_NEAR(n(1, 1), -2.0); CHECK_NEAR(n(2, 2), 1.0);
}
// matrix.rs:335-374
KAT(matrix_equality) {
  Matrix a = M4({1, 2, 3, 4, 5, 6, 7, 8, 9, 8, 7, 6, 5, 4, 3, 2});
  Matrix b = M4({1, 2, 3, 4, 5, 6, 7, 8, 9, 8, 7, 6, 5, 4, 3, 2});
  Matrix c = M4({2, 3, 4, 5, 6, 7, 8, 9, 8, 7, 6, 5, 4, 3, 2, 1});
  CHECK(a == b);
  CHECK(a != c);
}
// matrix.rs:376-402
KAT(matrix_multiply_two_matrices) {
  Matrix a = M4({1, 2, 3, 4, 5, 6, 7, 8, 9, 8, 7, 6, 5, 4, 3, 2});
  Matrix b = M4({-2, 1, 2, 3, 3, 2, 1, -1, 4, 3, 6, 5, 1, 2, 7, 8});
  Matrix e = M4({20, 22, 50, 48, 44, 54, 114, 108, 40, 58, 110, 102, 16, 26, 46, 42});
  CHECK(a * b == e);
}
// matrix.rs:404-426
KAT(matrix_multiply_by_point_and_vector) {
  Matrix a = M4({1, 2, 3, 4, 2, 4, 4, 2, 8, 6, 4, 1, 0, 0, 0, 1});
  CHECK(a * Point(1, 2, 3) == Point(18, 24, 33));
  CHECK(a * Vector(1, 2, 3) == Vector(14, 22, 32));
}
// matrix.rs:428-442
KAT(matrix_multiply_by_identity) {
  Matrix a = M4({0, 1, 2, 4, 1, 2, 4, 8, 2, 4, 8, 16, 4, 8, 16, 32});
  CHECK(a * Matrix::identity(4, 4) == a);
}
// matrix.rs:444-464
KAT(matrix_transpose) {
  Matrix a = M4({0, 9, 3, 0, 9, 8, 0, 8, 1, 8, 5, 3, 0, 0, 5, 8});
  Matrix e = M4({0, 9, 1, 0, 9, 8, 8, 0, 3, 0, 5, 5, 0, 8, 3, 8});
  CHECK(a.transpose() == e);
  Matrix id = Matrix::identity(4, 4);
  CHECK(id.transpose() == id);
}
// matrix.rs:466-470
KAT(matrix_determinant_2x2) { CHECK_NEAR(Matrix::from_list(2, 2, {1, 5, -3, 2}).determinant(), 17.0); }
// matrix.rs:472-500
KAT(matrix_submatrix) {
  Matrix a = Matrix::from_list(3, 3, {1, 5, 0, -3, 2, 7, 0, 6, 3});
  Matrix s = a.submatrix(0, 2);
  CHECK_EQ(s.rows, 2u); CHECK_EQ(s.columns, 2u);
  CHECK(s == Matrix::from_list(2, 2, {-3, 2, 0, 6}));
  Matrix b = M4({-6, 1, 1, 6, -8, 5, 8, 6, -1, 0, 8, 2, -7, 1, -1, 1});
  Matrix t = b.submatrix(2, 1);
  CHECK_EQ(t.rows, 3u); CHECK_EQ(t.columns, 3u);
  CHECK(t == Matrix::from_list(3, 3, {-6, 1, 6, -8, 8, 6, -7, -1, 1}));
}
// matrix.rs:502-518
KAT(matrix_minor_cofactor_3x3) {
  Matrix a = Matrix::from_list(3, 3, {3, 5, 0, 2, -1, -7, 6, -1, 5});
  CHECK_NEAR(a.submatrix(1, 0).determinant(), 25.0);
  CHECK_NEAR(a.minor(1, 0), 25.0);
  CHECK_NEAR(a.minor(0, 0), -12.0); CHECK_NEAR(a.cofactor(0, 0), -12.0);
  CHECK_NEAR(a.cofactor(1, 0), -25.0);
}
// matrix.rs:520-546
KAT(matrix_determinant_3x3_4x4) {
  Matrix a = Matrix::from_list(3, 3, {1, 2, 6, -5, 8, -4, 2, 6, 4});
  CHECK_NEAR(a.cofactor(0, 0), 56.0); CHECK_NEAR(a.cofactor(0, 1), 12.0); CHECK_NEAR(a.cofactor(0, 2), -46.0);
  CHECK_NEAR(a.determinant(), -196.0);
  Matrix b = M4({-2, -8, 3, 5, -3, 1, 7, 3, 1, 2, -9, 6, -6, 7, 7, -9});
  CHECK_NEAR(b.cofactor(0, 0), 690.0); CHECK_NEAR(b.cofactor(0, 1), 447.0); CHECK_NEAR(b.cofactor(0, 2), 210.0);
  CHECK_NEAR(b.cofactor(0, 3), 51.0);
  CHECK_NEAR(b.determinant(), -4071.0);
}
// matrix.rs:548-578
KAT(matrix_invertibility) {
  Matrix a = M4({6, 4, 4, 4, 5, 5, 7, 6, 4, -9, 3, -7, 9, 1, 7, -6});
  CHECK_NEAR(a.determinant(), -2120.0); CHECK(a.is_invertible());
  Matrix b = M4({-4, 2, -2, -3, 9, 6, 2, 6, 0, -5, 1, -5, 0, 0, 0, 0});
  CHECK_NEAR(b.determinant(), 0.0); CHECK(!b.is_invertible());
  CHECK_THROWS(b.inverse());  // matrix.rs:139: assert!(self.is_invertible())
}
// matrix.rs:580-612
KAT(matrix_inverse_1) {
  Matrix a = M4({-5, 2, 6, -8, 1, -5, 1, 8, 7, 7, -6, -7, 1, -3, 7, 4});
  Matrix b = a.inverse();
  CHECK_NEAR(a.determinant(), 532.0);
  CHECK_NEAR(a.cofactor(2, 3), -160.0); CHECK_NEAR(b(3, 2), -160.0 / 532.0);
  CHECK_NEAR(a.cofactor(3, 2), 105.0); CHECK_NEAR(b(2, 3), 105.0 / 532.0);
  Matrix e = M4({0.21805, 0.45113, 0.24060, -0.04511, -0.80827, -1.45677, -0.44361, 0.52068,
                 -0.07895, -0.22368, -0.05263, 0.19737, -0.52256, -0.81391, -0.30075, 0.30639});
  CHECK(b == e);
}
// matrix.rs:614-639
KAT(matrix_inverse_3) {
  Matrix a = M4({9, 3, 0, 9, -5, -2, -6, -3, -4, 9, 6, 4, -7, 6, 6, 2});
  Matrix e = M4({-0.04074, -0.07778, 0.14444, -0.22222, -0.07778, 0.03333, 0.36667, -0.33333,
                 -0.02901, -0.14630, -0.10926, 0.12963, 0.17778, 0.06667, -0.26667, 0.33333});
  CHECK(a.inverse() == e);
}
// matrix.rs:641-660
KAT(matrix_product_times_inverse) {
  Matrix a = M4({3, -9, 7, 3, 3, -8, 2, -9, -4, 4, 4, 1, -6, 5, -1, 1});
  Matrix b = M4({8, 2, 2, 2, 3, -1, 7, 0, 7, 0, 5, 4, 6, -2, 0, 5});
  Matrix c = a * b;
  CHECK(c * b.inverse() == a);
}
// matrix.rs:662-669
KAT(matrix_fluent_api_chaining) {
  Matrix t = Matrix::identity(4, 4).rotate_x(PI / 2.0).scale(5, 5, 5).translate(10, 5, 7);
  CHECK(t * Point(1, 0, 1) == Point(15, 0, 7));
}

// transform.rs:99-121
KAT(transform_translation) {
  Matrix t = translation(5, -3, 2);
  CHECK(t * Point(-3, 4, 5) == Point(2, 1, 7));
  CHECK(t.inverse() * Point(-3, 4, 5) == Point(-8, 7, 3));
  CHECK(t * Vector(-3, 4, 5) == Vector(-3, 4, 5));
}
// transform.rs:123-154
KAT(transform_scaling) {
  Matrix t = scaling(2, 3, 4);
  CHECK(t * Point(-4, 6, 8) == Point(-8, 18, 32));
  CHECK(t * Vector(-4, 6, 8) == Vector(-8, 18, 32));
  CHECK(t.inverse() * Vector(-4, 6, 8) == Vector(-2, 2, 2));
  CHECK(scaling(-1, 1, 1) * Point(2, 3, 4) == Point(-2, 3, 4));
}
// transform.rs:156-201
KAT(transform_rotations) {
  Point p(0, 1, 0);
  CHECK(rotation_x(PI / 4.0) * p == Point(0.0, S2 / 2.0, S2 / 2.0));
  CHECK(rotation_x(PI / 2.0) * p == Point(0, 0, 1));
  CHECK(rotation_x(PI / 4.0).inverse() * p == Point(0.0, S2 / 2.0, -(S2 / 2.0)));
  CHECK(rotation_y(PI / 4.0) * Point(0, 0, 1) == Point(S2 / 2.0, 0.0, S2 / 2.0));
  CHECK(rotation_y(PI / 2.0) * Point(0, 0, 1) == Point(1, 0, 0));
  CHECK(rotation_z(PI / 4.0) * p == Point(-(S2 / 2.0), S2 / 2.0, 0.0));
  CHECK(rotation_z(PI / 2.0) * p == Point(-1, 0, 0));
}
// transform.rs:203-249
KAT(transform_shearing) {
  Point p(2, 3, 4);
  CHECK(shearing(1, 0, 0, 0, 0, 0) * p == Point(5, 3, 4));
  CHECK(shearing(0, 1, 0, 0, 0, 0) * p == Point(6, 3, 4));
  CHECK(shearing(0, 0, 1, 0, 0, 0) * p == Point(2, 5, 4));
  CHECK(shearing(0, 0, 0, 1, 0, 0) * p == Point(2, 7, 4));
  CHECK(shearing(0, 0, 0, 0, 1, 0) * p == Point(2, 3, 6));
  CHECK(shearing(0, 0, 0, 0, 0, 1) * p == Point(2, 3, 7));
}
// transform.rs:251-277
KAT(transform_sequence_and_chain) {
  Point p(1, 0, 1);
  Matrix a = rotation_x(PI / 2.0), b = scaling(5, 5, 5), c = translation(10, 5, 7);
  Point p2 = a * p;
  CHECK(p2 == Point(1, -1, 0));
  Point p3 = b * p2;
  CHECK(p3 == Point(5, -5, 0));
  Point p4 = c * p3;
  CHECK(p4 == Point(15, 0, 7));
  Matrix t = (c * b) * a;
  CHECK(t * p == Point(15, 0, 7));
}
// transform.rs:279-326
KAT(transform_view_transform) {
  CHECK(view_transform(Point(0, 0, 0), Point(0, 0, -1), Vector(0, 1, 0)) == Matrix::identity(4, 4));
  CHECK(view_transform(Point(0, 0, 0), Point(0, 0, 1), Vector(0, 1, 0)) == scaling(-1, 1, -1));
  CHECK(view_transform(Point(0, 0, 8), Point(0, 0, 0), Vector(0, 1, 0)) == translation(0, 0, -8));
  Matrix t = view_transform(Point(1, 3, 2), Point(4, -2, 8), Vector(1, 1, 0));
  Matrix e = M4({-0.50709, 0.50709, 0.67612, -2.36643, 0.76772, 0.60609, 0.12122, -2.82843,
                 -0.35857, 0.59761, -0.71714, 0.0, 0.0, 0.0, 0.0, 1.0});
  CHECK(t == e);
}

// vector.rs:114-191
KAT(vector_arithmetic) {
  Vector a(4.3, -4.2, 3.1);
  CHECK_NEAR(a.x, 4.3); CHECK_NEAR(a.y, -4.2); CHECK_NEAR(a.z, 3.1);
  CHECK(Point(3, -2, 5) + Vector(-2, 3, 1) == Point(1, 1, 6));
  CHECK(Vector(1, -2, 3) + Vector(-2, -3, 4) == Vector(-1, -5, 7));
  CHECK(Vector(-2, -3, 4) + Vector(1, -2, 3) == Vector(-1, -5, 7));
  CHECK(Vector(3, 2, 1) - Vector(5, 6, 7) == Vector(-2, -4, -6));
  CHECK(Vector(0, 0, 0) - Vector(1, -2, 3) == Vector(-1, 2, -3));
  CHECK(-Vector(1, -2, 3) == Vector(-1, 2, -3));
  CHECK(Vector(1, -2, 3) * 3.5 == Vector(3.5, -7.0, 10.5));
  CHECK(Vector(1, -2, 3) * 0.5 == Vector(0.5, -1.0, 1.5));
  CHECK(Vector(1, -2, 3) / 2.0 == Vector(0.5, -1.0, 1.5));
}
// vector.rs:193-222
KAT(vector_magnitude_normalize) {
  CHECK_NEAR(Vector(1, 0, 0).magnitude(), 1.0); CHECK_NEAR(Vector(0, 1, 0).magnitude(), 1.0);
  CHECK_NEAR(Vector(0, 0, 1).magnitude(), 1.0);
  CHECK_NEAR(Vector(1, 2, 3).magnitude(), std::sqrt(14.0)); CHECK_NEAR(Vector(-1, -2, -3).magnitude(), std::sqrt(14.0));
  CHECK(Vector(4, 0, 0).normalize() == Vector(1, 0, 0));
  CHECK(Vector(1, 2, 3).normalize() == Vector(0.26726, 0.53452, 0.80178));
  CHECK_NEAR(Vector(1, 2, 3).normalize().magnitude(), 1.0);
}
// vector.rs:224-250
KAT(vector_dot_cross_reflect) {
  Vector a(1, 2, 3), b(2, 3, 4);
  CHECK_NEAR(dot(a, b), 20.0); CHECK_NEAR(dot(b, a), 20.0);
  CHECK(cross(a, b) == Vector(-1, 2, -1));
  CHECK(cross(b, a) == Vector(1, -2, 1));
  CHECK(reflect(Vector(1, -1, 0), Vector(0, 1, 0)) == Vector(1, 1, 0));
  CHECK(reflect(Vector(0, -1, 0), Vector(S2 / 2.0, S2 / 2.0, 0.0)) == Vector(1, 0, 0));
}
// point.rs:66-96
KAT(point_arithmetic) {
  Point a(4.3, -4.2, 3.1);
  CHECK_NEAR(a.x, 4.3); CHECK_NEAR(a.y, -4.2); CHECK_NEAR(a.z, 3.1);
  CHECK(Point(3, -2, 5) + Vector(-2, 3, 1) == Point(1, 1, 6));
  CHECK(Point(3, 2, 1) - Point(5, 6, 7) == Vector(-2, -4, -6));
  CHECK(Point(3, 2, 1) - Vector(5, 6, 7) == Point(-2, -4, -6));
}
// color.rs:109-150, color.rs:26-33 (average)
KAT(color_arithmetic) {
  Color c(-0.5, 0.4, 1.7);
  CHECK_NEAR(c.red, -0.5); CHECK_NEAR(c.green, 0.4); CHECK_NEAR(c.blue, 1.7);
  CHECK(Color(0.9, 0.6, 0.75) + Color(0.7, 0.1, 0.25) == Color(1.6, 0.7, 1.0));
  CHECK(Color(0.9, 0.6, 0.75) - Color(0.7, 0.1, 0.25) == Color(0.2, 0.5, 0.5));
  CHECK(Color(0.2, 0.3, 0.4) * 2.0 == Color(0.4, 0.6, 0.8));
  CHECK(Color(1.0, 0.2, 0.4) * Color(0.9, 1.0, 0.1) == Color(0.9, 0.2, 0.04));
  CHECK(color_average({Color(1, 0, 0), Color(0, 1, 0), Color(0, 0, 1), Color(1, 1, 1)}) == Color(0.5, 0.5, 0.5));
}
// ray.rs:37-72
KAT(ray_create_position_transform) {
  Ray r(Point(1, 2, 3), Vector(4, 5, 6));
  CHECK(r.origin == Point(1, 2, 3)); CHECK(r.direction == Vector(4, 5, 6));
  Ray q(Point(2, 3, 4), Vector(1, 0, 0));
  CHECK(q.position(0) == Point(2, 3, 4)); CHECK(q.position(1) == Point(3, 3, 4));
  CHECK(q.position(-1) == Point(1, 3, 4)); CHECK(q.position(2.5) == Point(4.5, 3, 4));
  Ray s(Point(1, 2, 3), Vector(0, 1, 0));
  Ray s2 = s.transform(translation(3, 4, 5));
  CHECK(s2.origin == Point(4, 6, 8)); CHECK(s2.direction == Vector(0, 1, 0));
  Ray s3 = s.transform(scaling(2, 3, 4));
  CHECK(s3.origin == Point(2, 6, 12)); CHECK(s3.direction == Vector(0, 3, 0));
}
// light.rs:30-37
KAT(light_create_point_light) {
  PointLight l(Point::origin(), Color(1, 1, 1));
  CHECK(l.position == Point::origin()); CHECK(l.intensity == Color(1, 1, 1));
}
// canvas.rs:59-99
KAT(canvas_create_index_write) {
  Canvas c(10, 20);
  CHECK_EQ(c.width, 10u); CHECK_EQ(c.height, 20u);
  for (const Color& p : c.pixels) CHECK(p == Color(0, 0, 0));
  CHECK_EQ(c.pixel_idx(5, 0), 5u); CHECK_EQ(c.pixel_idx(5, 1), 15u); CHECK_EQ(c.pixel_idx(0, 2), 20u);
  CHECK_EQ(c.pixel_idx(9, 19), 199u);
  CHECK_THROWS(c.pixel_idx(10, 1));
  CHECK_THROWS(c.pixel_idx(5, 200));
  c.set_pixel(2, 3, Color(1, 0, 0));
  CHECK(c.get_pixel(2, 3) == Color(1, 0, 0));
}
