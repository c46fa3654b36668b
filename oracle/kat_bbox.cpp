// ORACLE KATs — bounding_box.rs:171-509 test module restated.
#include "kat.hpp"
using namespace orc;

static const double PI = 3.14159265358979323846;
static const double SQRT_2 = std::sqrt(2.0);

// bounding_box.rs:187-222
KAT(bbox_create_add) {
  BoundingBox e;
  CHECK(std::isinf(e.min.x) && e.min.x > 0 && std::isinf(e.min.y) && e.min.y > 0 && std::isinf(e.min.z) && e.min.z > 0);
  CHECK(std::isinf(e.max.x) && e.max.x < 0 && std::isinf(e.max.y) && e.max.y < 0 && std::isinf(e.max.z) && e.max.z < 0);
  BoundingBox v(Point(-1, -2, -3), Point(3, 2, 1));
  CHECK_EQ(v.min, Point(-1, -2, -3)); CHECK_EQ(v.max, Point(3, 2, 1));
  BoundingBox bb;
  bb.add_point(Point(-5, 2, 0));
  bb.add_point(Point(7, 0, -3));
  CHECK_EQ(bb.min, Point(-5, 0, -3)); CHECK_EQ(bb.max, Point(7, 2, 0));
  BoundingBox box1(Point(-5, -2, 0), Point(7, 4, 4));
  box1.add_bounding_box(BoundingBox(Point(8, -7, -2), Point(14, 2, 8)));
  CHECK_EQ(box1.min, Point(-5, -7, -2)); CHECK_EQ(box1.max, Point(14, 4, 8));
}
// bounding_box.rs:225-264
KAT(bbox_contains) {
  BoundingBox bb(Point(5, -2, 0), Point(11, 4, 7));
  struct P { Point p; bool r; };
  P pts[] = {{Point(5, -2, 0), true}, {Point(11, 4, 7), true}, {Point(8, 1, 3), true}, {Point(3, 0, 3), false},
             {Point(8, -4, 3), false}, {Point(8, 1, -1), false}, {Point(13, 1, 3), false}, {Point(8, 5, 3), false},
             {Point(8, 1, 8), false}};
  for (auto& t : pts) CHECK_EQ(bb.contains_point(t.p), t.r);
  struct B { Point mn, mx; bool r; };
  B boxes[] = {{Point(5, -2, 0), Point(11, 4, 7), true}, {Point(6, -1, 1), Point(10, 3, 6), true},
               {Point(4, -3, -1), Point(10, 3, 6), false}, {Point(6, -1, 1), Point(12, 5, 8), false}};
  for (auto& t : boxes) CHECK_EQ(bb.contains_bounding_box(BoundingBox(t.mn, t.mx)), t.r);
}
// bounding_box.rs:267-285
KAT(bbox_transform_and_parent_space) {
  BoundingBox bb(Point(-1, -1, -1), Point(1, 1, 1));
  BoundingBox bb2 = bb.transform(rotation_x(PI / 4.0) * rotation_y(PI / 4.0));
  CHECK_EQ(bb2.min, Point(-SQRT_2, -1.70711, -1.70711));
  CHECK_EQ(bb2.max, Point(SQRT_2, 1.70711, 1.70711));
  Sphere shape;
  shape.set_transform(translation(1, -3, 5) * scaling(0.5, 2.0, 4.0));
  BoundingBox p = shape.parent_space_bounds();
  CHECK_EQ(p.min, Point(0.5, -5.0, 1.0));
  CHECK_EQ(p.max, Point(1.5, -1.0, 9.0));
}
// bounding_box.rs:287-368 — 26 ray/box cases
KAT(bbox_ray_intersection) {
  struct T { Point o; Vector d; bool r; };
  BoundingBox unit(Point(-1, -1, -1), Point(1, 1, 1));
  T a[] = {{Point(5.0, 0.5, 0.0), Vector(-1, 0, 0), true}, {Point(-5.0, 0.5, 0.0), Vector(1, 0, 0), true},
           {Point(0.5, 5.0, 0.0), Vector(0, -1, 0), true}, {Point(0.5, -5.0, 0.0), Vector(0, 1, 0), true},
           {Point(0.5, 0.0, 5.0), Vector(0, 0, -1), true}, {Point(0.5, 0.0, -5.0), Vector(0, 0, 1), true},
           {Point(0.0, 0.5, 0.0), Vector(0, 0, 1), true}, {Point(-2.0, 0.0, 0.0), Vector(2, 4, 6), false},
           {Point(0.0, -2.0, 0.0), Vector(6, 2, 4), false}, {Point(0.0, 0.0, -2.0), Vector(4, 6, 2), false},
           {Point(2.0, 0.0, 2.0), Vector(0, 0, -1), false}, {Point(0.0, 2.0, 2.0), Vector(0, -1, 0), false},
           {Point(2.0, 2.0, 0.0), Vector(-1, 0, 0), false}};
  for (auto& t : a) CHECK_EQ(unit.intersects(Ray(t.o, t.d.normalize())), t.r);
  BoundingBox nc(Point(5, -2, 0), Point(11, 4, 7));
  T b[] = {{Point(15, 1, 2), Vector(-1, 0, 0), true}, {Point(-5, -1, 4), Vector(1, 0, 0), true},
           {Point(7, 6, 5), Vector(0, -1, 0), true}, {Point(9, -5, 6), Vector(0, 1, 0), true},
           {Point(8, 2, 12), Vector(0, 0, -1), true}, {Point(6, 0, -5), Vector(0, 0, 1), true},
           {Point(8.0, 1.0, 3.5), Vector(0, 0, 1), true}, {Point(9, -1, -8), Vector(2, 4, 6), false},
           {Point(8, 3, -4), Vector(6, 2, 4), false}, {Point(9, -1, -2), Vector(4, 6, 2), false},
           {Point(4, 0, 9), Vector(0, 0, -1), false}, {Point(8, 6, -1), Vector(0, -1, 0), false},
           {Point(12, 5, 4), Vector(-1, 0, 0), false}};
  for (auto& t : b) CHECK_EQ(nc.intersects(Ray(t.o, t.d.normalize())), t.r);
}
// bounding_box.rs:373-452 — cull proofs with the TestShape double
KAT(bbox_group_and_csg_cull_children) {
  {
    Group shape;
    shape.add_child(std::make_unique<TestShape>());
    Intersections xs;
    shape.intersect(Ray(Point(0, 0, -5), Vector(0, 1, 0)), xs);
    const TestShape* child = static_cast<const TestShape*>(shape.children[0].get());
    CHECK_EQ(child->saved_ray.origin, Point::origin());
    CHECK_EQ(child->saved_ray.direction, Vector(0, 0, 0));
    Ray r(Point(0, 0, -5), Vector(0, 0, 1));
    shape.intersect(r, xs);
    CHECK_EQ(child->saved_ray.origin, r.origin);
    CHECK_EQ(child->saved_ray.direction, r.direction);
  }
  {
    Csg shape(Operation::Difference, std::make_unique<TestShape>(), std::make_unique<TestShape>());
    Intersections xs;
    shape.intersect(Ray(Point(0, 0, -5), Vector(0, 1, 0)), xs);
    const TestShape* l = static_cast<const TestShape*>(shape.left.get());
    const TestShape* rr = static_cast<const TestShape*>(shape.right.get());
    CHECK_EQ(l->saved_ray.direction, Vector(0, 0, 0));
    CHECK_EQ(rr->saved_ray.direction, Vector(0, 0, 0));
    Ray r(Point(0, 0, -5), Vector(0, 0, 1));
    shape.intersect(r, xs);
    CHECK_EQ(l->saved_ray.origin, r.origin); CHECK_EQ(l->saved_ray.direction, r.direction);
    CHECK_EQ(rr->saved_ray.origin, r.origin); CHECK_EQ(rr->saved_ray.direction, r.direction);
  }
}
// bounding_box.rs:456-508
KAT(bbox_split_and_primitive_divide) {
  auto h = BoundingBox(Point(-1, -4, -5), Point(9, 6, 5)).split();
  CHECK_EQ(h.first.min, Point(-1, -4, -5)); CHECK_EQ(h.first.max, Point(4, 6, 5));
  CHECK_EQ(h.second.min, Point(4, -4, -5)); CHECK_EQ(h.second.max, Point(9, 6, 5));
  h = BoundingBox(Point(-1, -2, -3), Point(9.0, 5.5, 3.0)).split();
  CHECK_EQ(h.first.max, Point(4.0, 5.5, 3.0)); CHECK_EQ(h.second.min, Point(4.0, -2.0, -3.0));
  h = BoundingBox(Point(-1, -2, -3), Point(5, 8, 3)).split();
  CHECK_EQ(h.first.max, Point(5, 3, 3)); CHECK_EQ(h.second.min, Point(-1, 3, -3));
  h = BoundingBox(Point(-1, -2, -3), Point(5, 3, 7)).split();
  CHECK_EQ(h.first.max, Point(5, 3, 2)); CHECK_EQ(h.second.min, Point(-1, -2, 2));
  Sphere s, d;
  s.divide(1);
  CHECK(s.equals(d));
}
// bounding_box.rs:76-93 with geometry/mod.rs:74-85 — quirks the flattener must reproduce:
// a plane's infinite box turns into the empty (inverted) box on any set_transform, and that empty
// box makes its group's box infinite; boxes of rotated shapes inflate on every re-parenting.
KAT(bbox_quirks_plane_and_inflation) {
  Plane p;
  p.set_transform(translation(0, 1, 0));
  CHECK(p.get_bounds().min.x == INF && p.get_bounds().max.x == -INF);
  CHECK(p.get_bounds().min.y == INF && p.get_bounds().max.y == -INF);
  Group g;
  g.add_child(std::make_unique<Plane>());
  CHECK(g.get_bounds().min.x == -INF && g.get_bounds().max.x == INF);
  CHECK(g.get_bounds().min.y == -INF && g.get_bounds().max.y == INF);
  CHECK(g.get_bounds().intersects(Ray(Point(0, 5, 0), Vector(0, 1, 0))));
  Sphere s;
  s.set_transform(rotation_y(PI / 4.0));
  double w0 = s.get_bounds().max.x;
  CHECK_NEAR(w0, SQRT_2);
  s.set_transform(rotation_y(PI / 4.0));  // same matrix again: un-rotate the AABB, re-rotate → x2
  CHECK_NEAR(s.get_bounds().max.x, 2.0 * SQRT_2);
}
