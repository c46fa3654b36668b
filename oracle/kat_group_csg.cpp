// ORACLE KATs — geometry/shape/group.rs:199-391 and csg.rs:142-358 test modules restated.
#include "kat.hpp"
using namespace orc;

static Intersections lisect(const Shape& s, const Ray& r) {
  Intersections xs;
  s.local_intersect(r, xs);
  return xs;
}
static std::unique_ptr<Sphere> sphere_at(const Matrix& m) {
  auto s = std::make_unique<Sphere>();
  s->set_transform(m);
  return s;
}
static const Group& as_group(const Shape& s) {
  CHECK(s.kind() == ShapeKind::Group);
  return static_cast<const Group&>(s);
}

// group.rs:215-238
KAT(group_create_add_empty) {
  Group g;
  CHECK(g.children.empty());
  CHECK(g.transform() == Matrix::identity(4, 4));
  CHECK(lisect(g, Ray(Point::origin(), Vector(0, 0, 1))).empty());
  g.add_child(std::make_unique<Sphere>());
  CHECK(!g.children.empty());
}
// group.rs:241-264
KAT(group_intersect_nonempty) {
  Group g;
  g.add_child(std::make_unique<Sphere>());
  g.add_child(sphere_at(translation(0, 0, -3)));
  g.add_child(sphere_at(translation(5, 0, 0)));
  const Shape* s1 = g.children[0].get();
  const Shape* s2 = g.children[1].get();
  auto xs = intersections(lisect(g, Ray(Point(0, 0, -5), Vector(0, 0, 1))));
  CHECK_EQ(xs.size(), 4u);
  CHECK(xs[0].object == s2); CHECK(xs[1].object == s2); CHECK(xs[2].object == s1); CHECK(xs[3].object == s1);
}
// group.rs:267-277
KAT(group_intersect_transformed) {
  Group g;
  g.set_transform(scaling(2, 2, 2));
  g.add_child(sphere_at(translation(5, 0, 0)));
  Intersections xs;
  g.intersect(Ray(Point(10, 0, -10), Vector(0, 0, 1)), xs);
  CHECK_EQ(xs.size(), 2u);
}
// group.rs:280-297
KAT(group_bounding_box_contains_children) {
  auto s = sphere_at(translation(2, 5, -3) * scaling(2, 2, 2));
  auto c = std::make_unique<Cylinder>(-2, 2, false);
  c->set_transform(translation(-4, -1, 4) * scaling(0.5, 1.0, 0.5));
  Group shape;
  shape.add_child(std::move(s));
  shape.add_child(std::move(c));
  CHECK_EQ(shape.get_bounds().min, Point(-4.5, -3.0, -5.0));
  CHECK_EQ(shape.get_bounds().max, Point(4.0, 7.0, 4.5));
}
// group.rs:300-325
KAT(group_partition_children) {
  Group g;
  g.add_child(sphere_at(translation(-2, 0, 0)));
  g.add_child(sphere_at(translation(2, 0, 0)));
  g.add_child(std::make_unique<Sphere>());
  std::vector<ShapeP> left, right;
  g.partition_children(left, right);
  CHECK_EQ(g.children.size(), 1u);
  CHECK(g.children[0]->transform() == Matrix::identity(4, 4));
  CHECK_EQ(left.size(), 1u);
  CHECK(left[0]->transform() == translation(-2, 0, 0));
  CHECK_EQ(right.size(), 1u);
  CHECK(right[0]->transform() == translation(2, 0, 0));
}
// group.rs:328-343
KAT(group_make_subgroup) {
  Group g;
  std::vector<ShapeP> v;
  v.push_back(std::make_unique<Sphere>());
  v.push_back(std::make_unique<Sphere>());
  g.make_subgroup(std::move(v));
  CHECK_EQ(g.children.size(), 1u);
  const Group& g0 = as_group(*g.children[0]);
  CHECK_EQ(g0.children.size(), 2u);
  Sphere d;
  CHECK(g0.children[0]->equals(d));
  CHECK(g0.children[1]->equals(d));
}
// group.rs:346-391
KAT(group_divide) {
  Group g;
  g.add_child(sphere_at(translation(-2, -2, 0)));
  g.add_child(sphere_at(translation(-2, 2, 0)));
  g.add_child(sphere_at(scaling(4, 4, 4)));
  g.divide(1);
  CHECK(g.children[0]->kind() == ShapeKind::Sphere);
  CHECK(g.children[0]->transform() == scaling(4, 4, 4));
  const Group& sub = as_group(*g.children[1]);
  CHECK_EQ(sub.children.size(), 2u);
  const Group& c0 = as_group(*sub.children[0]);
  CHECK(c0.children[0]->transform() == translation(-2, -2, 0));
  const Group& c1 = as_group(*sub.children[1]);
  CHECK(c1.children[0]->transform() == translation(-2, 2, 0));
}

// csg.rs:155-166
KAT(csg_create) {
  Csg c(Operation::Union, std::make_unique<Sphere>(), std::make_unique<Cube>());
  CHECK(c.operation == Operation::Union);
  CHECK(c.left->kind() == ShapeKind::Sphere);
  CHECK(c.right->kind() == ShapeKind::Cube);
}
// csg.rs:169-227 — the 24-row truth table
KAT(csg_truth_table) {
  struct T { Operation op; bool lhit, inl, inr, result; };
  const Operation U = Operation::Union, I = Operation::Intersection, D = Operation::Difference;
  T cases[] = {{U, true, true, true, false},   {U, true, true, false, true},   {U, true, false, true, false},
               {U, true, false, false, true},  {U, false, true, true, false},  {U, false, true, false, false},
               {U, false, false, true, true},  {U, false, false, false, true}, {I, true, true, true, true},
               {I, true, true, false, false},  {I, true, false, true, true},   {I, true, false, false, false},
               {I, false, true, true, true},   {I, false, true, false, true},  {I, false, false, true, false},
               {I, false, false, false, false}, {D, true, true, true, false},  {D, true, true, false, true},
               {D, true, false, true, false},  {D, true, false, false, true},  {D, false, true, true, true},
               {D, false, true, false, true},  {D, false, false, true, false}, {D, false, false, false, false}};
  for (auto& t : cases) CHECK_EQ(intersection_allowed(t.op, t.lhit, t.inl, t.inr), t.result);
}
// csg.rs:230-268
KAT(csg_filter_intersections) {
  struct T { Operation op; size_t x0, x1; };
  T cases[] = {{Operation::Union, 0, 3}, {Operation::Intersection, 1, 2}, {Operation::Difference, 0, 1}};
  for (auto& t : cases) {
    Csg c(t.op, std::make_unique<Sphere>(), std::make_unique<Cube>());
    const Shape* s1 = c.left.get();
    const Shape* s2 = c.right.get();
    Intersections xs = {Intersection(1.0, s1), Intersection(2.0, s2), Intersection(3.0, s1), Intersection(4.0, s2)};
    Intersections result = c.filter_intersections(xs);
    CHECK_EQ(result.size(), 2u);
    CHECK_NEAR(result[0].t, xs[t.x0].t);
    CHECK_NEAR(result[1].t, xs[t.x1].t);
  }
}
// csg.rs:271-302
KAT(csg_ray_miss_hit_bounds) {
  Csg c(Operation::Union, std::make_unique<Sphere>(), std::make_unique<Cube>());
  CHECK(lisect(c, Ray(Point(0, 2, -5), Vector(0, 0, 1))).empty());
  Csg c2(Operation::Union, std::make_unique<Sphere>(), sphere_at(translation(0.0, 0.0, 0.5)));
  auto xs = lisect(c2, Ray(Point(0, 0, -5), Vector(0, 0, 1)));
  CHECK_EQ(xs.size(), 2u); CHECK_NEAR(xs[0].t, 4.0); CHECK_NEAR(xs[1].t, 6.5);
  CHECK(xs[0].object == c2.left.get()); CHECK(xs[1].object == c2.right.get());
  Csg c3(Operation::Difference, std::make_unique<Sphere>(), sphere_at(translation(2, 3, 4)));
  CHECK_EQ(c3.get_bounds().min, Point(-1, -1, -1));
  CHECK_EQ(c3.get_bounds().max, Point(3, 4, 5));
}
// csg.rs:305-358
KAT(csg_divide_subdivides_children) {
  auto left = std::make_unique<Group>();
  left->add_child(sphere_at(translation(-1.5, 0.0, 0.0)));
  left->add_child(sphere_at(translation(1.5, 0.0, 0.0)));
  auto right = std::make_unique<Group>();
  right->add_child(sphere_at(translation(0.0, 0.0, -1.5)));
  right->add_child(sphere_at(translation(0.0, 0.0, 1.5)));
  Csg shape(Operation::Difference, std::move(left), std::move(right));
  shape.divide(1);
  const Group& l = as_group(*shape.left);
  const Group& r = as_group(*shape.right);
  CHECK(as_group(*l.children[0]).children[0]->transform() == translation(-1.5, 0.0, 0.0));
  CHECK(as_group(*l.children[1]).children[0]->transform() == translation(1.5, 0.0, 0.0));
  CHECK(as_group(*r.children[0]).children[0]->transform() == translation(0.0, 0.0, -1.5));
  CHECK(as_group(*r.children[1]).children[0]->transform() == translation(0.0, 0.0, 1.5));
}
