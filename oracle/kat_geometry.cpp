// ORACLE KATs — geometry/shape/{triangle,smooth_triangle}.rs, geometry/mod.rs and
// geometry/intersection.rs test modules restated.
#include "kat.hpp"
using namespace orc;

static const double PI = 3.14159265358979323846;
static const double S2 = std::sqrt(2.0);
static Intersections lisect(const Shape& s, const Ray& r) {
  Intersections xs;
  s.local_intersect(r, xs);
  return xs;
}
static Intersection dummy(const Shape& s) { return Intersection(-100.0, &s); }
static Triangle tri() { return Triangle(Point(0, 1, 0), Point(-1, 0, 0), Point(1, 0, 0)); }
static SmoothTriangle stri() {
  return SmoothTriangle(Point(0, 1, 0), Point(-1, 0, 0), Point(1, 0, 0), Vector(0, 1, 0), Vector(-1, 0, 0), Vector(1, 0, 0));
}

// triangle.rs:105-196
KAT(triangle_all) {
  Triangle t = tri();
  CHECK_EQ(t.p1, Point(0, 1, 0)); CHECK_EQ(t.p2, Point(-1, 0, 0)); CHECK_EQ(t.p3, Point(1, 0, 0));
  CHECK_EQ(t.e1, Vector(-1, -1, 0)); CHECK_EQ(t.e2, Vector(1, -1, 0)); CHECK_EQ(t.normal, Vector(0, 0, -1));
  CHECK_EQ(t.local_normal_at(Point(0.0, 0.5, 0.0), dummy(t)), t.normal);
  CHECK_EQ(t.local_normal_at(Point(-0.5, 0.75, 0.0), dummy(t)), t.normal);
  CHECK_EQ(t.local_normal_at(Point(0.5, 0.25, 0.0), dummy(t)), t.normal);
  CHECK_EQ(lisect(t, Ray(Point(0, -1, -2), Vector(0, 1, 0))).size(), 0u);  // parallel
  CHECK_EQ(lisect(t, Ray(Point(1, 1, -2), Vector(0, 0, 1))).size(), 0u);   // p1-p3 edge
  CHECK_EQ(lisect(t, Ray(Point(-1, 1, -2), Vector(0, 0, 1))).size(), 0u);  // p1-p2 edge
  CHECK_EQ(lisect(t, Ray(Point(0, -1, -2), Vector(0, 0, 1))).size(), 0u);  // p2-p3 edge
  auto xs = lisect(t, Ray(Point(0.0, 0.5, -2.0), Vector(0, 0, 1)));
  CHECK_EQ(xs.size(), 1u); CHECK_NEAR(xs[0].t, 2.0); CHECK(!xs[0].has_uv);
  Triangle s(Point(-3, 7, 2), Point(6, 2, -4), Point(2, -1, -1));
  CHECK_EQ(s.get_bounds().min, Point(-3, -1, -4));
  CHECK_EQ(s.get_bounds().max, Point(6, 7, 2));
}
// smooth_triangle.rs:113-200
KAT(smooth_triangle_all) {
  SmoothTriangle t = stri();
  CHECK_EQ(t.n1, Vector(0, 1, 0)); CHECK_EQ(t.n2, Vector(-1, 0, 0)); CHECK_EQ(t.n3, Vector(1, 0, 0));
  Ray r(Point(-0.2, 0.3, -2.0), Vector(0, 0, 1));
  auto xs = lisect(t, r);
  CHECK_EQ(xs.size(), 1u); CHECK(xs[0].has_uv);
  CHECK_NEAR(xs[0].u, 0.45); CHECK_NEAR(xs[0].v, 0.25);
  Intersection i(1.0, &t, 0.45, 0.25);
  CHECK_EQ(t.normal_at(Point::origin(), i), Vector(-0.5547, 0.83205, 0.0));
  Computations comps = prepare_computations(i, r, {i});
  CHECK_EQ(comps.normalv, Vector(-0.5547, 0.83205, 0.0));
  SmoothTriangle s(Point(-3, 7, 2), Point(6, 2, -4), Point(2, -1, -1), Vector(0, 1, 0), Vector(-1, 0, 0), Vector(1, 0, 0));
  CHECK_EQ(s.get_bounds().min, Point(-3, -1, -4));
  CHECK_EQ(s.get_bounds().max, Point(6, 7, 2));
}
// geometry/mod.rs:127-153 — group transforms are baked into the leaf
KAT(normal_on_child_object) {
  Group g1;
  g1.set_transform(rotation_y(PI / 2.0));
  auto g2 = std::make_unique<Group>();
  g2->set_transform(scaling(1, 2, 3));
  auto s = std::make_unique<Sphere>();
  s->set_transform(translation(5, 0, 0));
  g2->add_child(std::move(s));
  g1.add_child(std::move(g2));
  const Group* g2p = static_cast<const Group*>(g1.children[0].get());
  const Shape* sp = g2p->children[0].get();
  CHECK_EQ(sp->normal_at(Point(1.7321, 1.1547, -5.5774), dummy(*sp)), Vector(0.2857, 0.42854, -0.85716));
}
// intersection.rs:175-236
KAT(intersection_aggregate_and_hit) {
  Sphere s;
  Intersection i(3.5, &s);
  CHECK_NEAR(i.t, 3.5); CHECK(i.object == &s);
  Intersection i1(1.0, &s), i2(2.0, &s);
  auto xs = intersections({i1, i2});
  CHECK_EQ(xs.size(), 2u); CHECK_NEAR(xs[0].t, 1.0); CHECK_NEAR(xs[1].t, 2.0);
  CHECK(*hit(xs) == i1);
  Intersection m1(-1.0, &s), p1(1.0, &s);
  xs = intersections({m1, p1});
  CHECK(*hit(xs) == p1);
  xs = intersections({Intersection(-2.0, &s), Intersection(-1.0, &s)});
  CHECK(hit(xs) == nullptr);
  Intersection i4(2.0, &s);
  xs = intersections({Intersection(5.0, &s), Intersection(7.0, &s), Intersection(-3.0, &s), i4});
  CHECK(*hit(xs) == i4);
}
// intersection.rs:238-295
KAT(intersection_prepare_computations) {
  Ray r(Point(0, 0, -5), Vector(0, 0, 1));
  Sphere shape;
  Intersection i(4.0, &shape);
  Computations c = prepare_computations(i, r, {i});
  CHECK_NEAR(c.t, i.t); CHECK(c.object == i.object);
  CHECK_EQ(c.point, Point(0, 0, -1)); CHECK_EQ(c.eyev, Vector(0, 0, -1)); CHECK_EQ(c.normalv, Vector(0, 0, -1));
  CHECK_EQ(c.inside, false);
  Ray r2(Point(0, 0, 0), Vector(0, 0, 1));
  Intersection j(1.0, &shape);
  c = prepare_computations(j, r2, {j});
  CHECK_EQ(c.point, Point(0, 0, 1)); CHECK_EQ(c.eyev, Vector(0, 0, -1)); CHECK_EQ(c.inside, true);
  CHECK_EQ(c.normalv, Vector(0, 0, -1));
  Sphere moved;
  moved.set_transform(translation(0, 0, 1));
  Intersection k(5.0, &moved);
  c = prepare_computations(k, r, {k});
  CHECK(c.over_point.z < -EPSILON / 2.0); CHECK(c.point.z > c.over_point.z);
  Plane plane;
  Ray r3(Point(0, 1, -1), Vector(0.0, -(S2 / 2.0), S2 / 2.0));
  Intersection l(S2, &plane);
  c = prepare_computations(l, r3, {l});
  CHECK_EQ(c.reflectv, Vector(0.0, S2 / 2.0, S2 / 2.0));
}
// intersection.rs:298-335 — the n1/n2 table of three nested glass spheres
KAT(intersection_n1_n2) {
  auto a = Sphere::glass();
  a->set_transform(scaling(2, 2, 2));
  kat::edit_material(*a, [](Material& m) { m.refractive_index = 1.5; });
  auto b = Sphere::glass();
  b->set_transform(translation(0.0, 0.0, -0.25));
  kat::edit_material(*b, [](Material& m) { m.refractive_index = 2.0; });
  auto c = Sphere::glass();
  c->set_transform(translation(0.0, 0.0, 0.25));
  kat::edit_material(*c, [](Material& m) { m.refractive_index = 2.5; });
  Ray r(Point(0, 0, -4), Vector(0, 0, 1));
  auto xs = intersections({Intersection(2.0, a.get()), Intersection(2.75, b.get()), Intersection(3.25, c.get()),
                           Intersection(4.75, b.get()), Intersection(5.25, c.get()), Intersection(6.0, a.get())});
  double expected[6][2] = {{1.0, 1.5}, {1.5, 2.0}, {2.0, 2.5}, {2.5, 2.5}, {2.5, 1.5}, {1.5, 1.0}};
  for (int i = 0; i < 6; i++) {
    Computations comps = prepare_computations(xs[(size_t)i], r, xs);
    CHECK_NEAR(comps.n1, expected[i][0]);
    CHECK_NEAR(comps.n2, expected[i][1]);
  }
}
// intersection.rs:338-390
KAT(intersection_under_point_and_schlick) {
  Ray r(Point(0, 0, -5), Vector(0, 0, 1));
  auto shape = Sphere::glass();
  shape->set_transform(translation(0, 0, 1));
  Intersection i(5.0, shape.get());
  Computations c = prepare_computations(i, r, {i});
  CHECK(c.under_point.z > EPSILON / 2.0); CHECK(c.point.z < c.under_point.z);

  auto g = Sphere::glass();
  Ray r1(Point(0.0, 0.0, S2 / 2.0), Vector(0, 1, 0));
  Intersections xs = {Intersection(-(S2 / 2.0), g.get()), Intersection(S2 / 2.0, g.get())};
  c = prepare_computations(xs[1], r1, xs);
  CHECK_NEAR(c.schlick(), 1.0);
  Ray r2(Point::origin(), Vector(0, 1, 0));
  xs = intersections({Intersection(-1.0, g.get()), Intersection(1.0, g.get())});
  c = prepare_computations(xs[1], r2, xs);
  CHECK_NEAR(c.schlick(), 0.04);
  Ray r3(Point(0.0, 0.99, -2.0), Vector(0, 0, 1));
  xs = intersections({Intersection(1.8589, g.get())});
  c = prepare_computations(xs[0], r3, xs);
  CHECK_NEAR(c.schlick(), 0.48873);
}
// intersection.rs:393-418
KAT(intersection_uv_and_shadow_hit) {
  Triangle s = tri();
  Intersection i(3.5, &s, 0.2, 0.4);
  CHECK_NEAR(i.u, 0.2); CHECK_NEAR(i.v, 0.4);
  Sphere s1;
  s1.no_shadow();
  Sphere s2;
  Intersection i3(1.0, &s2);
  auto xs = intersections({Intersection(1.0, &s1), Intersection(2.0, &s1), i3, Intersection(2.0, &s2)});
  const Intersection* h = shadow_hit(xs);
  CHECK(h != nullptr && h->object == &s2 && h->t == 1.0);
}
