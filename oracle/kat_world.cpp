// ORACLE KATs — world.rs:160-521 and camera.rs:260-338 restated.
#include "kat.hpp"
using namespace orc;

static const double PI = 3.14159265358979323846;
static const double S2 = std::sqrt(2.0);

// world.rs:163-168
KAT(world_create) {
  World w;
  CHECK_EQ(w.objects.size(), 0u);
  CHECK_EQ(w.lights.size(), 0u);
}
// world.rs:170-184
KAT(world_default) {
  PointLight light(Point(-10, 10, -10), Color(1, 1, 1));
  Sphere s1;
  kat::edit_material(s1, [](Material& m) { m.color = Color(0.8, 1.0, 0.6); m.diffuse = 0.7; m.specular = 0.2; });
  Sphere s2;
  s2.set_transform(scaling(0.5, 0.5, 0.5));
  World w = World::default_world();
  CHECK(w.lights[0] == light);
  CHECK(w.objects[0]->equals(s1));
  CHECK(w.objects[1]->equals(s2));
  CHECK(!w.objects[0]->equals(s2));
}
// world.rs:186-195
KAT(world_intersect_with_ray) {
  World w = World::default_world();
  Ray r(Point(0, 0, -5), Vector(0, 0, 1));
  Intersections xs = w.intersect(r);
  CHECK_EQ(xs.size(), 4u);
  CHECK_NEAR(xs[0].t, 4.0);
  CHECK_NEAR(xs[1].t, 4.5);
  CHECK_NEAR(xs[2].t, 5.5);
  CHECK_NEAR(xs[3].t, 6.0);
}
// world.rs:198-206
KAT(world_shade_intersection) {
  World w = World::default_world();
  Ray r(Point(0, 0, -5), Vector(0, 0, 1));
  Intersection i(4.0, w.objects[0].get());
  Computations comps = prepare_computations(i, r, {i});
  CHECK_EQ(w.shade_hit(comps, MAX_RECURSION_DEPTH), Color(0.38066, 0.47583, 0.2855));
}
// world.rs:209-218
KAT(world_shade_intersection_inside) {
  World w = World::default_world();
  w.lights[0] = PointLight(Point(0.0, 0.25, 0.0), Color(1, 1, 1));
  Ray r(Point(0, 0, 0), Vector(0, 0, 1));
  Intersection i(0.5, w.objects[1].get());
  Computations comps = prepare_computations(i, r, {i});
  CHECK_EQ(w.shade_hit(comps, MAX_RECURSION_DEPTH), Color(0.90498, 0.90498, 0.90498));
}
// world.rs:221-234
KAT(world_color_ray_miss_and_hit) {
  World w = World::default_world();
  CHECK_EQ(w.color_at(Ray(Point(0, 0, -5), Vector(0, 1, 0)), MAX_RECURSION_DEPTH), Color::black());
  CHECK_EQ(w.color_at(Ray(Point(0, 0, -5), Vector(0, 0, 1)), MAX_RECURSION_DEPTH), Color(0.38066, 0.47583, 0.2855));
}
// world.rs:249-274
KAT(world_is_shadowed_four_cases) {
  World w = World::default_world();
  CHECK_EQ(w.is_shadowed(Point(0, 10, 0), w.lights[0]), false);
  CHECK_EQ(w.is_shadowed(Point(10, -10, 10), w.lights[0]), true);
  CHECK_EQ(w.is_shadowed(Point(-20, 20, -20), w.lights[0]), false);
  CHECK_EQ(w.is_shadowed(Point(-2, 2, -2), w.lights[0]), false);
}
// world.rs:277-293
KAT(world_shade_hit_in_shadow) {
  World w;
  w.add_light(PointLight(Point(0, 0, -10), Color(1, 1, 1)));
  w.add_object(std::make_unique<Sphere>());
  auto s2 = std::make_unique<Sphere>();
  s2->set_transform(translation(0, 0, 10));
  w.add_object(std::move(s2));
  Ray r(Point(0, 0, 5), Vector(0, 0, 1));
  Intersection i(4.0, w.objects[1].get());
  Computations comps = prepare_computations(i, r, {i});
  CHECK_EQ(w.shade_hit(comps, MAX_RECURSION_DEPTH), Color(0.1, 0.1, 0.1));
}
// world.rs:296-311
KAT(world_reflected_color_nonreflective) {
  World w = World::default_world();
  Ray r(Point::origin(), Vector(0, 0, 1));
  auto shape = std::make_unique<Sphere>();
  shape->set_transform(scaling(0.5, 0.5, 0.5));
  kat::edit_material(*shape, [](Material& m) { m.ambient = 1.0; });
  w.objects[1] = std::move(shape);
  Intersection i(1.0, w.objects[1].get());
  Computations comps = prepare_computations(i, r, {i});
  CHECK_EQ(w.reflected_color(comps, MAX_RECURSION_DEPTH), Color::black());
}
static World world_with_reflective_plane() {
  World w = World::default_world();
  auto shape = std::make_unique<Plane>();
  kat::edit_material(*shape, [](Material& m) { m.reflective = 0.5; });
  shape->set_transform(translation(0, -1, 0));
  w.add_object(std::move(shape));
  return w;
}
// world.rs:314-329, :332-347, :369-383
KAT(world_reflected_color_reflective) {
  World w = world_with_reflective_plane();
  Ray r(Point(0, 0, -3), Vector(0.0, -(S2 / 2.0), S2 / 2.0));
  Intersection i(S2, w.objects[2].get());
  Computations comps = prepare_computations(i, r, {i});
  CHECK_EQ(w.reflected_color(comps, MAX_RECURSION_DEPTH), Color(0.19033, 0.23791, 0.14274));
  CHECK_EQ(w.shade_hit(comps, MAX_RECURSION_DEPTH), Color(0.87676, 0.92435, 0.82918));
  Intersection i0(S2, w.objects[0].get());
  Computations c0 = prepare_computations(i0, r, {i0});
  CHECK_EQ(w.reflected_color(c0, 0), Color::black());
}
// world.rs:350-366 — must terminate
KAT(world_mutually_reflective_terminates) {
  World w;
  w.add_light(PointLight(Point::origin(), Color::white()));
  auto lower = std::make_unique<Plane>();
  kat::edit_material(*lower, [](Material& m) { m.reflective = 1.0; });
  lower->set_transform(translation(0, -1, 0));
  w.add_object(std::move(lower));
  auto upper = std::make_unique<Plane>();
  kat::edit_material(*upper, [](Material& m) { m.reflective = 1.0; });
  upper->set_transform(translation(0, 1, 0));
  w.add_object(std::move(upper));
  counters() = Counters();
  w.color_at(Ray(Point::origin(), Vector(0, 1, 0)), MAX_RECURSION_DEPTH);
  CHECK_EQ(counters().rays_reflect, 5u);
}
// world.rs:386-397
KAT(world_refracted_color_opaque) {
  World w = World::default_world();
  Shape* shape = w.objects[0].get();
  Ray r(Point(0, 0, -5), Vector(0, 0, 1));
  Intersections xs = intersections({Intersection(4.0, shape), Intersection(6.0, shape)});
  Computations comps = prepare_computations(xs[0], r, xs);
  CHECK_EQ(w.refracted_color(comps, MAX_RECURSION_DEPTH), Color::black());
}
// world.rs:400-416
KAT(world_refracted_color_max_depth) {
  World w = World::default_world();
  kat::edit_material(*w.objects[0], [](Material& m) { m.transparency = 1.0; m.refractive_index = 1.5; });
  Shape* shape = w.objects[0].get();
  Ray r(Point(0, 0, 5), Vector(0, 0, 1));
  Intersections xs = intersections({Intersection(4.0, shape), Intersection(6.0, shape)});
  Computations comps = prepare_computations(xs[0], r, xs);
  CHECK_EQ(w.refracted_color(comps, 0), Color::black());
}
// world.rs:419-438
KAT(world_refracted_color_total_internal_reflection) {
  World w = World::default_world();
  kat::edit_material(*w.objects[0], [](Material& m) { m.transparency = 1.0; m.refractive_index = 1.5; });
  Shape* shape = w.objects[0].get();
  Ray r(Point(0.0, 0.0, S2 / 2.0), Vector(0, 1, 0));
  Intersections xs = intersections({Intersection(-(S2 / 2.0), shape), Intersection(S2 / 2.0, shape)});
  Computations comps = prepare_computations(xs[1], r, xs);
  CHECK_EQ(w.refracted_color(comps, MAX_RECURSION_DEPTH), Color::black());
}
// world.rs:441-463
KAT(world_refracted_color_with_refracted_ray) {
  World w = World::default_world();
  kat::edit_material(*w.objects[0], [](Material& m) { m.ambient = 1.0; m.set_pattern(test_pattern()); });
  kat::edit_material(*w.objects[1], [](Material& m) { m.transparency = 1.0; m.refractive_index = 1.5; });
  Shape* a = w.objects[0].get();
  Shape* b = w.objects[1].get();
  Ray r(Point(0.0, 0.0, 0.1), Vector(0, 1, 0));
  Intersections xs = intersections({Intersection(-0.9899, a), Intersection(-0.4899, b), Intersection(0.4899, b),
                                    Intersection(0.9899, a)});
  Computations comps = prepare_computations(xs[2], r, xs);
  CHECK_EQ(w.refracted_color(comps, MAX_RECURSION_DEPTH), Color(0.0, 0.99887, 0.04722));
}
static World world_with_glass_floor(double reflective) {
  World w = World::default_world();
  auto floor = std::make_unique<Plane>();
  floor->set_transform(translation(0, -1, 0));
  kat::edit_material(*floor, [&](Material& m) { m.reflective = reflective; m.transparency = 0.5; m.refractive_index = 1.5; });
  w.add_object(std::move(floor));
  auto ball = std::make_unique<Sphere>();
  kat::edit_material(*ball, [](Material& m) { m.color = Color(1.0, 0.0, 0.0); m.ambient = 0.5; });
  ball->set_transform(translation(0.0, -3.5, -0.5));
  w.add_object(std::move(ball));
  return w;
}
// world.rs:466-491
KAT(world_shade_hit_transparent) {
  World w = world_with_glass_floor(0.0);
  Ray r(Point(0, 0, -3), Vector(0.0, -(S2 / 2.0), S2 / 2.0));
  Intersections xs = intersections({Intersection(S2, w.objects[2].get())});
  Computations comps = prepare_computations(xs[0], r, xs);
  CHECK_EQ(w.shade_hit(comps, MAX_RECURSION_DEPTH), Color(0.93642, 0.68642, 0.68642));
}
// world.rs:494-520
KAT(world_shade_hit_reflective_transparent) {
  World w = world_with_glass_floor(0.5);
  Ray r(Point(0, 0, -3), Vector(0.0, -(S2 / 2.0), S2 / 2.0));
  Intersections xs = intersections({Intersection(S2, w.objects[2].get())});
  Computations comps = prepare_computations(xs[0], r, xs);
  CHECK_EQ(w.shade_hit(comps, MAX_RECURSION_DEPTH), Color(0.93391, 0.69643, 0.69243));
}

// camera.rs:274-297
KAT(camera_create_and_pixel_size) {
  Camera c(160, 120, PI / 2.0);
  CHECK_EQ(c.hsize, 160u);
  CHECK_EQ(c.vsize, 120u);
  CHECK(c.transform == Matrix::identity(4, 4));
  CHECK_NEAR(Camera(200, 125, PI / 2.0).pixel_size, 0.01);
  CHECK_NEAR(Camera(125, 200, PI / 2.0).pixel_size, 0.01);
}
// camera.rs:300-325
KAT(camera_rays) {
  Camera c(201, 101, PI / 2.0);
  Ray r = c.ray_for_pixel(100, 50);
  CHECK_EQ(r.origin, Point::origin());
  CHECK_EQ(r.direction, Vector(0, 0, -1));
  r = c.ray_for_pixel(0, 0);
  CHECK_EQ(r.origin, Point::origin());
  CHECK_EQ(r.direction, Vector(0.66519, 0.33259, -0.66851));
  c.set_transform(rotation_y(PI / 4.0) * translation(0, -2, 5));
  r = c.ray_for_pixel(100, 50);
  CHECK_EQ(r.origin, Point(0, 2, -5));
  CHECK_EQ(r.direction, Vector(S2 / 2.0, 0.0, -(S2 / 2.0)));
}
// camera.rs:328-337 — the reference's one end-to-end render test
KAT(camera_render_default_world) {
  World w = World::default_world();
  Camera c(11, 11, PI / 2.0);
  c.set_transform(view_transform(Point(0, 0, -5), Point::origin(), Vector(0, 1, 0)));
  Canvas image = c.render(w);
  CHECK_EQ(image.get_pixel(5, 5), Color(0.38066, 0.47583, 0.2855));
  // the threaded path with AA x1 gives the same picture
  Camera c2 = c;
  c2.render_opts.num_threads = 3;
  Canvas image2 = c2.render_multithreaded(w);
  for (size_t i = 0; i < image.pixels.size(); i++) {
    CHECK(image.pixels[i].red == image2.pixels[i].red);
    CHECK(image.pixels[i].green == image2.pixels[i].green);
    CHECK(image.pixels[i].blue == image2.pixels[i].blue);
  }
}
// camera.rs:92-126 + color.rs:26-33
KAT(camera_aa_offsets) {
  CHECK_EQ(Camera::get_offsets(AASamples::X1).size(), 1u);
  CHECK_EQ(Camera::get_offsets(AASamples::X2).size(), 2u);
  CHECK_EQ(Camera::get_offsets(AASamples::X4).size(), 4u);
  CHECK_EQ(Camera::get_offsets(AASamples::X8).size(), 8u);
  CHECK_EQ(Camera::get_offsets(AASamples::X16).size(), 16u);
  auto o = Camera::get_offsets(AASamples::X16);
  CHECK(o[5].first == 0.375 && o[5].second == 0.375);
  CHECK(o[15].first == 0.875 && o[15].second == 0.875);
}
