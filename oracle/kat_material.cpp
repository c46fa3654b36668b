// ORACLE KATs — material.rs:88-256, pattern/{mod,stripe,gradient,ring,checkers,test_pattern}.rs and image/ppm.rs:76-150
// test modules restated. These pin the rows the headline scene exercises hardest (stripes + checkers + Phong).
#include "kat.hpp"
using namespace orc;

static const double S2 = std::sqrt(2.0);
static Color lit(const Material& m, Point light_pos, Vector eyev, bool shadow, Point pos = Point::origin()) {
  Sphere s;
  PointLight light(light_pos, Color(1, 1, 1));
  return m.lighting(s, light, pos, eyev, Vector(0, 0, -1), shadow);
}

// material.rs:94-102, :244-256
KAT(material_default) {
  Material m;
  CHECK(m.color == Color(1, 1, 1));
  CHECK_NEAR(m.ambient, 0.1); CHECK_NEAR(m.diffuse, 0.9); CHECK_NEAR(m.specular, 0.9); CHECK_NEAR(m.shininess, 200.0);
  CHECK_NEAR(m.reflective, 0.0); CHECK_NEAR(m.transparency, 0.0); CHECK_NEAR(m.refractive_index, 1.0);
}
// material.rs:104-120
KAT(material_lighting_eye_between_light_and_surface) {
  CHECK(lit(Material(), Point(0, 0, -10), Vector(0, 0, -1), false) == Color(1.9, 1.9, 1.9));
}
// material.rs:122-138
KAT(material_lighting_eye_offset_45deg) {
  CHECK(lit(Material(), Point(0, 0, -10), Vector(0.0, S2 / 2.0, -(S2 / 2.0)), false) == Color(1.0, 1.0, 1.0));
}
// material.rs:140-156
KAT(material_lighting_light_offset_45deg) {
  CHECK(lit(Material(), Point(0, 10, -10), Vector(0, 0, -1), false) == Color(0.7364, 0.7364, 0.7364));
}
// material.rs:158-174
KAT(material_lighting_eye_in_reflection_path) {
  CHECK(lit(Material(), Point(0, 10, -10), Vector(0.0, -(S2 / 2.0), -(S2 / 2.0)), false) == Color(1.6364, 1.6364, 1.6364));
}
// material.rs:176-192
KAT(material_lighting_light_behind_surface) {
  CHECK(lit(Material(), Point(0, 0, 10), Vector(0, 0, -1), false) == Color(0.1, 0.1, 0.1));
}
// material.rs:194-211
KAT(material_lighting_surface_in_shadow) {
  CHECK(lit(Material(), Point(0, 0, -10), Vector(0, 0, -1), true) == Color(0.1, 0.1, 0.1));
}
// material.rs:213-242
KAT(material_lighting_with_pattern) {
  Material m;
  m.set_pattern(stripe_pattern(Color::white(), Color::black()));
  m.ambient = 1.0; m.diffuse = 0.0; m.specular = 0.0;
  CHECK(lit(m, Point(0, 0, -10), Vector(0, 0, -1), false, Point(0.9, 0, 0)) == Color::white());
  CHECK(lit(m, Point(0, 0, -10), Vector(0, 0, -1), false, Point(1.1, 0, 0)) == Color::black());
}

// pattern/stripe.rs:33-76
KAT(pattern_stripe_color_at) {
  Color white = Color::white(), black = Color::black();
  Pattern p = stripe_pattern(white, black);
  CHECK(p.a == white); CHECK(p.b == black);
  CHECK(p.color_at(Point(0, 0, 0)) == white); CHECK(p.color_at(Point(0, 1, 0)) == white); CHECK(p.color_at(Point(0, 2, 0)) == white);
  CHECK(p.color_at(Point(0, 0, 1)) == white); CHECK(p.color_at(Point(0, 0, 2)) == white);
  CHECK(p.color_at(Point(0.9, 0, 0)) == white); CHECK(p.color_at(Point(1, 0, 0)) == black);
  CHECK(p.color_at(Point(-0.1, 0, 0)) == black); CHECK(p.color_at(Point(-1, 0, 0)) == black);
  CHECK(p.color_at(Point(-1.1, 0, 0)) == white);
}
// pattern/stripe.rs:78-117
KAT(pattern_stripe_transforms) {
  Color white = Color::white(), black = Color::black();
  Sphere scaled;
  scaled.set_transform(scaling(2, 2, 2));
  Sphere plain;
  Pattern p = stripe_pattern(white, black);
  CHECK(p.color_at_shape(scaled, Point(1.5, 0, 0)) == white);
  Pattern q = stripe_pattern(white, black);
  q.set_transform(scaling(2, 2, 2));
  CHECK(q.color_at_shape(plain, Point(1.5, 0, 0)) == white);
  Pattern r = stripe_pattern(white, black);
  r.set_transform(translation(0.5, 0, 0));
  CHECK(r.color_at_shape(scaled, Point(2.5, 0, 0)) == white);
}
// pattern/gradient.rs:29-46
KAT(pattern_gradient_interpolates) {
  Pattern p = gradient_pattern(Color::white(), Color::black());
  CHECK(p.color_at(Point::origin()) == Color::white());
  CHECK(p.color_at(Point(0.25, 0, 0)) == Color(0.75, 0.75, 0.75));
  CHECK(p.color_at(Point(0.5, 0, 0)) == Color(0.5, 0.5, 0.5));
  CHECK(p.color_at(Point(0.75, 0, 0)) == Color(0.25, 0.25, 0.25));
}
// pattern/ring.rs:32-40
KAT(pattern_ring_extends_in_x_and_z) {
  Pattern p = ring_pattern(Color::white(), Color::black());
  CHECK(p.color_at(Point::origin()) == Color::white());
  CHECK(p.color_at(Point(1, 0, 0)) == Color::black());
  CHECK(p.color_at(Point(0, 0, 1)) == Color::black());
  CHECK(p.color_at(Point(0.708, 0, 0.708)) == Color::black());
}
// pattern/checkers.rs:32-59
KAT(pattern_checkers_repeat) {
  Pattern p = checkers_pattern(Color::white(), Color::black());
  CHECK(p.color_at(Point::origin()) == Color::white());
  CHECK(p.color_at(Point(0.99, 0, 0)) == Color::white()); CHECK(p.color_at(Point(1.01, 0, 0)) == Color::black());
  CHECK(p.color_at(Point(0, 0.99, 0)) == Color::white()); CHECK(p.color_at(Point(0, 1.01, 0)) == Color::black());
  CHECK(p.color_at(Point(0, 0, 0.99)) == Color::white()); CHECK(p.color_at(Point(0, 0, 1.01)) == Color::black());
}
// pattern/test_pattern.rs:24-68
KAT(pattern_test_pattern_transforms) {
  Pattern d = test_pattern();
  CHECK(d.transform == Matrix::identity(4, 4)); CHECK(d.transform_inverse == Matrix::identity(4, 4));
  Pattern t = test_pattern();
  t.set_transform(translation(1, 2, 3));
  CHECK(t.transform == translation(1, 2, 3)); CHECK(t.transform_inverse == translation(1, 2, 3).inverse());
  Sphere scaled;
  scaled.set_transform(scaling(2, 2, 2));
  Sphere plain;
  CHECK(test_pattern().color_at_shape(scaled, Point(2, 3, 4)) == Color(1.0, 1.5, 2.0));
  Pattern s = test_pattern();
  s.set_transform(scaling(2, 2, 2));
  CHECK(s.color_at_shape(plain, Point(2, 3, 4)) == Color(1.0, 1.5, 2.0));
  Pattern u = test_pattern();
  u.set_transform(translation(0.5, 1.0, 1.5));
  CHECK(u.color_at_shape(scaled, Point(2.5, 3.0, 3.5)) == Color(0.75, 0.5, 0.25));
}

static std::vector<std::string> lines_of(const std::string& s) {
  std::vector<std::string> out;
  std::istringstream in(s);
  std::string l;
  while (std::getline(in, l)) out.push_back(l);
  return out;
}
// image/ppm.rs:80-87
KAT(ppm_header) {
  auto l = lines_of(canvas_to_ppm(Canvas(5, 3)));
  CHECK(l.size() >= 3);
  CHECK_EQ(l[0], std::string("P3")); CHECK_EQ(l[1], std::string("5 3")); CHECK_EQ(l[2], std::string("255"));
}
// image/ppm.rs:89-109
KAT(ppm_pixel_data) {
  Canvas c(5, 3);
  c.set_pixel(0, 0, Color(1.5, 0, 0));
  c.set_pixel(2, 1, Color(0, 0.5, 0));
  c.set_pixel(4, 2, Color(-0.5, 0, 1));
  auto l = lines_of(canvas_to_ppm(c));
  CHECK_EQ(l.size(), 6u);
  CHECK_EQ(l[3], std::string("255 0 0 0 0 0 0 0 0 0 0 0 0 0 0"));
  CHECK_EQ(l[4], std::string("0 0 0 0 0 0 0 128 0 0 0 0 0 0 0"));
  CHECK_EQ(l[5], std::string("0 0 0 0 0 0 0 0 0 0 0 0 0 0 255"));
}
// image/ppm.rs:111-125
KAT(ppm_color_component_scaling) {
  CHECK_EQ((int)scale_color_component(0.0), 0); CHECK_EQ((int)scale_color_component(255.0), 255);
  CHECK_EQ((int)scale_color_component(-0.5), 0); CHECK_EQ((int)scale_color_component(1.5), 255);
  CHECK_EQ((int)scale_color_component(0.5), 128);
}
// image/ppm.rs:127-150
KAT(ppm_split_long_lines) {
  Canvas c(10, 2);
  for (size_t j = 0; j < c.height; j++)
    for (size_t i = 0; i < c.width; i++) c.set_pixel(i, j, Color(1.0, 0.8, 0.6));
  auto l = lines_of(canvas_to_ppm(c));
  CHECK_EQ(l.size(), 7u);
  const std::string a = "255 204 153 255 204 153 255 204 153 255 204 153 255 204 153 255 204";
  const std::string b = "153 255 204 153 255 204 153 255 204 153 255 204 153";
  CHECK_EQ(l[3], a); CHECK_EQ(l[4], b); CHECK_EQ(l[5], a); CHECK_EQ(l[6], b);
}
