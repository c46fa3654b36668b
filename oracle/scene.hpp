// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement of the reference's scene graph and ray/shape arithmetic, kept in the
// reference's own shape (trait objects → virtual classes, Vec<Intersection> → std::vector,
// collect-everything-then-sort), quirks included. Paths relative to /root/reference/raytracer/src.
//
//   bounding_box.rs:5-169                BoundingBox
//   light.rs:4-24                        PointLight
//   pattern/mod.rs:17-91 + kinds         Pattern (test/stripe/gradient/ring/checkers)
//   material.rs:11-87                    Material, lighting (Phong)
//   geometry/mod.rs:13-114               BaseShape, trait Shape (intersect, normal_at, set_transform …)
//   geometry/intersection.rs:11-162      Intersection, prepare_computations, intersections, hit,
//                                        shadow_hit, Computations::schlick
//   geometry/shape/{sphere,plane,cube,cylinder,cone,triangle,smooth_triangle,group,csg,test_shape}.rs
#pragma once
#include <algorithm>
#include <memory>
#include <string>

#include "math.hpp"

namespace orc {

// ---------------------------------------------------------------------------------------------
// Per-thread call counters: the exact, scene-specific "algorithmic work" figures of SURVEY §8-d.
// ---------------------------------------------------------------------------------------------
struct Counters {
  uint64_t rays_primary = 0, rays_shadow = 0, rays_reflect = 0, rays_refract = 0;
  uint64_t ray_gen = 0, ray_xform = 0;
  uint64_t sphere_miss = 0, sphere_hit = 0, plane = 0, cube = 0, aabb = 0;
  uint64_t cyl = 0, cone = 0, caps = 0;
  uint64_t tri_exit0 = 0, tri_exit1 = 0, tri_exit2 = 0, tri_hit = 0;
  uint64_t normal_at = 0, prepare = 0, lighting = 0, lighting_pattern = 0;
  uint64_t shadow_setup = 0, refract = 0, schlick = 0, shade = 0, pixels = 0;
  uint64_t sorted_elems = 0, nodes_visited = 0;
  void add(const Counters& o) {
    const uint64_t* s = reinterpret_cast<const uint64_t*>(&o);
    uint64_t* d = reinterpret_cast<uint64_t*>(this);
    for (size_t i = 0; i < sizeof(Counters) / sizeof(uint64_t); i++) d[i] += s[i];
  }
  // flops per call: SURVEY §8-a constants.
  double flops() const {
    return 54.0 * ray_gen + 33.0 * ray_xform + 24.0 * sphere_miss + 32.0 * sphere_hit + 1.0 * plane +
           16.0 * cube + 16.0 * aabb + 28.0 * cyl + 34.0 * cone + 10.0 * caps + 14.0 * tri_exit0 +
           24.0 * tri_exit1 + 40.0 * tri_exit2 + 46.0 * tri_hit + 45.0 * normal_at + 37.0 * prepare +
           65.0 * lighting + 40.0 * lighting_pattern + 18.0 * shadow_setup + 25.0 * refract +
           20.0 * schlick + 9.0 * shade;
  }
};
inline Counters& counters() {
  static thread_local Counters c;
  return c;
}

// ---------------------------------------------------------------------------------------------
// bounding_box.rs
// ---------------------------------------------------------------------------------------------
struct BoundingBox {
  Point min{INF, INF, INF};      // bounding_box.rs:11-18 (Default = empty, inverted infinite box)
  Point max{-INF, -INF, -INF};
  BoundingBox() = default;
  BoundingBox(Point mn, Point mx) : min(mn), max(mx) {}

  bool operator==(const BoundingBox& o) const {  // bounding_box.rs:20-24
    return min.eq_ignore_inf(o.min) && max.eq_ignore_inf(o.max);
  }
  void add_point(Point p) {  // bounding_box.rs:39-59
    if (p.x > max.x) max.x = p.x;
    if (p.y > max.y) max.y = p.y;
    if (p.z > max.z) max.z = p.z;
    if (p.x < min.x) min.x = p.x;
    if (p.y < min.y) min.y = p.y;
    if (p.z < min.z) min.z = p.z;
  }
  void add_bounding_box(const BoundingBox& b) {  // bounding_box.rs:61-64
    add_point(b.min);
    add_point(b.max);
  }
  bool contains_point(Point p) const {  // bounding_box.rs:66-70 (inclusive ranges)
    return (min.x <= p.x && p.x <= max.x) && (min.y <= p.y && p.y <= max.y) && (min.z <= p.z && p.z <= max.z);
  }
  bool contains_bounding_box(const BoundingBox& b) const {  // bounding_box.rs:72-74
    return contains_point(b.min) && contains_point(b.max);
  }
  BoundingBox transform(const Matrix& m) const {  // bounding_box.rs:76-93
    Point ps[8] = {min,
                   Point(min.x, min.y, max.z),
                   Point(min.x, max.y, min.z),
                   Point(min.x, max.y, max.z),
                   Point(max.x, min.y, min.z),
                   Point(max.x, min.y, max.z),
                   Point(max.x, max.y, min.z),
                   max};
    BoundingBox nb;
    for (const Point& p : ps) nb.add_point(m * p);
    return nb;
  }
  static void check_axis(double origin, double direction, double mn, double mx, double& tmin, double& tmax) {
    // bounding_box.rs:118-136
    double tmin_numerator = mn - origin;
    double tmax_numerator = mx - origin;
    if (std::fabs(direction) >= EPSILON) {
      tmin = tmin_numerator / direction;
      tmax = tmax_numerator / direction;
    } else {
      tmin = tmin_numerator * INF;
      tmax = tmax_numerator * INF;
    }
    if (tmin > tmax) std::swap(tmin, tmax);
  }
  bool intersects(const Ray& r) const {  // bounding_box.rs:95-116 — no t-range: tmin <= tmax only
    counters().aabb++;
    double xtmin, xtmax, ytmin, ytmax, ztmin, ztmax;
    check_axis(r.origin.x, r.direction.x, min.x, max.x, xtmin, xtmax);
    check_axis(r.origin.y, r.direction.y, min.y, max.y, ytmin, ytmax);
    check_axis(r.origin.z, r.direction.z, min.z, max.z, ztmin, ztmax);
    double tmin = rmax(rmax(xtmin, ytmin), ztmin);
    double tmax = rmin(rmin(xtmax, ytmax), ztmax);
    return tmin <= tmax;
  }
  std::pair<BoundingBox, BoundingBox> split() const {  // bounding_box.rs:138-169
    double dx = std::fabs(max.x - min.x);
    double dy = std::fabs(max.y - min.y);
    double dz = std::fabs(max.z - min.z);
    double greatest = rmax(rmax(dx, dy), dz);
    double x0 = min.x, y0 = min.y, z0 = min.z;
    double x1 = max.x, y1 = max.y, z1 = max.z;
    if (equal(greatest, dx)) {
      double midx = x0 + dx / 2.0;
      x0 = midx; x1 = midx;
    } else if (equal(greatest, dy)) {
      double midy = y0 + dy / 2.0;
      y0 = midy; y1 = midy;
    } else {
      double midz = z0 + dz / 2.0;
      z0 = midz; z1 = midz;
    }
    Point mid_min(x0, y0, z0), mid_max(x1, y1, z1);
    return {BoundingBox(min, mid_max), BoundingBox(mid_min, max)};
  }
};

// ---------------------------------------------------------------------------------------------
// light.rs
// ---------------------------------------------------------------------------------------------
struct PointLight {
  Color intensity;
  Point position;
  PointLight(Point p, Color i) : intensity(i), position(p) {}
  bool operator==(const PointLight& o) const { return intensity == o.intensity && position == o.position; }
};

// ---------------------------------------------------------------------------------------------
// pattern/*.rs
// ---------------------------------------------------------------------------------------------
enum class PatternKind : int { Test = 0, Stripe = 1, Gradient = 2, Ring = 3, Checkers = 4 };

struct Shape;

struct Pattern {
  Matrix transform = Matrix::identity(4, 4);
  Matrix transform_inverse = Matrix::identity(4, 4);
  PatternKind kind = PatternKind::Test;  // pattern/mod.rs:23-31: Default is the test pattern
  Color a, b;

  void set_transform(const Matrix& t) {  // pattern/mod.rs:34-37
    transform = t;
    transform_inverse = transform.inverse();
  }
  // Rust `as isize` — saturating, NaN → 0
  static long long as_isize(double v) {
    if (std::isnan(v)) return 0;
    if (v >= 9223372036854775807.0) return std::numeric_limits<long long>::max();
    if (v <= -9223372036854775808.0) return std::numeric_limits<long long>::min();
    return (long long)v;
  }
  Color color_at(Point p) const {
    switch (kind) {
      case PatternKind::Test:  // test_pattern.rs:7-9
        return Color(p.x, p.y, p.z);
      case PatternKind::Stripe:  // stripe.rs:14-20 — Rust `%` on f64 is fmod
        return std::fmod(std::floor(p.x), 2.0) == 0.0 ? a : b;
      case PatternKind::Gradient: {  // gradient.rs:14-18
        Color distance = b - a;
        double fraction = p.x - std::floor(p.x);
        return a + distance * fraction;
      }
      case PatternKind::Ring: {  // ring.rs:14-21
        double distance = std::floor(std::sqrt(p.x * p.x + p.z * p.z));
        return std::fmod(distance, 2.0) == 0.0 ? a : b;
      }
      case PatternKind::Checkers: {  // checkers.rs:14-21
        double distance = std::floor(p.x) + std::floor(p.y) + std::floor(p.z);
        return as_isize(distance) % 2 == 0 ? a : b;
      }
    }
    return Color::black();
  }
  Color color_at_shape(const Shape& shape, Point world_point) const;  // pattern/mod.rs:39-49
  bool operator==(const Pattern& o) const {  // derived PartialEq
    if (!(transform == o.transform) || !(transform_inverse == o.transform_inverse)) return false;
    if (kind != o.kind) return false;
    if (kind == PatternKind::Test) return true;
    return a == o.a && b == o.b;
  }
};
inline Pattern test_pattern() { return Pattern(); }
inline Pattern make_pattern(PatternKind k, Color a, Color b) {
  Pattern p;
  p.kind = k; p.a = a; p.b = b;
  return p;
}
inline Pattern stripe_pattern(Color a, Color b) { return make_pattern(PatternKind::Stripe, a, b); }
inline Pattern gradient_pattern(Color a, Color b) { return make_pattern(PatternKind::Gradient, a, b); }
inline Pattern ring_pattern(Color a, Color b) { return make_pattern(PatternKind::Ring, a, b); }
inline Pattern checkers_pattern(Color a, Color b) { return make_pattern(PatternKind::Checkers, a, b); }

// ---------------------------------------------------------------------------------------------
// material.rs
// ---------------------------------------------------------------------------------------------
struct Material {
  Color color{1.0, 1.0, 1.0};  // material.rs:24-36
  double ambient = 0.1, diffuse = 0.9, specular = 0.9, shininess = 200.0;
  double reflective = 0.0, transparency = 0.0, refractive_index = 1.0;
  bool has_pattern = false;
  Pattern pattern;

  void set_pattern(const Pattern& p) { has_pattern = true; pattern = p; }  // material.rs:84-86
  bool operator==(const Material& o) const {  // derived PartialEq: Color via epsilon, f64 exact
    return color == o.color && ambient == o.ambient && diffuse == o.diffuse && specular == o.specular &&
           shininess == o.shininess && reflective == o.reflective && transparency == o.transparency &&
           refractive_index == o.refractive_index && has_pattern == o.has_pattern &&
           (!has_pattern || pattern == o.pattern);
  }
  Color lighting(const Shape& object, const PointLight& light, Point point, Vector eyev, Vector normalv,
                 bool in_shadow) const;  // material.rs:38-82
};

// ---------------------------------------------------------------------------------------------
// geometry/intersection.rs:11-50
// ---------------------------------------------------------------------------------------------
struct Intersection {
  double t = 0;
  const Shape* object = nullptr;
  bool has_uv = false;
  double u = 0, v = 0;
  Intersection() = default;
  Intersection(double t_, const Shape* o) : t(t_), object(o) {}
  Intersection(double t_, const Shape* o, double u_, double v_) : t(t_), object(o), has_uv(true), u(u_), v(v_) {}
  bool operator==(const Intersection& o) const;  // derived PartialEq (t exact, object by value, u, v)
};
using Intersections = std::vector<Intersection>;

struct Computations {  // intersection.rs:132-144
  const Shape* object = nullptr;
  double t = 0;
  Point point, over_point, under_point;
  Vector eyev, normalv;
  bool inside = false;
  Vector reflectv;
  double n1 = -1.0, n2 = -1.0;
  double schlick() const;  // intersection.rs:147-162
};

// ---------------------------------------------------------------------------------------------
// geometry/mod.rs — BaseShape and trait Shape
// ---------------------------------------------------------------------------------------------
enum class ShapeKind : int {
  Sphere = 0, Plane = 1, Cube = 2, Cylinder = 3, Cone = 4, Triangle = 5, SmoothTriangle = 6,
  Group = 7, Csg = 8, TestShape = 9
};

// transform + inverse + inverse-transpose, shared between shapes whose matrices are bit-identical
// (purely a memory/time saving for 1M-triangle meshes; values are exactly those of
// geometry/mod.rs:74-85 recomputed per shape).
struct Xform {
  Matrix transform = Matrix::identity(4, 4);
  Matrix inverse = Matrix::identity(4, 4);
  Matrix inverse_transpose = Matrix::identity(4, 4);
};
using XformP = std::shared_ptr<const Xform>;
inline XformP identity_xform() {
  static XformP id = std::make_shared<const Xform>();
  return id;
}
// One-entry memo: consecutive set_transform calls with bit-identical matrices (all triangles of a
// mesh) share the computed inverse.
inline XformP make_xform(const Matrix& t) {
  static thread_local XformP last;
  if (last && last->transform.bits_equal(t)) return last;
  auto x = std::make_shared<Xform>();
  x->transform = t;
  x->inverse = t.inverse();
  x->inverse_transpose = x->inverse.transpose();
  last = x;
  return last;
}

struct BaseShape {  // geometry/mod.rs:13-36
  XformP xf = identity_xform();
  std::shared_ptr<const Material> material_p;
  BoundingBox bounding_box;
  bool shadow = true;
  const Matrix& transform() const { return xf->transform; }
  const Matrix& transform_inverse() const { return xf->inverse; }
  const Matrix& transform_inverse_transpose() const { return xf->inverse_transpose; }
  const Material& material() const {
    static const Material dflt;
    return material_p ? *material_p : dflt;
  }
  bool operator==(const BaseShape& o) const {  // derived PartialEq
    return transform() == o.transform() && transform_inverse() == o.transform_inverse() &&
           transform_inverse_transpose() == o.transform_inverse_transpose() && material() == o.material() &&
           bounding_box == o.bounding_box && shadow == o.shadow;
  }
};

struct Shape {
  BaseShape base;
  virtual ~Shape() = default;
  virtual ShapeKind kind() const = 0;
  virtual void local_intersect(const Ray& ray, Intersections& out) const = 0;
  virtual Vector local_normal_at(Point p, const Intersection& i) const = 0;
  virtual bool equals(const Shape& other) const = 0;

  // geometry/mod.rs:46-49
  virtual void intersect(const Ray& ray, Intersections& out) const {
    counters().ray_xform++;
    Ray local_ray = ray.transform(base.transform_inverse());
    local_intersect(local_ray, out);
  }
  // geometry/mod.rs:51-56
  Vector normal_at(Point point, const Intersection& i) const {
    counters().normal_at++;
    Point local_point = base.transform_inverse() * point;
    Vector local_normal = local_normal_at(local_point, i);
    Vector world_normal = base.transform_inverse_transpose() * local_normal;
    return world_normal.normalize();
  }
  const Material& material() const { return base.material(); }
  virtual void set_material(const Material& m) { base.material_p = std::make_shared<const Material>(m); }
  virtual void set_material_shared(std::shared_ptr<const Material> m) { base.material_p = std::move(m); }
  const Matrix& transform() const { return base.transform(); }
  // geometry/mod.rs:74-85 — un-transform the stored AABB by the OLD inverse, then re-transform by
  // the new matrix (boxes of rotated shapes inflate on every call; infinite boxes become empty).
  virtual void set_transform(const Matrix& t) {
    base.bounding_box = base.bounding_box.transform(base.transform_inverse());
    base.xf = make_xform(t);
    base.bounding_box = base.bounding_box.transform(base.transform());
  }
  // geometry/mod.rs:87-89 — pointer identity
  virtual bool includes(const Shape& other) const { return &base == &other.base; }
  const BoundingBox& get_bounds() const { return base.bounding_box; }
  BoundingBox parent_space_bounds() const { return base.bounding_box.transform(Matrix::identity(4, 4)); }
  virtual void divide(size_t) {}
  bool has_shadow() const { return base.shadow; }
  void no_shadow() { base.shadow = false; }
};
using ShapeP = std::unique_ptr<Shape>;

inline bool Intersection::operator==(const Intersection& o) const {
  if (!(t == o.t)) return false;
  if (!object->equals(*o.object)) return false;
  if (has_uv != o.has_uv) return false;
  if (has_uv && !(u == o.u && v == o.v)) return false;
  return true;
}

inline Color Pattern::color_at_shape(const Shape& shape, Point world_point) const {
  Point object_point = shape.base.transform_inverse() * world_point;
  Point pattern_point = transform_inverse * object_point;
  return color_at(pattern_point);
}

inline Color Material::lighting(const Shape& object, const PointLight& light, Point point, Vector eyev,
                                Vector normalv, bool in_shadow) const {
  counters().lighting++;
  Color c;
  if (has_pattern) {
    counters().lighting_pattern++;
    c = pattern.color_at_shape(object, point);
  } else {
    c = color;
  }
  Color effective_color = c * light.intensity;
  Vector lightv = (light.position - point).normalize();
  Color ambient_c = effective_color * ambient;
  if (in_shadow) return ambient_c;
  double light_dot_normal = dot(lightv, normalv);
  Color diffuse_c, specular_c;
  if (light_dot_normal < 0.0) {
    diffuse_c = Color::black();
    specular_c = Color::black();
  } else {
    diffuse_c = effective_color * diffuse * light_dot_normal;
    Vector reflectv = reflect(-lightv, normalv);
    double reflect_dot_eye = dot(reflectv, eyev);
    if (reflect_dot_eye <= 0.0) {
      specular_c = Color::black();
    } else {
      double factor = std::pow(reflect_dot_eye, shininess);  // powf → libm pow
      specular_c = light.intensity * specular * factor;
    }
  }
  return ambient_c + diffuse_c + specular_c;
}

// intersection.rs:147-162. powi(2) = x*x; powi(5) = x*((x*x)*(x*x)) (LLVM powi expansion).
inline double Computations::schlick() const {
  counters().schlick++;
  double c = dot(eyev, normalv);
  if (n1 > n2) {
    double n = n1 / n2;
    double sin2_t = n * n * (1.0 - c * c);
    if (sin2_t > 1.0) return 1.0;
    double cos_t = std::sqrt(1.0 - sin2_t);
    c = cos_t;
  }
  double q = (n1 - n2) / (n1 + n2);
  double r0 = q * q;
  double x = 1.0 - c;
  double x2 = x * x;
  double x4 = x2 * x2;
  return r0 + (1.0 - r0) * (x * x4);
}

// intersection.rs:108-116: clone + sort by t. sort_unstable_by leaves the order of equal t
// unspecified in the reference ("parity unpinned"); the oracle DEFINES it as stable (traversal
// order), which is also what Rust's insertion sort does for the <=20-element lists of most scenes.
// A NaN t makes the reference panic (partial_cmp().unwrap()); the oracle throws.
inline Intersections intersections(const Intersections& xs) {
  Intersections v(xs);
  for (const Intersection& i : v)
    if (std::isnan(i.t)) throw std::runtime_error("intersections: NaN t (reference panics here)");
  std::stable_sort(v.begin(), v.end(), [](const Intersection& a, const Intersection& b) { return a.t < b.t; });
  counters().sorted_elems += v.size();
  return v;
}
// intersection.rs:118-120
inline const Intersection* hit(const Intersections& xs) {
  for (const Intersection& i : xs)
    if (i.t >= 0.0) return &i;
  return nullptr;
}
// intersection.rs:122-125
inline const Intersection* shadow_hit(const Intersections& xs) {
  for (const Intersection& i : xs)
    if (i.t >= 0.0 && i.object->has_shadow()) return &i;
  return nullptr;
}

// intersection.rs:53-105
inline Computations prepare_computations(const Intersection& self, const Ray& ray, const Intersections& xs) {
  counters().prepare++;
  Point point = ray.position(self.t);
  Vector eyev = -ray.direction;
  Vector normalv = self.object->normal_at(point, self);
  bool inside = false;
  if (dot(normalv, eyev) < 0.0) {
    inside = true;
    normalv = -normalv;
  }
  std::vector<const Shape*> containers;
  double n1 = -1.0, n2 = -1.0;
  for (const Intersection& i : xs) {
    bool is_self = (i == self);
    if (is_self) {
      n1 = containers.empty() ? 1.0 : containers.back()->material().refractive_index;
    }
    // containers.contains(&i.object) / position(): VALUE equality (dyn Shape PartialEq → equals)
    size_t idx = containers.size();
    for (size_t k = 0; k < containers.size(); k++)
      if (containers[k]->equals(*i.object)) { idx = k; break; }
    if (idx < containers.size()) containers.erase(containers.begin() + (long)idx);
    else containers.push_back(i.object);
    if (is_self) {
      n2 = containers.empty() ? 1.0 : containers.back()->material().refractive_index;
      break;
    }
  }
  Computations c;
  c.object = self.object;
  c.t = self.t;
  c.point = point;
  c.over_point = point + normalv * EPSILON;
  c.under_point = point - normalv * EPSILON;
  c.eyev = eyev;
  c.normalv = normalv;
  c.inside = inside;
  c.reflectv = reflect(ray.direction, normalv);
  c.n1 = n1;
  c.n2 = n2;
  return c;
}

// ---------------------------------------------------------------------------------------------
// Leaf shapes
// ---------------------------------------------------------------------------------------------
template <class T>
inline const T* downcast(const Shape& s, ShapeKind k) { return s.kind() == k ? static_cast<const T*>(&s) : nullptr; }

struct Sphere : Shape {  // sphere.rs
  Sphere() { base.bounding_box = BoundingBox(Point(-1, -1, -1), Point(1, 1, 1)); }
  static std::unique_ptr<Sphere> glass() {  // sphere.rs:70-76
    auto s = std::make_unique<Sphere>();
    Material m;
    m.transparency = 1.0;
    m.refractive_index = 1.5;
    s->set_material(m);
    return s;
  }
  ShapeKind kind() const override { return ShapeKind::Sphere; }
  bool equals(const Shape& o) const override { return o.kind() == ShapeKind::Sphere && base == o.base; }
  void local_intersect(const Ray& ray, Intersections& out) const override {  // sphere.rs:47-62
    Vector sphere_to_ray = ray.origin - Point::origin();
    double a = dot(ray.direction, ray.direction);
    double b = 2.0 * dot(ray.direction, sphere_to_ray);
    double c = dot(sphere_to_ray, sphere_to_ray) - 1.0;
    double discriminant = b * b - 4.0 * a * c;
    if (discriminant < 0.0) {
      counters().sphere_miss++;
      return;
    }
    counters().sphere_hit++;
    double t1 = (-b - std::sqrt(discriminant)) / (2.0 * a);
    double t2 = (-b + std::sqrt(discriminant)) / (2.0 * a);
    out.emplace_back(t1, this);
    out.emplace_back(t2, this);
  }
  Vector local_normal_at(Point p, const Intersection&) const override { return p - Point::origin(); }
};

struct Plane : Shape {  // plane.rs
  Plane() { base.bounding_box = BoundingBox(Point(-INF, 0.0, -INF), Point(INF, 0.0, INF)); }
  ShapeKind kind() const override { return ShapeKind::Plane; }
  bool equals(const Shape& o) const override { return o.kind() == ShapeKind::Plane && base == o.base; }
  void local_intersect(const Ray& ray, Intersections& out) const override {  // plane.rs:53-60
    counters().plane++;
    if (std::fabs(ray.direction.y) < EPSILON) return;
    double t = -ray.origin.y / ray.direction.y;
    out.emplace_back(t, this);
  }
  Vector local_normal_at(Point, const Intersection&) const override { return Vector(0, 1, 0); }
};

struct Cube : Shape {  // cube.rs
  Cube() { base.bounding_box = BoundingBox(Point(-1, -1, -1), Point(1, 1, 1)); }
  ShapeKind kind() const override { return ShapeKind::Cube; }
  bool equals(const Shape& o) const override { return o.kind() == ShapeKind::Cube && base == o.base; }
  static void check_axis(double origin, double direction, double& tmin, double& tmax) {  // cube.rs:30-48
    double tmin_numerator = -1.0 - origin;
    double tmax_numerator = 1.0 - origin;
    if (std::fabs(direction) >= EPSILON) {
      tmin = tmin_numerator / direction;
      tmax = tmax_numerator / direction;
    } else {
      tmin = tmin_numerator * INF;
      tmax = tmax_numerator * INF;
    }
    if (tmin > tmax) std::swap(tmin, tmax);
  }
  void local_intersect(const Ray& ray, Intersections& out) const override {  // cube.rs:71-93
    counters().cube++;
    double xtmin, xtmax, ytmin, ytmax, ztmin, ztmax;
    check_axis(ray.origin.x, ray.direction.x, xtmin, xtmax);
    check_axis(ray.origin.y, ray.direction.y, ytmin, ytmax);
    check_axis(ray.origin.z, ray.direction.z, ztmin, ztmax);
    double tmin = rmax(rmax(xtmin, ytmin), ztmin);
    double tmax = rmin(rmin(xtmax, ytmax), ztmax);
    if (tmin > tmax) return;
    out.emplace_back(tmin, this);
    out.emplace_back(tmax, this);
  }
  Vector local_normal_at(Point p, const Intersection&) const override {  // cube.rs:95-105
    double maxc = rmax(rmax(std::fabs(p.x), std::fabs(p.y)), std::fabs(p.z));
    if (equal(maxc, std::fabs(p.x))) return Vector(p.x, 0.0, 0.0);
    if (equal(maxc, std::fabs(p.y))) return Vector(0.0, p.y, 0.0);
    return Vector(0.0, 0.0, p.z);
  }
};

struct Cylinder : Shape {  // cylinder.rs
  double minimum, maximum;
  bool closed;
  Cylinder(double mn = -INF, double mx = INF, bool cl = false) : minimum(mn), maximum(mx), closed(cl) {
    base.bounding_box = BoundingBox(Point(-1.0, mn, -1.0), Point(1.0, mx, 1.0));  // cylinder.rs:27-40
  }
  ShapeKind kind() const override { return ShapeKind::Cylinder; }
  bool equals(const Shape& o) const override {
    auto c = downcast<Cylinder>(o, ShapeKind::Cylinder);
    return c && base == c->base && minimum == c->minimum && maximum == c->maximum && closed == c->closed;
  }
  bool check_cap(const Ray& ray, double t) const {  // cylinder.rs:42-46
    double x = ray.origin.x + t * ray.direction.x;
    double z = ray.origin.z + t * ray.direction.z;
    return (x * x + z * z) <= 1.0;
  }
  void intersect_caps(const Ray& ray, Intersections& out) const {  // cylinder.rs:48-65 — no |d.y|<EPS guard
    if (!closed) return;
    counters().caps++;
    double t = (minimum - ray.origin.y) / ray.direction.y;
    if (check_cap(ray, t)) out.emplace_back(t, this);
    t = (maximum - ray.origin.y) / ray.direction.y;
    if (check_cap(ray, t)) out.emplace_back(t, this);
  }
  void local_intersect(const Ray& ray, Intersections& out) const override {  // cylinder.rs:88-119
    counters().cyl++;
    double a = ray.direction.x * ray.direction.x + ray.direction.z * ray.direction.z;
    if (std::fabs(a) < EPSILON) {
      intersect_caps(ray, out);
      return;
    }
    double b = 2.0 * ray.origin.x * ray.direction.x + 2.0 * ray.origin.z * ray.direction.z;
    double c = ray.origin.x * ray.origin.x + ray.origin.z * ray.origin.z - 1.0;
    double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return;  // before caps — kept
    double t0 = (-b - std::sqrt(disc)) / (2.0 * a);
    double t1 = (-b + std::sqrt(disc)) / (2.0 * a);
    double y0 = ray.origin.y + t0 * ray.direction.y;
    if (minimum < y0 && y0 < maximum) out.emplace_back(t0, this);
    double y1 = ray.origin.y + t1 * ray.direction.y;
    if (minimum < y1 && y1 < maximum) out.emplace_back(t1, this);
    intersect_caps(ray, out);
  }
  Vector local_normal_at(Point p, const Intersection&) const override {  // cylinder.rs:121-130
    double dist = p.x * p.x + p.z * p.z;
    if (dist < 1.0 && p.y >= maximum - EPSILON) return Vector(0, 1, 0);
    if (dist < 1.0 && p.y <= minimum + EPSILON) return Vector(0, -1, 0);
    return Vector(p.x, 0.0, p.z);
  }
};

struct Cone : Shape {  // cone.rs
  double minimum, maximum;
  bool closed;
  Cone(double mn = -INF, double mx = INF, bool cl = false) : minimum(mn), maximum(mx), closed(cl) {
    double limit = rmax(std::fabs(mn), std::fabs(mx));  // cone.rs:27-46
    base.bounding_box = BoundingBox(Point(-limit, mn, -limit), Point(limit, mx, limit));
  }
  ShapeKind kind() const override { return ShapeKind::Cone; }
  bool equals(const Shape& o) const override {
    auto c = downcast<Cone>(o, ShapeKind::Cone);
    return c && base == c->base && minimum == c->minimum && maximum == c->maximum && closed == c->closed;
  }
  bool check_cap(const Ray& ray, double t, double radius) const {  // cone.rs:67-71
    double x = ray.origin.x + t * ray.direction.x;
    double z = ray.origin.z + t * ray.direction.z;
    return (x * x + z * z) <= radius * radius;
  }
  void intersect_caps(const Ray& ray, Intersections& out) const {  // cone.rs:48-65
    if (!closed) return;
    counters().caps++;
    double t = (minimum - ray.origin.y) / ray.direction.y;
    if (check_cap(ray, t, minimum)) out.emplace_back(t, this);
    t = (maximum - ray.origin.y) / ray.direction.y;
    if (check_cap(ray, t, maximum)) out.emplace_back(t, this);
  }
  void local_intersect(const Ray& ray, Intersections& out) const override {  // cone.rs:94-134
    counters().cone++;
    const Point& o = ray.origin;
    const Vector& d = ray.direction;
    double a = d.x * d.x - d.y * d.y + d.z * d.z;
    double b = 2.0 * o.x * d.x - 2.0 * o.y * d.y + 2.0 * o.z * d.z;
    double c = o.x * o.x - o.y * o.y + o.z * o.z;
    if (std::fabs(a) < EPSILON) {
      if (std::fabs(b) < EPSILON) {
        intersect_caps(ray, out);
        return;
      }
      double t = -c / 2.0 * b;  // cone.rs:106 — (−c/2)·b, the reference's bug, kept
      out.emplace_back(t, this);
      intersect_caps(ray, out);
      return;
    }
    double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return;
    double t0 = (-b - std::sqrt(disc)) / (2.0 * a);
    double t1 = (-b + std::sqrt(disc)) / (2.0 * a);
    double y0 = o.y + t0 * d.y;
    if (minimum < y0 && y0 < maximum) out.emplace_back(t0, this);
    double y1 = o.y + t1 * d.y;
    if (minimum < y1 && y1 < maximum) out.emplace_back(t1, this);
    intersect_caps(ray, out);
  }
  Vector local_normal_at(Point p, const Intersection&) const override {  // cone.rs:136-149 (dist<1.0 kept)
    double dist = p.x * p.x + p.z * p.z;
    if (dist < 1.0 && p.y >= maximum - EPSILON) return Vector(0, 1, 0);
    if (dist < 1.0 && p.y <= minimum + EPSILON) return Vector(0, -1, 0);
    double y = std::sqrt(p.x * p.x + p.z * p.z);
    if (p.y > 0.0) y = -y;
    return Vector(p.x, y, p.z);
  }
};

struct Triangle : Shape {  // triangle.rs
  Point p1, p2, p3;
  Vector e1, e2, normal;
  Triangle(Point a, Point b, Point c) : p1(a), p2(b), p3(c) {  // triangle.rs:24-45
    BoundingBox bb;
    bb.add_point(p1); bb.add_point(p2); bb.add_point(p3);
    base.bounding_box = bb;
    e1 = p2 - p1;
    e2 = p3 - p1;
    normal = cross(e2, e1).normalize();
  }
  ShapeKind kind() const override { return ShapeKind::Triangle; }
  bool equals(const Shape& o) const override {
    auto t = downcast<Triangle>(o, ShapeKind::Triangle);
    return t && base == t->base && p1 == t->p1 && p2 == t->p2 && p3 == t->p3 && e1 == t->e1 && e2 == t->e2 &&
           normal == t->normal;
  }
  // triangle.rs:67-90 / smooth_triangle.rs:73-96 (Möller–Trumbore, object space)
  static bool moller_trumbore(const Ray& ray, Point p1, Vector e1, Vector e2, double& t, double& u, double& v) {
    Vector dir_cross_e2 = cross(ray.direction, e2);
    double det = dot(e1, dir_cross_e2);
    if (std::fabs(det) < EPSILON) { counters().tri_exit0++; return false; }
    double f = 1.0 / det;
    Vector p1_to_origin = ray.origin - p1;
    u = f * dot(p1_to_origin, dir_cross_e2);
    if (!(0.0 <= u && u <= 1.0)) { counters().tri_exit1++; return false; }
    Vector origin_cross_e1 = cross(p1_to_origin, e1);
    v = f * dot(ray.direction, origin_cross_e1);
    if (v < 0.0 || (u + v) > 1.0) { counters().tri_exit2++; return false; }
    t = f * dot(e2, origin_cross_e1);
    counters().tri_hit++;
    return true;
  }
  void local_intersect(const Ray& ray, Intersections& out) const override {
    double t, u, v;
    if (moller_trumbore(ray, p1, e1, e2, t, u, v)) out.emplace_back(t, this);
  }
  Vector local_normal_at(Point, const Intersection&) const override { return normal; }
};

struct SmoothTriangle : Shape {  // smooth_triangle.rs
  Point p1, p2, p3;
  Vector n1, n2, n3;
  Vector e1, e2, normal;
  SmoothTriangle(Point a, Point b, Point c, Vector na, Vector nb, Vector nc)
      : p1(a), p2(b), p3(c), n1(na), n2(nb), n3(nc) {  // smooth_triangle.rs:27-51
    BoundingBox bb;
    bb.add_point(p1); bb.add_point(p2); bb.add_point(p3);
    base.bounding_box = bb;
    e1 = p2 - p1;
    e2 = p3 - p1;
    normal = cross(e2, e1).normalize();
  }
  ShapeKind kind() const override { return ShapeKind::SmoothTriangle; }
  bool equals(const Shape& o) const override {
    auto t = downcast<SmoothTriangle>(o, ShapeKind::SmoothTriangle);
    return t && base == t->base && p1 == t->p1 && p2 == t->p2 && p3 == t->p3 && n1 == t->n1 && n2 == t->n2 &&
           n3 == t->n3 && e1 == t->e1 && e2 == t->e2 && normal == t->normal;
  }
  void local_intersect(const Ray& ray, Intersections& out) const override {
    double t, u, v;
    if (Triangle::moller_trumbore(ray, p1, e1, e2, t, u, v)) out.emplace_back(t, this, u, v);
  }
  Vector local_normal_at(Point, const Intersection& hit) const override {  // smooth_triangle.rs:98-102
    return n2 * hit.u + n3 * hit.v + n1 * (1.0 - hit.u - hit.v);
  }
};

// test_shape.rs — test double: remembers the object-space ray, never intersects.
struct TestShape : Shape {
  mutable Ray saved_ray{Point::origin(), Vector(0, 0, 0)};
  TestShape() { base.bounding_box = BoundingBox(Point(-1, -1, -1), Point(1, 1, 1)); }
  ShapeKind kind() const override { return ShapeKind::TestShape; }
  bool equals(const Shape& o) const override { return base == o.base; }  // test_shape.rs:42-44
  void local_intersect(const Ray& ray, Intersections&) const override { saved_ray = ray; }
  Vector local_normal_at(Point p, const Intersection&) const override { return Vector(p.x, p.y, p.z); }
};

// ---------------------------------------------------------------------------------------------
// group.rs
// ---------------------------------------------------------------------------------------------
struct Group : Shape {
  std::vector<ShapeP> children;
  ShapeKind kind() const override { return ShapeKind::Group; }
  bool equals(const Shape& o) const override {  // derived PartialEq: base + children pairwise by value
    auto g = downcast<Group>(o, ShapeKind::Group);
    if (!g || !(base == g->base) || children.size() != g->children.size()) return false;
    for (size_t i = 0; i < children.size(); i++)
      if (!children[i]->equals(*g->children[i])) return false;
    return true;
  }
  // group.rs:49-58 — world-space AABB vs the UNtransformed ray; then every child in order.
  void intersect(const Ray& ray, Intersections& out) const override {
    counters().nodes_visited++;
    if (!get_bounds().intersects(ray)) return;
    for (const ShapeP& c : children) c->intersect(ray, out);
  }
  void local_intersect(const Ray& ray, Intersections& out) const override {  // group.rs:60-65
    for (const ShapeP& c : children) c->intersect(ray, out);
  }
  Vector local_normal_at(Point, const Intersection&) const override {
    throw std::logic_error("Group::local_normal_at is unreachable (group.rs:67-69)");
  }
  // group.rs:71-94
  void set_transform(const Matrix& transform) override {
    Matrix inverse = base.transform_inverse();
    for (ShapeP& child : children) child->set_transform(inverse * child->base.transform());
    base.xf = make_xform(transform);
    Matrix t = base.transform();
    BoundingBox nb;
    for (ShapeP& child : children) {
      child->set_transform(t * child->base.transform());
      nb.add_bounding_box(child->get_bounds());
    }
    base.bounding_box = nb;
  }
  // group.rs:96-102
  void set_material(const Material& m) override { set_material_shared(std::make_shared<const Material>(m)); }
  void set_material_shared(std::shared_ptr<const Material> m) override {
    base.material_p = m;
    for (ShapeP& child : children) child->set_material_shared(m);
  }
  bool includes(const Shape& other) const override {  // group.rs:104-106
    for (const ShapeP& c : children)
      if (c->includes(other)) return true;
    return false;
  }
  // group.rs:128-133
  void add_child(ShapeP shape) {
    shape->set_transform(base.transform() * shape->base.transform());
    BoundingBox cbox = shape->parent_space_bounds();
    base.bounding_box.add_bounding_box(cbox);
    children.push_back(std::move(shape));
  }
  // group.rs:135-157 — the reference removes from a Vec one element at a time (O(n²)); a stable
  // two-pass partition yields the same three lists in the same order.
  void partition_children(std::vector<ShapeP>& left, std::vector<ShapeP>& right) {
    auto halves = get_bounds().split();
    std::vector<ShapeP> rest;
    rest.reserve(children.size());
    for (ShapeP& c : children) {
      if (halves.first.contains_bounding_box(c->parent_space_bounds())) left.push_back(std::move(c));
      else rest.push_back(std::move(c));
    }
    children.clear();
    for (ShapeP& c : rest) {
      if (halves.second.contains_bounding_box(c->parent_space_bounds())) right.push_back(std::move(c));
      else children.push_back(std::move(c));
    }
  }
  void make_subgroup(std::vector<ShapeP> shapes) {  // group.rs:191-197
    auto g = std::make_unique<Group>();
    for (ShapeP& s : shapes) g->add_child(std::move(s));
    children.push_back(std::move(g));
  }
  void divide(size_t threshold) override {  // group.rs:108-122
    if (threshold <= children.size()) {
      std::vector<ShapeP> left, right;
      partition_children(left, right);
      if (!left.empty()) make_subgroup(std::move(left));
      if (!right.empty()) make_subgroup(std::move(right));
    }
    for (ShapeP& c : children) c->divide(threshold);
  }
};

// ---------------------------------------------------------------------------------------------
// csg.rs
// ---------------------------------------------------------------------------------------------
enum class Operation : int { Union = 0, Intersection = 1, Difference = 2 };

inline bool intersection_allowed(Operation op, bool lhit, bool inl, bool inr) {  // csg.rs:22-40
  switch (op) {
    case Operation::Union: return (lhit && !inr) || (!lhit && !inl);
    case Operation::Intersection: return (lhit && inr) || (!lhit && inl);
    case Operation::Difference: return (lhit && !inr) || (!lhit && inl);
  }
  return false;
}

struct Csg : Shape {
  Operation operation;
  ShapeP left, right;
  Csg(Operation op, ShapeP l, ShapeP r) : operation(op), left(std::move(l)), right(std::move(r)) {  // csg.rs:52-69
    BoundingBox bb;
    bb.add_bounding_box(left->parent_space_bounds());
    bb.add_bounding_box(right->parent_space_bounds());
    base.bounding_box = bb;
  }
  ShapeKind kind() const override { return ShapeKind::Csg; }
  bool equals(const Shape& o) const override {  // csg.rs:107-113 (operation is NOT compared)
    auto c = downcast<Csg>(o, ShapeKind::Csg);
    return c && base == o.base && left->equals(*c->left) && right->equals(*c->right);
  }
  Intersections filter_intersections(const Intersections& xs) const {  // csg.rs:71-91
    bool inl = false, inr = false;
    Intersections result;
    for (const Intersection& i : xs) {
      bool lhit = left->includes(*i.object);
      if (intersection_allowed(operation, lhit, inl, inr)) result.push_back(i);
      if (lhit) inl = !inl;
      else inr = !inr;
    }
    return result;
  }
  // csg.rs:115-126 — the AABB (moved to PARENT space by set_transform) is tested against the
  // LOCAL ray: the reference's quirk, kept.
  void local_intersect(const Ray& ray, Intersections& out) const override {
    counters().nodes_visited++;
    if (!get_bounds().intersects(ray)) return;
    Intersections leftxs;
    left->intersect(ray, leftxs);
    right->intersect(ray, leftxs);
    Intersections xs = intersections(leftxs);
    Intersections f = filter_intersections(xs);
    out.insert(out.end(), f.begin(), f.end());
  }
  Vector local_normal_at(Point, const Intersection&) const override {
    throw std::logic_error("Csg::local_normal_at is unreachable (csg.rs:128-130)");
  }
  bool includes(const Shape& other) const override { return left->includes(other) || right->includes(other); }
  void divide(size_t threshold) override {  // csg.rs:136-139
    left->divide(threshold);
    right->divide(threshold);
  }
};

}  // namespace orc
