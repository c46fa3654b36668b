// rth_scene.hpp — host mirror of the reference's scene API for the B200 backend:
// Shape (tagged node, no virtuals), Group (add_child / set_transform / set_material / divide),
// Csg, Material, Pattern, PointLight, World, Camera. Names and semantics follow the reference so a
// scene written against raytracer-challenge-rs translates line by line:
//   geometry/mod.rs:13-108 (BaseShape, Shape::set_transform/no_shadow/parent_space_bounds),
//   geometry/shape/*.rs constructors (default bounding boxes), group.rs:71-197, csg.rs:52-69,136-139,
//   material.rs:11-36,84-86, pattern/mod.rs:17-91, light.rs:4-16, world.rs:18-29,83-93,137-151,
//   camera.rs:20-55,128-131,220-252.
// Nothing here intersects rays: rendering is `Camera::render(world)` → flatten (rth_flatten.hpp) →
// rtc_render (include/rtc.h) on the GPU.
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/rtc.h"
#include "rth_math.hpp"

namespace rth {

struct Xf {  // transform + inverse of a shape (the inverse-transpose is the same 3x3 block transposed)
  Mat4 m = Mat4::identity();
  Mat4 inv = Mat4::identity();
};
using XfP = std::shared_ptr<const Xf>;
inline XfP identity_xf() {
  static XfP id = std::make_shared<const Xf>();
  return id;
}
inline XfP make_xf(const Mat4& m) {
  auto x = std::make_shared<Xf>();
  x->m = m;
  x->inv = m.inverse();
  return x;
}
// a * child->m (matrix.rs:210-230) and its inverse, memoised on the child's shared Xf, so that
// shapes sharing an Xf (all triangles of a mesh) cost one product and one inverse. The keys are
// shared_ptrs: a key's address cannot be recycled while the memo lives.
struct XfMemo {
  XfP base;  // the left factor this memo is valid for
  std::unordered_map<XfP, XfP> map;
  XfP product(const XfP& a, const XfP& child, bool use_inverse = false) {
    if (base != a) { map.clear(); base = a; }
    auto it = map.find(child);
    if (it != map.end()) return it->second;
    XfP r = make_xf((use_inverse ? a->inv : a->m) * child->m);
    map.emplace(child, r);
    return r;
  }
};

struct Pattern {  // pattern/mod.rs:17-31
  int kind = RTC_PAT_TEST;
  double a[3] = {0, 0, 0}, b[3] = {0, 0, 0};
  Mat4 transform = Mat4::identity();
  Mat4 inverse = Mat4::identity();
  void set_transform(const Mat4& t) { transform = t; inverse = t.inverse(); }
};
inline Pattern make_pattern(int kind, double ar, double ag, double ab, double br, double bg, double bb) {
  Pattern p;
  p.kind = kind;
  p.a[0] = ar; p.a[1] = ag; p.a[2] = ab; p.b[0] = br; p.b[1] = bg; p.b[2] = bb;
  return p;
}

struct Material {  // material.rs:11-36
  double color[3] = {1.0, 1.0, 1.0};
  double ambient = 0.1, diffuse = 0.9, specular = 0.9, shininess = 200.0;
  double reflective = 0.0, transparency = 0.0, refractive_index = 1.0;
  bool has_pattern = false;
  Pattern pattern;
  void set_pattern(const Pattern& p) { has_pattern = true; pattern = p; }
};
using MaterialP = std::shared_ptr<const Material>;
inline MaterialP default_material() {
  static MaterialP d = std::make_shared<const Material>();
  return d;
}

struct TriData {
  Vec3 p1, p2, p3, e1, e2, normal, n1, n2, n3;
};

struct Shape;
using ShapeP = std::unique_ptr<Shape>;

struct Shape {
  int kind = RTC_SPHERE;
  XfP xf = identity_xf();
  MaterialP material = default_material();
  BBox box;
  bool shadow = true;
  double minimum = -INF, maximum = INF;  // cylinder / cone
  bool closed = false;
  int op = RTC_CSG_UNION;                // csg
  std::unique_ptr<TriData> tri;          // (smooth) triangle
  std::vector<ShapeP> children;          // group: children; csg: {left, right}

  // ---- constructors (default bounding boxes of each reference shape) --------------------------
  static ShapeP sphere() { return unit_box_shape(RTC_SPHERE); }  // sphere.rs:16-25
  static ShapeP cube() { return unit_box_shape(RTC_CUBE); }      // cube.rs:18-27
  static ShapeP glass_sphere() {                                  // sphere.rs:70-76
    ShapeP s = sphere();
    Material m;
    m.transparency = 1.0;
    m.refractive_index = 1.5;
    s->set_material(m);
    return s;
  }
  static ShapeP plane() {  // plane.rs:19-31
    ShapeP s(new Shape);
    s->kind = RTC_PLANE;
    s->box.min = {-INF, 0.0, -INF};
    s->box.max = {INF, 0.0, INF};
    return s;
  }
  static ShapeP cylinder(double mn = -INF, double mx = INF, bool closed = false) {  // cylinder.rs:27-40
    ShapeP s(new Shape);
    s->kind = RTC_CYLINDER;
    s->minimum = mn; s->maximum = mx; s->closed = closed;
    s->box.min = {-1.0, mn, -1.0};
    s->box.max = {1.0, mx, 1.0};
    return s;
  }
  static ShapeP cone(double mn = -INF, double mx = INF, bool closed = false) {  // cone.rs:27-46
    ShapeP s(new Shape);
    s->kind = RTC_CONE;
    s->minimum = mn; s->maximum = mx; s->closed = closed;
    double limit = std::fmax(std::fabs(mn), std::fabs(mx));
    s->box.min = {-limit, mn, -limit};
    s->box.max = {limit, mx, limit};
    return s;
  }
  static ShapeP triangle(Vec3 p1, Vec3 p2, Vec3 p3) {  // triangle.rs:24-45
    ShapeP s(new Shape);
    s->kind = RTC_TRIANGLE;
    s->init_tri(p1, p2, p3);
    return s;
  }
  static ShapeP smooth_triangle(Vec3 p1, Vec3 p2, Vec3 p3, Vec3 n1, Vec3 n2, Vec3 n3) {  // smooth_triangle.rs:27-51
    ShapeP s(new Shape);
    s->kind = RTC_SMOOTH_TRIANGLE;
    s->init_tri(p1, p2, p3);
    s->tri->n1 = n1; s->tri->n2 = n2; s->tri->n3 = n3;
    return s;
  }
  static ShapeP group() {  // group.rs:20-27: BaseShape::default() → empty (inverted) box
    ShapeP s(new Shape);
    s->kind = RTC_GROUP;
    return s;
  }
  static ShapeP csg(int op, ShapeP left, ShapeP right) {  // csg.rs:52-69
    ShapeP s(new Shape);
    s->kind = RTC_CSG;
    s->op = op;
    s->box.add_box(left->parent_space_bounds());
    s->box.add_box(right->parent_space_bounds());
    s->children.push_back(std::move(left));
    s->children.push_back(std::move(right));
    return s;
  }

  bool is_group() const { return kind == RTC_GROUP; }
  BBox parent_space_bounds() const { return box.transform(Mat4::identity()); }  // geometry/mod.rs:95-97
  void no_shadow() { shadow = false; }

  // geometry/mod.rs:74-85 for leaves and CSG; group.rs:71-94 for groups
  void set_transform(const Mat4& t) { set_transform_xf(make_xf(t)); }
  void set_transform_xf(XfP nx) {
    if (!is_group()) {
      box = box.transform(xf->inv);  // un-transform the stored AABB with the OLD inverse …
      xf = std::move(nx);
      box = box.transform(xf->m);    // … and re-transform with the new matrix (boxes inflate)
      return;
    }
    // remove the current transform from the children, then apply the new one; children that share
    // an Xf share the result (all triangles of a mesh → one product, one inverse)
    {
      XfMemo undo;
      for (ShapeP& c : children) c->set_transform_xf(undo.product(xf, c->xf, /*use_inverse=*/true));
    }
    xf = std::move(nx);
    BBox nb;
    XfMemo apply;
    for (ShapeP& c : children) {
      c->set_transform_xf(apply.product(xf, c->xf));
      nb.add_box(c->box);
    }
    box = nb;
    add_memo_.reset();
  }
  // geometry/mod.rs:66-68 / group.rs:96-102
  void set_material(const Material& m) { set_material_p(std::make_shared<const Material>(m)); }
  void set_material_p(const MaterialP& m) {
    material = m;
    if (is_group())
      for (ShapeP& c : children) c->set_material_p(m);
  }
  // group.rs:128-133
  void add_child(ShapeP c) {
    if (!is_group()) throw std::logic_error("add_child: not a group");
    if (!add_memo_) add_memo_.reset(new XfMemo);
    c->set_transform_xf(add_memo_->product(xf, c->xf));
    box.add_box(c->parent_space_bounds());
    children.push_back(std::move(c));
  }
  // group.rs:108-122 / csg.rs:136-139 / geometry/mod.rs:99
  void divide(size_t threshold) {
    if (kind == RTC_CSG) {
      children[0]->divide(threshold);
      children[1]->divide(threshold);
      return;
    }
    if (!is_group()) return;
    if (threshold <= children.size()) {
      std::vector<ShapeP> left, right;
      partition_children(left, right);
      if (!left.empty()) make_subgroup(std::move(left));
      if (!right.empty()) make_subgroup(std::move(right));
    }
    for (ShapeP& c : children) c->divide(threshold);
  }
  // group.rs:135-157: same three lists, same order, as the reference's remove-in-a-loop; O(n)
  void partition_children(std::vector<ShapeP>& left, std::vector<ShapeP>& right) {
    std::pair<BBox, BBox> halves = box.split();
    std::vector<ShapeP> rest;
    rest.reserve(children.size());
    for (ShapeP& c : children) {
      if (halves.first.contains_box(c->parent_space_bounds())) left.push_back(std::move(c));
      else rest.push_back(std::move(c));
    }
    children.clear();
    for (ShapeP& c : rest) {
      if (halves.second.contains_box(c->parent_space_bounds())) right.push_back(std::move(c));
      else children.push_back(std::move(c));
    }
  }
  void make_subgroup(std::vector<ShapeP> shapes) {  // group.rs:191-197
    ShapeP g = group();
    for (ShapeP& s : shapes) g->add_child(std::move(s));
    g->add_memo_.reset();
    children.push_back(std::move(g));
  }

 private:
  static ShapeP unit_box_shape(int k) {
    ShapeP s(new Shape);
    s->kind = k;
    s->box.min = {-1, -1, -1};
    s->box.max = {1, 1, 1};
    return s;
  }
  void init_tri(Vec3 p1, Vec3 p2, Vec3 p3) {
    tri.reset(new TriData);
    tri->p1 = p1; tri->p2 = p2; tri->p3 = p3;
    box.add_point(p1); box.add_point(p2); box.add_point(p3);
    tri->e1 = vsub(p2, p1);
    tri->e2 = vsub(p3, p1);
    tri->normal = vnormalize(vcross(tri->e2, tri->e1));
    tri->n1 = tri->n2 = tri->n3 = Vec3{0, 0, 0};
  }
  // add_child is called once per child by parsers and make_subgroup; a run of children sharing an
  // Xf then costs one product (dropped whenever the group's own transform changes).
  std::unique_ptr<XfMemo> add_memo_;
};

struct PointLight {  // light.rs:4-16
  double position[3];
  double intensity[3];
};

struct World {  // world.rs:18-29
  std::vector<ShapeP> objects;
  std::vector<PointLight> lights;
  void add_object(ShapeP s) { objects.push_back(std::move(s)); }
  void add_light(const PointLight& l) { lights.push_back(l); }
  static World default_world() {  // world.rs:137-151
    World w;
    w.add_light(PointLight{{-10, 10, -10}, {1.0, 1.0, 1.0}});
    ShapeP s1 = Shape::sphere();
    Material m;
    m.color[0] = 0.8; m.color[1] = 1.0; m.color[2] = 0.6;
    m.diffuse = 0.7;
    m.specular = 0.2;
    s1->set_material(m);
    ShapeP s2 = Shape::sphere();
    s2->set_transform(scaling(0.5, 0.5, 0.5));
    w.add_object(std::move(s1));
    w.add_object(std::move(s2));
    return w;
  }
};

struct RenderOpts {  // camera.rs:220-252
  size_t num_threads = 1;  // kept for API parity; the GPU backend ignores it
  int aa_samples = 1;      // AASamples::X1..X16
};

struct Camera {  // camera.rs:20-55
  size_t hsize, vsize;
  double field_of_view;
  Mat4 transform = Mat4::identity();
  Mat4 transform_inverse = Mat4::identity();
  double pixel_size, half_width, half_height;
  RenderOpts render_opts;
  Camera(size_t h, size_t v, double fov) : hsize(h), vsize(v), field_of_view(fov) {
    double half_view = std::tan(fov / 2.0);
    double aspect = (double)hsize / (double)vsize;
    if (aspect >= 1.0) { half_width = half_view; half_height = half_view / aspect; }
    else { half_width = half_view * aspect; half_height = half_view; }
    pixel_size = half_width * 2.0 / (double)hsize;
  }
  void set_transform(const Mat4& t) { transform = t; transform_inverse = t.inverse(); }  // camera.rs:128-131
  rtc_camera to_rtc() const {
    rtc_camera c;
    c.hsize = (uint32_t)hsize; c.vsize = (uint32_t)vsize;
    c.pixel_size = pixel_size; c.half_width = half_width; c.half_height = half_height;
    std::memcpy(c.inverse, transform_inverse.m, sizeof c.inverse);
    return c;
  }
};

}  // namespace rth
