// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// Handle-based C API over the CPU restatement, for ctypes (tests/, bench.py's cpu_baseline /
// --impl reference leg, __graft_entry__.smoke()). The product's host library exposes the same
// calls under the `rth_` prefix so one Python scene script can build the scene in both.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>

#include "obj.hpp"
#include "world.hpp"

using namespace orc;

namespace {
thread_local std::string g_err;
int fail(const std::exception& e) {
  g_err = e.what();
  return -1;
}
Matrix mat16(const double* m) { return Matrix::from_slice(4, 4, m, 16); }
}  // namespace

extern "C" {

struct orc_material {
  double color[3];
  double ambient, diffuse, specular, shininess, reflective, transparency, refractive_index;
  int32_t has_pattern;   // bit 0: material has a pattern; bit 1: the pattern had set_transform called
  int32_t pattern_kind;  // 0 test, 1 stripe, 2 gradient, 3 ring, 4 checkers
  double pattern_a[3], pattern_b[3];
  double pattern_transform[16];
};

const char* orc_last_error() { return g_err.c_str(); }

void* orc_shape_new(int kind, double minimum, double maximum, int closed) {
  switch ((ShapeKind)kind) {
    case ShapeKind::Sphere: return new Sphere();
    case ShapeKind::Plane: return new Plane();
    case ShapeKind::Cube: return new Cube();
    case ShapeKind::Cylinder: return new Cylinder(minimum, maximum, closed != 0);
    case ShapeKind::Cone: return new Cone(minimum, maximum, closed != 0);
    case ShapeKind::Group: return new Group();
    case ShapeKind::TestShape: return new TestShape();
    default: g_err = "orc_shape_new: bad kind"; return nullptr;
  }
}
void* orc_triangle_new(const double* p) {
  return new Triangle(Point(p[0], p[1], p[2]), Point(p[3], p[4], p[5]), Point(p[6], p[7], p[8]));
}
void* orc_smooth_triangle_new(const double* p, const double* n) {
  return new SmoothTriangle(Point(p[0], p[1], p[2]), Point(p[3], p[4], p[5]), Point(p[6], p[7], p[8]),
                            Vector(n[0], n[1], n[2]), Vector(n[3], n[4], n[5]), Vector(n[6], n[7], n[8]));
}
void* orc_csg_new(int op, void* left, void* right) {
  return new Csg((Operation)op, ShapeP((Shape*)left), ShapeP((Shape*)right));
}
void orc_shape_free(void* s) { delete (Shape*)s; }
int orc_shape_kind(void* s) { return (int)((Shape*)s)->kind(); }

int orc_group_add_child(void* g, void* child) {
  Shape* s = (Shape*)g;
  if (s->kind() != ShapeKind::Group) { g_err = "orc_group_add_child: not a group"; return -1; }
  try {
    static_cast<Group*>(s)->add_child(ShapeP((Shape*)child));
  } catch (const std::exception& e) { return fail(e); }
  return 0;
}
// Bulk mesh: n triangles, 9 doubles each (+9 normals each when `normals` != NULL), appended to
// the group with add_child exactly as the OBJ parser would.
int orc_group_add_triangles(void* g, size_t n, const double* p, const double* normals) {
  Shape* s = (Shape*)g;
  if (s->kind() != ShapeKind::Group) { g_err = "orc_group_add_triangles: not a group"; return -1; }
  Group* grp = static_cast<Group*>(s);
  try {
    for (size_t i = 0; i < n; i++) {
      const double* q = p + 9 * i;
      if (normals) grp->add_child(ShapeP((Shape*)orc_smooth_triangle_new(q, normals + 9 * i)));
      else grp->add_child(ShapeP((Shape*)orc_triangle_new(q)));
    }
  } catch (const std::exception& e) { return fail(e); }
  return 0;
}
int orc_shape_set_transform(void* s, const double* m) {
  try {
    ((Shape*)s)->set_transform(mat16(m));
  } catch (const std::exception& e) { return fail(e); }
  return 0;
}
void orc_shape_get_transform(void* s, double* m, double* inv) {
  Shape* sh = (Shape*)s;
  if (m) std::memcpy(m, sh->base.transform().e.data(), 16 * sizeof(double));
  if (inv) std::memcpy(inv, sh->base.transform_inverse().e.data(), 16 * sizeof(double));
}
int orc_shape_set_material(void* s, const orc_material* m) {
  try {
    Material mat;
    mat.color = Color(m->color[0], m->color[1], m->color[2]);
    mat.ambient = m->ambient; mat.diffuse = m->diffuse; mat.specular = m->specular; mat.shininess = m->shininess;
    mat.reflective = m->reflective; mat.transparency = m->transparency; mat.refractive_index = m->refractive_index;
    if (m->has_pattern) {
      Pattern p = make_pattern((PatternKind)m->pattern_kind, Color(m->pattern_a[0], m->pattern_a[1], m->pattern_a[2]),
                               Color(m->pattern_b[0], m->pattern_b[1], m->pattern_b[2]));
      if (m->has_pattern & 2) p.set_transform(mat16(m->pattern_transform));  // bit 1: set_transform was called
      mat.set_pattern(p);
    }
    ((Shape*)s)->set_material(mat);
  } catch (const std::exception& e) { return fail(e); }
  return 0;
}
void orc_shape_no_shadow(void* s) { ((Shape*)s)->no_shadow(); }
void orc_shape_divide(void* s, size_t threshold) { ((Shape*)s)->divide(threshold); }
void orc_shape_bounds(void* s, double* out) {
  const BoundingBox& b = ((Shape*)s)->get_bounds();
  out[0] = b.min.x; out[1] = b.min.y; out[2] = b.min.z; out[3] = b.max.x; out[4] = b.max.y; out[5] = b.max.z;
}
size_t orc_shape_child_count(void* s) {
  Shape* sh = (Shape*)s;
  if (sh->kind() == ShapeKind::Group) return static_cast<Group*>(sh)->children.size();
  if (sh->kind() == ShapeKind::Csg) return 2;
  return 0;
}
void* orc_shape_child(void* s, size_t i) {  // borrowed
  Shape* sh = (Shape*)s;
  if (sh->kind() == ShapeKind::Group) return static_cast<Group*>(sh)->children.at(i).get();
  if (sh->kind() == ShapeKind::Csg) return i == 0 ? static_cast<Csg*>(sh)->left.get() : static_cast<Csg*>(sh)->right.get();
  return nullptr;
}
void* orc_obj_parse(const char* text, size_t len) {
  try {
    ObjParser p = parse_obj_string(std::string(text, len));
    return p.as_group().release();
  } catch (const std::exception& e) { fail(e); return nullptr; }
}

void* orc_world_new() { return new World(); }
void orc_world_free(void* w) { delete (World*)w; }
void orc_world_add_object(void* w, void* s) { ((World*)w)->add_object(ShapeP((Shape*)s)); }
void orc_world_add_light(void* w, const double* pos, const double* intensity) {
  ((World*)w)->add_light(PointLight(Point(pos[0], pos[1], pos[2]), Color(intensity[0], intensity[1], intensity[2])));
}

void* orc_camera_new(size_t hsize, size_t vsize, double fov) { return new Camera(hsize, vsize, fov); }
void orc_camera_free(void* c) { delete (Camera*)c; }
int orc_camera_set_transform(void* c, const double* m) {
  try {
    ((Camera*)c)->set_transform(mat16(m));
  } catch (const std::exception& e) { return fail(e); }
  return 0;
}
// pixel_size, half_width, half_height, inverse[16]
void orc_camera_params(void* c, double* out) {
  Camera* cam = (Camera*)c;
  out[0] = cam->pixel_size; out[1] = cam->half_width; out[2] = cam->half_height;
  std::memcpy(out + 3, cam->transform_inverse.e.data(), 16 * sizeof(double));
}

// stats layout (uint64): the Counters struct verbatim, then [n] = elapsed microseconds.
size_t orc_counters_len() { return sizeof(Counters) / sizeof(uint64_t) + 1; }
double orc_counters_flops(const uint64_t* c) {
  Counters k;
  std::memcpy((void*)&k, c, sizeof(Counters));
  return k.flops();
}

// Render rows [row0,row1) of the frame into rgb (hsize*vsize*3 doubles, row-major y*w+x like
// Canvas). threads==0 → Camera::render semantics (serial, no AA averaging; needs aa==1);
// threads>=1 → render_multithreaded semantics over the row range (contiguous blocks, last thread
// takes the remainder), aa ∈ {1,2,4,8,16}.
int orc_render(void* cam_, void* world_, int aa, size_t threads, size_t row0, size_t row1, double* rgb,
               uint64_t* stats) {
  try {
    Camera cam = *(Camera*)cam_;
    const World& world = *(World*)world_;
    if (row1 > cam.vsize) row1 = cam.vsize;
    cam.render_opts.aa_samples = (AASamples)aa;
    Counters total;
    auto t0 = std::chrono::steady_clock::now();
    if (threads == 0) {
      if (aa != 1) throw std::invalid_argument("orc_render: serial Camera::render has no AA");
      counters() = Counters();
      for (size_t y = row0; y < row1; y++)
        for (size_t x = 0; x < cam.hsize; x++) {
          Ray ray = cam.ray_for_pixel(x, y);
          counters().rays_primary++;
          counters().pixels++;
          Color c = world.color_at(ray, MAX_RECURSION_DEPTH);
          double* o = rgb + (y * cam.hsize + x) * 3;
          o[0] = c.red; o[1] = c.green; o[2] = c.blue;
        }
      total = counters();
    } else {
      size_t rows = row1 > row0 ? row1 - row0 : 0;
      if (threads > rows && rows > 0) threads = rows;
      if (rows == 0) threads = 1;
      size_t rpt = rows / threads;
      std::vector<std::thread> handles;
      std::vector<Counters> per(threads);
      std::vector<std::string> errs(threads);
      for (size_t i = 0; i < threads; i++) {
        handles.emplace_back([&, i]() {
          try {
            size_t start = row0 + i * rpt, end = row0 + i * rpt + rpt;
            if (i == threads - 1) end = row1;
            counters() = Counters();
            for (size_t y = start; y < end; y++)
              for (size_t x = 0; x < cam.hsize; x++) {
                Color c = cam.pixel_aa(world, x, y);
                double* o = rgb + (y * cam.hsize + x) * 3;
                o[0] = c.red; o[1] = c.green; o[2] = c.blue;
              }
            per[i] = counters();
          } catch (const std::exception& e) { errs[i] = e.what(); }
        });
      }
      for (auto& h : handles) h.join();
      for (auto& e : errs)
        if (!e.empty()) throw std::runtime_error(e);
      for (auto& c : per) total.add(c);
    }
    auto t1 = std::chrono::steady_clock::now();
    if (stats) {
      std::memcpy(stats, &total, sizeof(Counters));
      stats[sizeof(Counters) / sizeof(uint64_t)] =
          (uint64_t)std::chrono::duration_cast<std::chrono::microseconds>(t1 - t0).count();
    }
  } catch (const std::exception& e) { return fail(e); }
  return 0;
}

// Ray-batch probe: colour = World::color_at(ray, remaining); t = hit t (or -1 when no hit).
int orc_trace_rays(void* world_, size_t n, const double* origins, const double* dirs, int remaining, double* rgb,
                   double* t_out) {
  try {
    const World& world = *(World*)world_;
    for (size_t i = 0; i < n; i++) {
      Ray r(Point(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
            Vector(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
      if (rgb) {
        Color c = world.color_at(r, (size_t)remaining);
        rgb[3 * i] = c.red; rgb[3 * i + 1] = c.green; rgb[3 * i + 2] = c.blue;
      }
      if (t_out) {
        Intersections xs = world.intersect(r);
        const Intersection* h = hit(xs);
        t_out[i] = h ? h->t : -1.0;
      }
    }
  } catch (const std::exception& e) { return fail(e); }
  return 0;
}

// Hit-point probe for one ray: out = [hit(0/1), t, point(3), normal(3), over(3), under(3), reflectv(3), n1, n2, inside]
int orc_probe_hit(void* world_, const double* o, const double* d, double* out) {
  try {
    const World& world = *(World*)world_;
    Ray r(Point(o[0], o[1], o[2]), Vector(d[0], d[1], d[2]));
    Intersections xs = world.intersect(r);
    const Intersection* h = hit(xs);
    for (int i = 0; i < 20; i++) out[i] = 0.0;
    if (!h) return 0;
    Computations c = prepare_computations(*h, r, xs);
    out[0] = 1.0; out[1] = c.t;
    out[2] = c.point.x; out[3] = c.point.y; out[4] = c.point.z;
    out[5] = c.normalv.x; out[6] = c.normalv.y; out[7] = c.normalv.z;
    out[8] = c.over_point.x; out[9] = c.over_point.y; out[10] = c.over_point.z;
    out[11] = c.under_point.x; out[12] = c.under_point.y; out[13] = c.under_point.z;
    out[14] = c.reflectv.x; out[15] = c.reflectv.y; out[16] = c.reflectv.z;
    out[17] = c.n1; out[18] = c.n2; out[19] = c.inside ? 1.0 : 0.0;
  } catch (const std::exception& e) { return fail(e); }
  return 0;
}

int orc_is_shadowed(void* world_, const double* p, size_t light_index) {
  const World& world = *(World*)world_;
  try {
    return world.is_shadowed(Point(p[0], p[1], p[2]), world.lights.at(light_index)) ? 1 : 0;
  } catch (const std::exception& e) { return fail(e); }
}

// canvas (hsize*vsize*3 doubles) → P3 PPM text (ppm.rs:24-51). Returns the byte length; copies at
// most `cap` bytes into `out`.
size_t orc_canvas_to_ppm(const double* rgb, size_t width, size_t height, char* out, size_t cap) {
  Canvas c(width, height);
  for (size_t i = 0; i < width * height; i++) c.pixels[i] = Color(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
  std::string s = canvas_to_ppm(c);
  if (out && cap) std::memcpy(out, s.data(), s.size() < cap ? s.size() : cap);
  return s.size();
}
void orc_quantize(const double* rgb, size_t n, uint8_t* out) {
  for (size_t i = 0; i < n; i++) out[i] = scale_color_component(rgb[i]);
}

}  // extern "C"
