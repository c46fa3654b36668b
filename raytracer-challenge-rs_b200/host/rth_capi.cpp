// rth_capi.cpp — librth_host.so: the host mirror of the reference's scene API (rth_scene.hpp),
// the flattener, the OBJ front end and the output stage behind a handle-based C API, plus
// rth_render = `Camera::render(&World) -> Canvas` (camera.rs:133) on the GPU:
//     flatten(World) → rtc_scene_create → rtc_render → rtc_scene_destroy      (include/rtc.h)
// There is no CPU render path in this library.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "rth_flatten.hpp"
#include "rth_obj.hpp"
#include "rth_yaml.hpp"

using namespace rth;

namespace {
thread_local std::string g_err;
int fail(const std::string& m) {
  g_err = m;
  return -1;
}
Vec3 v3(const double* p) { return Vec3{p[0], p[1], p[2]}; }
}  // namespace

extern "C" {

struct rth_material {  // same layout as the oracle's C API so one Python struct serves both
  double color[3];
  double ambient, diffuse, specular, shininess, reflective, transparency, refractive_index;
  int32_t has_pattern;   // bit 0: material has a pattern; bit 1: the pattern had set_transform called
  int32_t pattern_kind;
  double pattern_a[3], pattern_b[3];
  double pattern_transform[16];
};

const char* rth_last_error() { return g_err.c_str(); }

// kind: rtc_kind 0..4 (sphere, plane, cube, cylinder, cone) or 7 (group)
void* rth_shape_new(int kind, double minimum, double maximum, int closed) {
  switch (kind) {
    case RTC_SPHERE: return Shape::sphere().release();
    case RTC_PLANE: return Shape::plane().release();
    case RTC_CUBE: return Shape::cube().release();
    case RTC_CYLINDER: return Shape::cylinder(minimum, maximum, closed != 0).release();
    case RTC_CONE: return Shape::cone(minimum, maximum, closed != 0).release();
    case RTC_GROUP: return Shape::group().release();
    default: g_err = "rth_shape_new: bad kind"; return nullptr;
  }
}
void* rth_triangle_new(const double* p) { return Shape::triangle(v3(p), v3(p + 3), v3(p + 6)).release(); }
void* rth_smooth_triangle_new(const double* p, const double* n) {
  return Shape::smooth_triangle(v3(p), v3(p + 3), v3(p + 6), v3(n), v3(n + 3), v3(n + 6)).release();
}
void* rth_csg_new(int op, void* left, void* right) {
  return Shape::csg(op, ShapeP((Shape*)left), ShapeP((Shape*)right)).release();
}
void rth_shape_free(void* s) { delete (Shape*)s; }
int rth_shape_kind(void* s) { return ((Shape*)s)->kind; }
int rth_group_add_child(void* g, void* child) {
  try {
    ((Shape*)g)->add_child(ShapeP((Shape*)child));
  } catch (const std::exception& e) { return fail(e.what()); }
  return 0;
}
int rth_group_add_triangles(void* g, size_t n, const double* p, const double* normals) {
  try {
    Shape* grp = (Shape*)g;
    for (size_t i = 0; i < n; i++) {
      const double* q = p + 9 * i;
      if (normals) {
        const double* m = normals + 9 * i;
        grp->add_child(Shape::smooth_triangle(v3(q), v3(q + 3), v3(q + 6), v3(m), v3(m + 3), v3(m + 6)));
      } else {
        grp->add_child(Shape::triangle(v3(q), v3(q + 3), v3(q + 6)));
      }
    }
  } catch (const std::exception& e) { return fail(e.what()); }
  return 0;
}
int rth_shape_set_transform(void* s, const double* m) {
  try {
    ((Shape*)s)->set_transform(Mat4::from(m));
  } catch (const std::exception& e) { return fail(e.what()); }
  return 0;
}
void rth_shape_get_transform(void* s, double* m, double* inv) {
  Shape* sh = (Shape*)s;
  if (m) std::memcpy(m, sh->xf->m.m, 16 * sizeof(double));
  if (inv) std::memcpy(inv, sh->xf->inv.m, 16 * sizeof(double));
}
int rth_shape_set_material(void* s, const rth_material* m) {
  try {
    Material mat;
    std::memcpy(mat.color, m->color, sizeof mat.color);
    mat.ambient = m->ambient; mat.diffuse = m->diffuse; mat.specular = m->specular; mat.shininess = m->shininess;
    mat.reflective = m->reflective; mat.transparency = m->transparency; mat.refractive_index = m->refractive_index;
    if (m->has_pattern) {
      Pattern p = make_pattern(m->pattern_kind, m->pattern_a[0], m->pattern_a[1], m->pattern_a[2], m->pattern_b[0],
                               m->pattern_b[1], m->pattern_b[2]);
      if (m->has_pattern & 2) p.set_transform(Mat4::from(m->pattern_transform));  // bit 1: set_transform was called
      mat.set_pattern(p);
    }
    ((Shape*)s)->set_material(mat);
  } catch (const std::exception& e) { return fail(e.what()); }
  return 0;
}
void rth_shape_no_shadow(void* s) { ((Shape*)s)->no_shadow(); }
void rth_shape_divide(void* s, size_t threshold) { ((Shape*)s)->divide(threshold); }
void rth_shape_bounds(void* s, double* out) {
  const BBox& b = ((Shape*)s)->box;
  out[0] = b.min.x; out[1] = b.min.y; out[2] = b.min.z; out[3] = b.max.x; out[4] = b.max.y; out[5] = b.max.z;
}
size_t rth_shape_child_count(void* s) { return ((Shape*)s)->children.size(); }
void* rth_shape_child(void* s, size_t i) { return ((Shape*)s)->children.at(i).get(); }
void* rth_obj_parse(const char* text, size_t len) {
  try {
    ObjParser p;
    p.parse(text, len);
    return p.as_group().release();
  } catch (const std::exception& e) { fail(e.what()); return nullptr; }
}

// SceneParser::load_file (scene-parser/src/lib.rs:90-119). On success *world_out owns the World
// (rth_world_free) and *camera_out the Camera (rth_camera_free; NULL when the file adds none — the
// reference panics only later, at render, lib.rs:273). flags bit 0: accept the extended `add:` kinds
// (cylinder, cone, group, csg, obj); base_dir (may be NULL): where relative `obj` file paths start.
int rth_yaml_load(const char* text, size_t len, uint32_t flags, const char* base_dir, void** world_out,
                  void** camera_out, int* skipped) {
  try {
    std::unique_ptr<YamlScene> sc =
        load_yaml_scene(std::string(text, len), (flags & 1u) != 0, base_dir ? std::string(base_dir) : std::string());
    if (skipped) *skipped = sc->skipped;
    *camera_out = sc->camera.release();
    *world_out = new World(std::move(sc->world));
    return 0;
  } catch (const std::exception& e) { return fail(e.what()); }
}

void* rth_world_new() { return new World(); }
void rth_world_free(void* w) { delete (World*)w; }
void rth_world_add_object(void* w, void* s) { ((World*)w)->add_object(ShapeP((Shape*)s)); }
void rth_world_add_light(void* w, const double* pos, const double* intensity) {
  PointLight l;
  std::memcpy(l.position, pos, sizeof l.position);
  std::memcpy(l.intensity, intensity, sizeof l.intensity);
  ((World*)w)->add_light(l);
}

void* rth_camera_new(size_t hsize, size_t vsize, double fov) { return new Camera(hsize, vsize, fov); }
void rth_camera_free(void* c) { delete (Camera*)c; }
int rth_camera_set_transform(void* c, const double* m) {
  try {
    ((Camera*)c)->set_transform(Mat4::from(m));
  } catch (const std::exception& e) { return fail(e.what()); }
  return 0;
}
void rth_camera_params(void* c, double* out) {
  Camera* cam = (Camera*)c;
  out[0] = cam->pixel_size; out[1] = cam->half_width; out[2] = cam->half_height;
  std::memcpy(out + 3, cam->transform_inverse.m, 16 * sizeof(double));
}
void rth_camera_dims(void* c, size_t* hsize, size_t* vsize, double* fov) {
  const Camera* cam = (const Camera*)c;
  *hsize = cam->hsize; *vsize = cam->vsize; *fov = cam->field_of_view;
}
void rth_world_counts(void* w, size_t* n_objects, size_t* n_lights) {
  *n_objects = ((World*)w)->objects.size();
  *n_lights = ((World*)w)->lights.size();
}
void* rth_world_object(void* w, size_t i) { return ((World*)w)->objects.at(i).get(); }
void rth_camera_to_rtc(void* c, rtc_camera* out) { *out = ((Camera*)c)->to_rtc(); }

// ---- flattener --------------------------------------------------------------------------------
void* rth_flatten(void* world) {
  try {
    return new FlatScene(flatten(*(World*)world));
  } catch (const std::exception& e) { fail(e.what()); return nullptr; }
}
void rth_flat_free(void* f) { delete (FlatScene*)f; }
void rth_flat_desc(void* f, rtc_scene_desc* out) { *out = ((FlatScene*)f)->desc(); }
// info[0..7] = nodes, leaves, groups, csgs, xforms, materials, triangles, bytes
void rth_flat_info(void* f, uint64_t* info) {
  FlatScene* s = (FlatScene*)f;
  info[0] = s->nodes.size(); info[1] = s->n_leaves; info[2] = s->n_groups; info[3] = s->n_csg;
  info[4] = s->xforms.size() / 12; info[5] = s->materials.size(); info[6] = s->tri_geom.size() / 9; info[7] = s->bytes();
}

// ---- Camera::render on the GPU ------------------------------------------------------------------
// rgb: hsize*vsize*3 doubles (host), Canvas layout. aa = 1 → Camera::render (camera.rs:133);
// aa > 1 or flags & RTC_OPT_AVERAGE → render_multithreaded semantics (camera.rs:150).
// times_ms[0..3] (optional) = flatten, upload (rtc_scene_create), render (rtc_render), total.
int rth_render(void* cam_, void* world_, const rtc_opts* opts, int device, double* rgb, rtc_stats* stats, double* times_ms) {
  using clk = std::chrono::steady_clock;
  auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
  try {
    auto t0 = clk::now();
    FlatScene flat = flatten(*(World*)world_);
    auto t1 = clk::now();
    rtc_scene_desc d = flat.desc();
    rtc_scene* scene = nullptr;
    int rc = rtc_scene_create(&d, device, &scene);
    if (rc) return fail(std::string("rtc_scene_create: ") + rtc_last_error()), rc;
    auto t2 = clk::now();
    rtc_camera rc_cam = ((Camera*)cam_)->to_rtc();
    rc = rtc_render(scene, &rc_cam, opts, rgb, stats);
    auto t3 = clk::now();
    std::string msg = rc ? rtc_last_error() : "";
    rtc_scene_destroy(scene);
    if (times_ms) { times_ms[0] = ms(t0, t1); times_ms[1] = ms(t1, t2); times_ms[2] = ms(t2, t3); times_ms[3] = ms(t0, clk::now()); }
    if (rc) return fail("rtc_render: " + msg), rc;
  } catch (const std::exception& e) { return fail(e.what()); }
  return 0;
}

// ---- output stage (image/ppm.rs:24-75) ----------------------------------------------------------
static inline uint8_t quant(double v) {  // (v*255).round() as u8
  double r = std::round(v * 255.0);
  if (r != r || r <= 0.0) return 0;
  if (r >= 255.0) return 255;
  return (uint8_t)r;
}
void rth_quantize(const double* rgb, size_t n, uint8_t* out) {
  for (size_t i = 0; i < n; i++) out[i] = quant(rgb[i]);
}
// P3 text with the reference's 70-column wrapping. Returns the length; writes at most cap bytes.
size_t rth_canvas_to_ppm(const double* rgb, size_t width, size_t height, char* out, size_t cap) {
  std::string ppm;
  ppm.reserve(width * height * 12 + 32);
  char buf[64];
  std::snprintf(buf, sizeof buf, "P3\n%zu %zu\n255\n", width, height);
  ppm += buf;
  for (size_t j = 0; j < height; j++) {
    size_t line_len = 0;       // characters on the current output line
    size_t line_start = ppm.size();
    for (size_t i = 0; i < width; i++) {
      for (int c = 0; c < 3; c++) {
        int n = std::snprintf(buf, sizeof buf, "%d", (int)quant(rgb[(j * width + i) * 3 + (size_t)c]));
        if (line_len + (size_t)n > 70) {
          while (ppm.size() > line_start && ppm.back() == ' ') ppm.pop_back();  // trim_end
          ppm += '\n';
          line_start = ppm.size();
          line_len = 0;
        }
        ppm.append(buf, (size_t)n);
        line_len += (size_t)n;
        if (c < 2) { ppm += ' '; line_len++; }
      }
      if (i + 1 < width) { ppm += ' '; line_len++; }
    }
    ppm += '\n';
  }
  if (out && cap) std::memcpy(out, ppm.data(), ppm.size() < cap ? ppm.size() : cap);
  return ppm.size();
}
int rth_save_ppm(const double* rgb, size_t width, size_t height, const char* path) {
  size_t n = rth_canvas_to_ppm(rgb, width, height, nullptr, 0);
  std::string s(n, '\0');
  rth_canvas_to_ppm(rgb, width, height, &s[0], n);
  std::ofstream f(path, std::ios::binary);
  if (!f) return fail(std::string("cannot open ") + path);
  f.write(s.data(), (std::streamsize)s.size());
  return f ? 0 : fail("write failed");
}

// ---- 8-bit image files (image/png.rs:13-31; binary PPM) -------------------------------------------
// Both take already quantised pixels (rtc_render_rgb8 / rtc_quantize_device / rth_quantize).
int rth_save_p6(const uint8_t* rgb8, size_t width, size_t height, const char* path) {
  std::ofstream f(path, std::ios::binary);
  if (!f) return fail(std::string("cannot open ") + path);
  char head[64];
  int n = std::snprintf(head, sizeof head, "P6\n%zu %zu\n255\n", width, height);
  f.write(head, n);
  f.write((const char*)rgb8, (std::streamsize)(width * height * 3));
  return f ? 0 : fail("write failed");
}

namespace {
struct Crc32 {
  uint32_t table[256];
  Crc32() {
    for (uint32_t i = 0; i < 256; i++) {
      uint32_t c = i;
      for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
      table[i] = c;
    }
  }
  uint32_t run(uint32_t crc, const uint8_t* p, size_t n) const {
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return crc;
  }
};
void put_be32(std::string& s, uint32_t v) {
  for (int k = 3; k >= 0; k--) s.push_back((char)((v >> (8 * k)) & 0xFFu));
}
void png_chunk(std::string& out, const char type[4], const std::string& data) {
  static const Crc32 crc;
  put_be32(out, (uint32_t)data.size());
  std::string body(type, 4);
  body += data;
  out += body;
  put_be32(out, crc.run(0xFFFFFFFFu, (const uint8_t*)body.data(), body.size()) ^ 0xFFFFFFFFu);
}
}  // namespace

// 8-bit RGB PNG, filter 0 on every row, zlib stream of stored (uncompressed) deflate blocks: any PNG
// reader decodes it to exactly the bytes handed in, which is all the parity contract asks of the file.
int rth_save_png(const uint8_t* rgb8, size_t width, size_t height, const char* path) {
  if (width == 0 || height == 0 || width > 0x7FFFFFFFu || height > 0x7FFFFFFFu) return fail("bad image size");
  std::string raw;
  raw.reserve(height * (width * 3 + 1));
  for (size_t y = 0; y < height; y++) {
    raw.push_back(0);
    raw.append((const char*)rgb8 + y * width * 3, width * 3);
  }
  std::string z("\x78\x01", 2);
  uint32_t a = 1, b = 0;  // adler32
  for (size_t off = 0; off < raw.size();) {
    size_t n = std::min<size_t>(65535, raw.size() - off);
    z.push_back(off + n == raw.size() ? 1 : 0);
    z.push_back((char)(n & 0xFF)); z.push_back((char)(n >> 8));
    z.push_back((char)(~n & 0xFF)); z.push_back((char)((~n >> 8) & 0xFF));
    z.append(raw, off, n);
    for (size_t i = 0; i < n; i++) {
      a += (uint8_t)raw[off + i];
      if (a >= 65521) a -= 65521;
      b += a;
      if (b >= 65521) b -= 65521;
    }
    off += n;
  }
  put_be32(z, (b << 16) | a);
  std::string out("\x89PNG\r\n\x1a\n", 8), ihdr;
  put_be32(ihdr, (uint32_t)width);
  put_be32(ihdr, (uint32_t)height);
  ihdr += std::string("\x08\x02\x00\x00\x00", 5);  // 8 bit, colour type 2 (RGB), deflate, filter 0, no interlace
  png_chunk(out, "IHDR", ihdr);
  png_chunk(out, "IDAT", z);
  png_chunk(out, "IEND", std::string());
  std::ofstream f(path, std::ios::binary);
  if (!f) return fail(std::string("cannot open ") + path);
  f.write(out.data(), (std::streamsize)out.size());
  return f ? 0 : fail("write failed");
}

// render_scene <scene-file> <output-file> (scene-parser/src/bin/render_scene.rs + lib.rs:90-119,263-283)
// on the GPU: load the YAML, flatten, upload, render with 8-bit output, write .png / .ppm (P3, the
// reference's PPM) / anything else as binary P6. flags bit 0: extended `add:` kinds; aa: samples.
int rth_render_scene_file(const char* scene_path, const char* out_path, uint32_t flags, uint32_t aa, int device,
                          rtc_stats* stats) {
  try {
    std::ifstream in(scene_path, std::ios::binary);
    if (!in) return fail(std::string("cannot open ") + scene_path);
    std::string text((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    std::string dir(scene_path);
    size_t slash = dir.find_last_of('/');
    dir = slash == std::string::npos ? std::string(".") : dir.substr(0, slash);
    std::unique_ptr<YamlScene> sc = load_yaml_scene(text, (flags & 1u) != 0, dir);
    if (!sc->camera) return fail("the scene adds no camera");  // the reference unwraps None at lib.rs:273
    FlatScene flat = flatten(sc->world);
    rtc_scene_desc d = flat.desc();
    rtc_scene* scene = nullptr;
    int rc = rtc_scene_create(&d, device, &scene);
    if (rc) return fail(std::string("rtc_scene_create: ") + rtc_last_error()), rc;
    rtc_camera cam = sc->camera->to_rtc();
    rtc_opts o;
    std::memset(&o, 0, sizeof o);
    o.aa_samples = aa ? aa : 1; o.max_depth = 5; o.nranks = 1;
    std::vector<uint8_t> img((size_t)cam.hsize * cam.vsize * 3);
    rc = rtc_render_rgb8(scene, &cam, &o, img.data(), stats);
    std::string msg = rc ? rtc_last_error() : "";
    rtc_scene_destroy(scene);
    if (rc) return fail("rtc_render_rgb8: " + msg), rc;
    std::string out(out_path);
    auto ends = [&](const char* e) { std::string s(e); return out.size() >= s.size() && out.compare(out.size() - s.size(), s.size(), s) == 0; };
    if (ends(".png")) return rth_save_png(img.data(), cam.hsize, cam.vsize, out_path);
    if (ends(".ppm")) {
      std::vector<double> rgb(img.size());
      for (size_t i = 0; i < img.size(); i++) rgb[i] = img[i] / 255.0;  // quant(v/255*255) == v for v in 0..255
      return rth_save_ppm(rgb.data(), cam.hsize, cam.vsize, out_path);
    }
    return rth_save_p6(img.data(), cam.hsize, cam.vsize, out_path);
  } catch (const std::exception& e) { return fail(e.what()); }
}


}  // extern "C"
