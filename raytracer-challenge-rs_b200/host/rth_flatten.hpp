// rth_flatten.hpp — World → flat SoA tables (rtc_scene_desc) for the GPU.
//
// Pre-order walk of World.objects (world.rs:19) in insertion order; a Group's children follow it in
// order (group.rs:17), a Csg's left subtree then right subtree (csg.rs:47-48); every node gets a
// skip pointer so the device walk needs no stack. Transforms and materials are de-duplicated by
// exact bit pattern (all triangles of a mesh share one 96-byte inverse). Value-equality classes
// (`impl PartialEq for dyn Shape`, geometry/mod.rs:110-114, used by the containers walk of
// geometry/intersection.rs:63-90) are assigned by exact-bit hashing of everything the reference's
// derived PartialEq compares: kind, transform, material, bounding box, shadow flag and the
// kind-specific fields. Shapes that are equal only within the reference's 1e-5 epsilon but not
// bit-equal get distinct classes (documented deviation).
#pragma once
#include <cstring>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "rth_scene.hpp"

namespace rth {

struct FlatScene {
  std::vector<rtc_node> nodes;
  std::vector<double> xforms, boxes, limits, tri_geom, tri_normals;
  std::vector<rtc_material> materials;
  std::vector<rtc_pattern> patterns;
  std::vector<rtc_light> lights;
  size_t n_leaves = 0, n_groups = 0, n_csg = 0;

  rtc_scene_desc desc() const {
    rtc_scene_desc d;
    std::memset(&d, 0, sizeof d);
    d.abi_version = RTC_ABI_VERSION;
    d.n_nodes = (uint32_t)nodes.size();
    d.n_xforms = (uint32_t)(xforms.size() / 12);
    d.n_boxes = (uint32_t)(boxes.size() / 6);
    d.n_limits = (uint32_t)(limits.size() / 2);
    d.n_triangles = (uint32_t)(tri_geom.size() / 9);
    d.n_materials = (uint32_t)materials.size();
    d.n_patterns = (uint32_t)patterns.size();
    d.n_lights = (uint32_t)lights.size();
    d.nodes = nodes.data();
    d.xforms = xforms.data();
    d.boxes = boxes.data();
    d.limits = limits.data();
    d.tri_p1e1e2 = tri_geom.data();
    d.tri_normals = tri_normals.data();
    d.materials = materials.data();
    d.patterns = patterns.data();
    d.lights = lights.data();
    return d;
  }
  size_t bytes() const {
    return nodes.size() * sizeof(rtc_node) + (xforms.size() + boxes.size() + limits.size() + tri_geom.size() + tri_normals.size()) * 8 +
           materials.size() * sizeof(rtc_material) + patterns.size() * sizeof(rtc_pattern) + lights.size() * sizeof(rtc_light);
  }
};

class Flattener {
 public:
  FlatScene flatten(const World& w) {
    out_ = FlatScene();
    for (const PointLight& l : w.lights) {
      rtc_light rl;
      std::memcpy(rl.position, l.position, sizeof rl.position);
      std::memcpy(rl.intensity, l.intensity, sizeof rl.intensity);
      out_.lights.push_back(rl);
    }
    // Size the tables before the walk: every node is some aggregate's child (or a top-level object), so visiting
    // the aggregates alone (1/8 of a divided mesh) gives the exact node count; the triangle tables are bounded by
    // it. Growing 180 MB of vectors by doubling re-copies them twice and faults every page twice.
    size_t n_nodes = w.objects.size();
    for (const ShapeP& s : w.objects) n_nodes += count_below(*s);
    out_.nodes.reserve(n_nodes);
    class_reserve(n_nodes);
    class_reps_.reserve(n_nodes);
    if (n_nodes > 4096) {
      out_.tri_geom.reserve(9 * n_nodes);
      out_.tri_normals.reserve(9 * n_nodes);
    }
    for (const ShapeP& s : w.objects) emit(*s);
    return std::move(out_);
  }
  static size_t count_below(const Shape& s) {
    if (s.kind != RTC_GROUP && s.kind != RTC_CSG) return 0;
    size_t n = s.children.size();
    for (const ShapeP& c : s.children)
      if (c->kind == RTC_GROUP || c->kind == RTC_CSG) n += count_below(*c);
    return n;
  }

 private:
  FlatScene out_;
  std::unordered_map<const Xf*, int32_t> xf_by_ptr_;
  std::unordered_map<std::string, int32_t> xf_by_bits_, mat_by_bits_, pat_by_bits_;
  std::unordered_map<const Material*, int32_t> mat_by_ptr_;
  const Xf* last_xf_ = nullptr;
  const Material* last_mat_ = nullptr;
  int32_t last_xf_id_ = -1, last_mat_id_ = -1;

  static void put(std::string& k, const void* p, size_t n) { k.append((const char*)p, n); }

  int32_t xform_id(const XfP& x) {
    if (x.get() == last_xf_) return last_xf_id_;  // runs of triangles share one Xf
    const int32_t id = xform_lookup(x);
    last_xf_ = x.get();
    last_xf_id_ = id;
    return id;
  }
  int32_t xform_lookup(const XfP& x) {
    auto it = xf_by_ptr_.find(x.get());
    if (it != xf_by_ptr_.end()) return it->second;
    std::string key((const char*)x->inv.m, 12 * sizeof(double));
    auto jt = xf_by_bits_.find(key);
    int32_t id;
    if (jt != xf_by_bits_.end()) {
      id = jt->second;
    } else {
      id = (int32_t)(out_.xforms.size() / 12);
      out_.xforms.insert(out_.xforms.end(), x->inv.m, x->inv.m + 12);
      xf_by_bits_.emplace(std::move(key), id);
    }
    xf_by_ptr_.emplace(x.get(), id);
    return id;
  }
  int32_t pattern_id(const Pattern& p) {
    rtc_pattern rp;
    std::memset(&rp, 0, sizeof rp);
    rp.kind = p.kind;
    std::memcpy(rp.a, p.a, sizeof rp.a);
    std::memcpy(rp.b, p.b, sizeof rp.b);
    std::memcpy(rp.inverse, p.inverse.m, sizeof rp.inverse);
    std::string key((const char*)&rp, sizeof rp);
    auto it = pat_by_bits_.find(key);
    if (it != pat_by_bits_.end()) return it->second;
    int32_t id = (int32_t)out_.patterns.size();
    out_.patterns.push_back(rp);
    pat_by_bits_.emplace(std::move(key), id);
    return id;
  }
  int32_t material_id(const MaterialP& m) {
    if (m.get() == last_mat_) return last_mat_id_;
    const int32_t id = material_lookup(m);
    last_mat_ = m.get();
    last_mat_id_ = id;
    return id;
  }
  int32_t material_lookup(const MaterialP& m) {
    auto it = mat_by_ptr_.find(m.get());
    if (it != mat_by_ptr_.end()) return it->second;
    rtc_material rm;
    std::memset(&rm, 0, sizeof rm);
    std::memcpy(rm.color, m->color, sizeof rm.color);
    rm.ambient = m->ambient; rm.diffuse = m->diffuse; rm.specular = m->specular; rm.shininess = m->shininess;
    rm.reflective = m->reflective; rm.transparency = m->transparency; rm.refractive_index = m->refractive_index;
    rm.pattern = m->has_pattern ? pattern_id(m->pattern) : -1;
    std::string key((const char*)&rm, sizeof rm);
    auto jt = mat_by_bits_.find(key);
    int32_t id;
    if (jt != mat_by_bits_.end()) {
      id = jt->second;
    } else {
      id = (int32_t)out_.materials.size();
      out_.materials.push_back(rm);
      mat_by_bits_.emplace(std::move(key), id);
    }
    mat_by_ptr_.emplace(m.get(), id);
    return id;
  }
  int32_t box_id(const BBox& b) {
    int32_t id = (int32_t)(out_.boxes.size() / 6);
    const double v[6] = {b.min.x, b.min.y, b.min.z, b.max.x, b.max.y, b.max.z};
    out_.boxes.insert(out_.boxes.end(), v, v + 6);
    return id;
  }
  // exact-bit equality class of a leaf: an open-addressing table keyed by a hash of everything the
  // reference's PartialEq compares; a hash match is confirmed field by field against the earlier shape,
  // so classes are exact and a million-triangle mesh costs no per-leaf allocation.
  struct ClassSlot {      // 16 bytes: a 2 M-slot table for a million-triangle mesh is 32 MB, one cache miss per leaf
    uint64_t hash = 0;
    int32_t id = -1;      // -1: empty
    int32_t _pad = 0;
  };
  struct ClassRep { const Shape* s; int32_t xf, mat; };  // the first shape of every class, by class id
  std::vector<ClassSlot> class_slots_;
  std::vector<ClassRep> class_reps_;

  static uint64_t mix(uint64_t h, const void* p, size_t n) {
    const unsigned char* c = (const unsigned char*)p;
    for (; n >= 8; n -= 8, c += 8) {
      uint64_t w;
      std::memcpy(&w, c, 8);
      h = (h ^ w) * 0x9E3779B97F4A7C15ull;
      h ^= h >> 29;
    }
    if (n) {
      uint64_t w = 0;
      std::memcpy(&w, c, n);
      h = (h ^ w) * 0x9E3779B97F4A7C15ull;
      h ^= h >> 29;
    }
    return h;
  }
  static void box6(const Shape& s, double bb[6]) {
    bb[0] = s.box.min.x; bb[1] = s.box.min.y; bb[2] = s.box.min.z;
    bb[3] = s.box.max.x; bb[4] = s.box.max.y; bb[5] = s.box.max.z;
  }
  static bool same_bits(const void* a, const void* b, size_t n) { return std::memcmp(a, b, n) == 0; }
  static bool same_class(const Shape& a, int32_t axf, int32_t amat, const ClassRep& e) {
    const Shape& b = *e.s;
    if (a.kind != b.kind || axf != e.xf || amat != e.mat || a.shadow != b.shadow) return false;
    if (!same_bits(a.xf->m.m, b.xf->m.m, sizeof a.xf->m.m)) return false;
    double ba[6], bb[6];
    box6(a, ba);
    box6(b, bb);
    if (!same_bits(ba, bb, sizeof ba)) return false;
    if (a.kind == RTC_CYLINDER || a.kind == RTC_CONE) {
      if (!same_bits(&a.minimum, &b.minimum, 8) || !same_bits(&a.maximum, &b.maximum, 8) || a.closed != b.closed) return false;
    }
    if ((a.tri != nullptr) != (b.tri != nullptr)) return false;
    if (a.tri && !same_bits(a.tri.get(), b.tri.get(), sizeof(TriData))) return false;
    return true;
  }
  // table of at least 2 x `leaves` slots (a power of two), sized once: growing by doubling re-inserts every class
  // eleven times over for a million leaves
  void class_reserve(size_t leaves) {
    size_t want = 1024;
    while (want < 2 * leaves + 2) want <<= 1;
    if (want <= class_slots_.size()) return;
    std::vector<ClassSlot> old;
    old.swap(class_slots_);
    class_slots_.assign(want, ClassSlot());
    const size_t mask = want - 1;
    for (const ClassSlot& e : old) {
      if (e.id < 0) continue;
      size_t k = (size_t)e.hash & mask;
      while (class_slots_[k].id >= 0) k = (k + 1) & mask;
      class_slots_[k] = e;
    }
  }
  int32_t class_id(const Shape& s, int32_t xf, int32_t mat) {
    if ((class_reps_.size() + 1) * 2 > class_slots_.size()) class_reserve(class_reps_.size() * 2 + 512);
    const int32_t head[4] = {s.kind, xf, mat, s.shadow ? 1 : 0};
    uint64_t h = mix(0x243F6A8885A308D3ull, head, sizeof head);
    if (s.tri) {
      // a mesh is a million of these: the three vertices tell triangles apart (same_class compares everything)
      h = mix(h, &s.tri->p1, 3 * sizeof(Vec3));
    } else {
      h = mix(h, s.xf->m.m, sizeof s.xf->m.m);
      double bb[6];
      box6(s, bb);
      h = mix(h, bb, sizeof bb);
      if (s.kind == RTC_CYLINDER || s.kind == RTC_CONE) {
        const double lim[3] = {s.minimum, s.maximum, s.closed ? 1.0 : 0.0};
        h = mix(h, lim, sizeof lim);
      }
    }
    const size_t mask = class_slots_.size() - 1;
    size_t k = (size_t)h & mask;
    while (class_slots_[k].id >= 0) {
      if (class_slots_[k].hash == h && same_class(s, xf, mat, class_reps_[(size_t)class_slots_[k].id])) return class_slots_[k].id;
      k = (k + 1) & mask;
    }
    const int32_t id = (int32_t)class_reps_.size();
    class_slots_[k].hash = h;
    class_slots_[k].id = id;
    class_reps_.push_back(ClassRep{&s, xf, mat});
    return id;
  }

  void emit(const Shape& s) {
    const int32_t self = (int32_t)out_.nodes.size();
    rtc_node nd;
    nd.kind = s.kind; nd.skip = self + 1; nd.xform = -1; nd.material = -1; nd.aux = -1;
    nd.flags = 0; nd.right = -1; nd.eq_class = -1;
    out_.nodes.push_back(nd);
    if (s.kind == RTC_GROUP) {
      out_.n_groups++;
      nd.aux = box_id(s.box);
      // the children are separate heap objects (and a triangle's geometry one more): ask for them a few ahead
      const size_t nc = s.children.size();
      for (size_t i = 0; i < nc; i++) {
        if (i + 8 < nc) __builtin_prefetch(s.children[i + 8].get());
        if (i + 4 < nc && s.children[i + 4]->tri) __builtin_prefetch(s.children[i + 4]->tri.get());
        emit(*s.children[i]);
      }
      nd.skip = (int32_t)out_.nodes.size();
    } else if (s.kind == RTC_CSG) {
      out_.n_csg++;
      nd.aux = box_id(s.box);
      nd.xform = xform_id(s.xf);
      nd.flags = ((uint32_t)s.op & 3u) << 8;
      emit(*s.children[0]);
      nd.right = (int32_t)out_.nodes.size();
      emit(*s.children[1]);
      nd.skip = (int32_t)out_.nodes.size();
    } else {
      out_.n_leaves++;
      nd.xform = xform_id(s.xf);
      nd.material = material_id(s.material);
      if (s.shadow) nd.flags |= RTC_FLAG_SHADOW;
      if (s.kind == RTC_CYLINDER || s.kind == RTC_CONE) {
        if (s.closed) nd.flags |= RTC_FLAG_CLOSED;
        nd.aux = (int32_t)(out_.limits.size() / 2);
        out_.limits.push_back(s.minimum);
        out_.limits.push_back(s.maximum);
      } else if (s.kind == RTC_TRIANGLE || s.kind == RTC_SMOOTH_TRIANGLE) {
        const TriData& t = *s.tri;
        nd.aux = (int32_t)(out_.tri_geom.size() / 9);
        const double g[9] = {t.p1.x, t.p1.y, t.p1.z, t.e1.x, t.e1.y, t.e1.z, t.e2.x, t.e2.y, t.e2.z};
        out_.tri_geom.insert(out_.tri_geom.end(), g, g + 9);
        if (s.kind == RTC_TRIANGLE) {
          const double n[9] = {t.normal.x, t.normal.y, t.normal.z, 0, 0, 0, 0, 0, 0};
          out_.tri_normals.insert(out_.tri_normals.end(), n, n + 9);
        } else {
          const double n[9] = {t.n1.x, t.n1.y, t.n1.z, t.n2.x, t.n2.y, t.n2.z, t.n3.x, t.n3.y, t.n3.z};
          out_.tri_normals.insert(out_.tri_normals.end(), n, n + 9);
        }
      }
      nd.eq_class = class_id(s, nd.xform, nd.material);
    }
    out_.nodes[(size_t)self] = nd;
  }
};

inline FlatScene flatten(const World& w) {
  Flattener f;
  return f.flatten(w);
}

}  // namespace rth
