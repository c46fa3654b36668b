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
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "rth_scene.hpp"

namespace rth {

// resize() without the zero fill: the big tables of a merged scene are sized by one thread and WRITTEN by the workers,
// so their pages are first touched (faulted in) in parallel instead of by a 200 MB memset on the calling thread
template <class T>
struct default_init_allocator : std::allocator<T> {
  template <class U> struct rebind { typedef default_init_allocator<U> other; };
  default_init_allocator() = default;
  template <class U> default_init_allocator(const default_init_allocator<U>&) {}
  template <class U> void construct(U* p) { ::new ((void*)p) U; }
  template <class U, class... A> void construct(U* p, A&&... a) { ::new ((void*)p) U(std::forward<A>(a)...); }
};
template <class T> using raw_vector = std::vector<T, default_init_allocator<T>>;

struct FlatScene {
  raw_vector<rtc_node> nodes;
  std::vector<double> xforms;
  raw_vector<double> boxes, limits, tri_geom, tri_normals;
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
  // The same tables from worker threads (a million-triangle world is a million pointer-chased heap objects: 0.28 s on
  // one thread). The pre-order sequence is cut into JOBS — runs of whole subtrees of at most `grain` nodes — between
  // the few aggregates too big for one job, whose own nodes ("spine") the calling thread writes. Every job is flattened
  // by its own Flattener into local tables with local ids; the merge walks the pieces in pre-order, so transforms,
  // materials, patterns and equality classes get the ids of their first appearance in the walk — exactly what the
  // serial walk assigns (tests/test_cpu_parity.py holds the two byte-identical).
  FlatScene flatten_parallel(const World& w, int threads, size_t grain) {
    out_ = FlatScene();
    for (const PointLight& l : w.lights) {
      rtc_light rl;
      std::memcpy(rl.position, l.position, sizeof rl.position);
      std::memcpy(rl.intensity, l.intensity, sizeof rl.intensity);
      out_.lights.push_back(rl);
    }
    struct Job {
      std::vector<const Shape*> roots;
      size_t node_start = 0, n_nodes = 0;
      Flattener* f = nullptr;
      std::vector<int32_t> xf_map, mat_map, class_map;
      size_t box_off = 0, lim_off = 0, tri_off = 0;
    };
    // a piece is a job (a run of sibling subtrees handed to a worker) or one spine node; `end` / `right` name the piece
    // that follows the spine node's subtree / starts its right child (resolved to node indices once the jobs have run:
    // nobody knows a subtree's size before it has been walked, and counting first would chase every pointer twice)
    struct Piece { int job; const Shape* spine; size_t end, right; size_t node_index; };
    std::vector<Job> jobs;
    std::vector<Piece> pieces;
    struct Planner {
      std::vector<Job>& jobs; std::vector<Piece>& pieces; int max_depth; size_t range;
      void job_of(const ShapeP* first, size_t count) {
        jobs.emplace_back();
        for (size_t k = 0; k < count; k++) jobs.back().roots.push_back(first[k].get());
        pieces.push_back(Piece{(int)jobs.size() - 1, nullptr, 0, 0, 0});
      }
      // children [first, first + count) of one parent, `depth` levels below the world
      void plan(const ShapeP* first, size_t count, int depth) {
        size_t k = 0;
        while (k < count) {
          if (depth >= max_depth || count > 65536) {
            // below the opened top, or a wide parent: ranges of siblings, cut without looking at a single child (a
            // flat group of a million triangles must not be pointer-chased by the planner)
            const size_t run = std::min(range, count - k);
            job_of(first + k, run);
            k += run;
            continue;
          }
          const Shape& s = *first[k];
          const bool aggregate = s.kind == RTC_GROUP || s.kind == RTC_CSG;
          if (aggregate && !s.children.empty()) {
            const size_t pi = pieces.size();
            pieces.push_back(Piece{-1, &s, 0, 0, 0});
            if (s.kind == RTC_CSG) {
              plan(&s.children[0], 1, depth + 1);
              pieces[pi].right = pieces.size();
              plan(&s.children[1], 1, depth + 1);
            } else {
              plan(s.children.data(), s.children.size(), depth + 1);
            }
            pieces[pi].end = pieces.size();
            k++;
          } else {
            size_t run = 1;
            while (k + run < count && run < range && !(first[k + run]->kind == RTC_GROUP || first[k + run]->kind == RTC_CSG)) run++;
            job_of(first + k, run);
            k += run;
          }
        }
      }
    } planner{jobs, pieces, 0, grain};
    // open the top of the tree until there are about eight jobs per thread
    { int d = 1; while ((1 << d) < 8 * threads && d < 12) d++; planner.max_depth = d; }
    const auto t_begin = std::chrono::steady_clock::now();
    planner.plan(w.objects.data(), w.objects.size(), 0);
    const auto t_planned = std::chrono::steady_clock::now();
    // ---- the jobs, on `threads` threads (largest first)
    std::vector<Flattener> workers(jobs.size());
    std::vector<size_t> order(jobs.size());
    for (size_t i = 0; i < order.size(); i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](size_t a, size_t b) { return jobs[a].roots.size() > jobs[b].roots.size(); });
    std::atomic<size_t> next{0};
    auto work = [&]() {
      for (;;) {
        const size_t k = next.fetch_add(1);
        if (k >= order.size()) return;
        Job& j = jobs[order[k]];
        Flattener& f = workers[order[k]];
        j.f = &f;
        f.out_ = FlatScene();
        for (const Shape* r : j.roots) f.emit(*r);
        j.n_nodes = f.out_.nodes.size();
      }
    };
    {
      std::vector<std::thread> pool;
      const int nt = std::max(1, std::min(threads, (int)jobs.size()));
      for (int t = 1; t < nt; t++) pool.emplace_back(work);
      work();
      for (auto& t : pool) t.join();
    }
    const auto t_jobs = std::chrono::steady_clock::now();
    // ---- node indices of the pieces, now that every job knows its size
    size_t n_nodes = 0;
    for (Piece& pc : pieces) {
      pc.node_index = n_nodes;
      if (pc.job < 0) n_nodes += 1;
      else { jobs[(size_t)pc.job].node_start = n_nodes; n_nodes += jobs[(size_t)pc.job].n_nodes; }
    }
    auto index_of_piece = [&](size_t pi) { return pi < pieces.size() ? pieces[pi].node_index : n_nodes; };
    // ---- merge in pre-order: ids in order of first appearance, table offsets per job
    out_.nodes.resize(n_nodes);
    class_reserve(n_nodes);
    class_reps_.reserve(n_nodes);
    class_hash_.reserve(n_nodes);
    size_t n_tri = 0, n_lim = 0, n_box = 0;
    for (const Piece& pc : pieces) {
      if (pc.job < 0) n_box++;
      else { const FlatScene& l = jobs[(size_t)pc.job].f->out_; n_tri += l.tri_geom.size() / 9; n_lim += l.limits.size() / 2; n_box += l.boxes.size() / 6; }
    }
    out_.tri_geom.resize(9 * n_tri);
    out_.tri_normals.resize(9 * n_tri);
    out_.limits.resize(2 * n_lim);
    out_.boxes.reserve(6 * n_box);
    size_t tri_at = 0, lim_at = 0;
    const auto t_alloc = std::chrono::steady_clock::now();
    if (std::getenv("RTC_TIMING")) std::fprintf(stderr, "flatten_parallel: merge tables sized in %.1f ms\n", std::chrono::duration<double, std::milli>(t_alloc - t_jobs).count());
    for (const Piece& pc : pieces) {
      if (pc.job < 0) {
        const Shape& s = *pc.spine;
        rtc_node nd;
        nd.kind = s.kind; nd.skip = (int32_t)index_of_piece(pc.end); nd.xform = -1; nd.material = -1; nd.aux = box_id(s.box);
        nd.flags = 0; nd.right = -1; nd.eq_class = -1;
        if (s.kind == RTC_CSG) {
          out_.n_csg++;
          nd.xform = xform_id(s.xf);
          nd.flags = ((uint32_t)s.op & 3u) << 8;
          nd.right = (int32_t)index_of_piece(pc.right);
        } else {
          out_.n_groups++;
        }
        out_.nodes[pc.node_index] = nd;
        continue;
      }
      Job& j = jobs[(size_t)pc.job];
      Flattener& f = *j.f;
      const FlatScene& l = f.out_;
      j.box_off = out_.boxes.size() / 6;
      out_.boxes.insert(out_.boxes.end(), l.boxes.begin(), l.boxes.end());
      j.tri_off = tri_at;  tri_at += l.tri_geom.size() / 9;   // (the triangle and limit tables are copied by the workers below)
      j.lim_off = lim_at;  lim_at += l.limits.size() / 2;
      j.xf_map.resize(l.xforms.size() / 12);
      for (size_t k = 0; k < j.xf_map.size(); k++) j.xf_map[k] = xform_bits_id(&l.xforms[12 * k]);
      std::vector<int32_t> pat_map(l.patterns.size());
      for (size_t k = 0; k < pat_map.size(); k++) pat_map[k] = pattern_bits_id(l.patterns[k]);
      j.mat_map.resize(l.materials.size());
      for (size_t k = 0; k < j.mat_map.size(); k++) {
        rtc_material rm = l.materials[k];
        if (rm.pattern >= 0) rm.pattern = pat_map[(size_t)rm.pattern];
        j.mat_map[k] = material_bits_id(rm);
      }
      j.class_map.resize(f.class_reps_.size());
      const size_t mask = class_slots_.size() - 1;
      for (size_t k = 0; k < j.class_map.size(); k++) {
        if (k + 8 < j.class_map.size()) __builtin_prefetch(&class_slots_[(size_t)f.class_hash_[k + 8] & mask]);
        const ClassRep& r = f.class_reps_[k];
        j.class_map[k] = class_insert(f.class_hash_[k], *r.s, j.xf_map[(size_t)r.xf], j.mat_map[(size_t)r.mat]);
      }
      out_.n_leaves += l.n_leaves; out_.n_groups += l.n_groups; out_.n_csg += l.n_csg;
    }
    const auto t_merged = std::chrono::steady_clock::now();
    // ---- the nodes and the big tables, remapped / copied on all threads
    next = 0;
    auto fix = [&]() {
      for (;;) {
        const size_t k = next.fetch_add(1);
        if (k >= order.size()) return;
        const Job& j = jobs[order[k]];
        const FlatScene& l = j.f->out_;
        const int32_t base = (int32_t)j.node_start;
        for (size_t q = 0; q < l.nodes.size(); q++) {
          rtc_node nd = l.nodes[q];
          nd.skip += base;
          if (nd.right >= 0) nd.right += base;
          if (nd.kind == RTC_GROUP || nd.kind == RTC_CSG) nd.aux += (int32_t)j.box_off;
          else if (nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE) nd.aux += (int32_t)j.tri_off;
          else if (nd.kind == RTC_CYLINDER || nd.kind == RTC_CONE) nd.aux += (int32_t)j.lim_off;
          if (nd.xform >= 0) nd.xform = j.xf_map[(size_t)nd.xform];
          if (nd.material >= 0) nd.material = j.mat_map[(size_t)nd.material];
          if (nd.eq_class >= 0) nd.eq_class = j.class_map[(size_t)nd.eq_class];
          out_.nodes[j.node_start + q] = nd;
        }
        std::copy(l.tri_geom.begin(), l.tri_geom.end(), out_.tri_geom.begin() + 9 * j.tri_off);
        std::copy(l.tri_normals.begin(), l.tri_normals.end(), out_.tri_normals.begin() + 9 * j.tri_off);
        std::copy(l.limits.begin(), l.limits.end(), out_.limits.begin() + 2 * j.lim_off);
      }
    };
    {
      std::vector<std::thread> pool;
      const int nt = std::max(1, std::min(threads, (int)jobs.size()));
      for (int t = 1; t < nt; t++) pool.emplace_back(fix);
      fix();
      for (auto& t : pool) t.join();
    }
    if (std::getenv("RTC_TIMING")) {
      auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count();
      };
      std::fprintf(stderr, "flatten_parallel: %zu nodes, %zu jobs + %zu spine nodes on %d threads: plan %.1f ms, jobs %.1f ms, merge %.1f ms, remap + copy %.1f ms\n",
                   n_nodes, jobs.size(), pieces.size() - jobs.size(), threads, ms(t_begin, t_planned), ms(t_planned, t_jobs),
                   ms(t_jobs, t_merged), ms(t_merged, std::chrono::steady_clock::now()));
    }
    return std::move(out_);
  }
  // does the world have at least `want` nodes? (stops counting there: a small world answers in microseconds, a big
  // one after `want` children instead of after all of them)
  static bool at_least(const World& w, size_t want) {
    size_t n = w.objects.size();
    if (n >= want) return true;
    std::vector<const Shape*> todo;
    for (const ShapeP& s : w.objects)
      if (s->kind == RTC_GROUP || s->kind == RTC_CSG) todo.push_back(s.get());
    while (!todo.empty()) {
      const Shape* s = todo.back();
      todo.pop_back();
      n += s->children.size();
      if (n >= want) return true;
      for (const ShapeP& c : s->children)
        if (c->kind == RTC_GROUP || c->kind == RTC_CSG) todo.push_back(c.get());
    }
    return false;
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
  int32_t xform_bits_id(const double* inv12) {
    std::string key((const char*)inv12, 12 * sizeof(double));
    auto jt = xf_by_bits_.find(key);
    if (jt != xf_by_bits_.end()) return jt->second;
    const int32_t id = (int32_t)(out_.xforms.size() / 12);
    out_.xforms.insert(out_.xforms.end(), inv12, inv12 + 12);
    xf_by_bits_.emplace(std::move(key), id);
    return id;
  }
  int32_t material_bits_id(const rtc_material& rm) {
    std::string key((const char*)&rm, sizeof rm);
    auto jt = mat_by_bits_.find(key);
    if (jt != mat_by_bits_.end()) return jt->second;
    const int32_t id = (int32_t)out_.materials.size();
    out_.materials.push_back(rm);
    mat_by_bits_.emplace(std::move(key), id);
    return id;
  }
  int32_t pattern_id(const Pattern& p) {
    rtc_pattern rp;
    std::memset(&rp, 0, sizeof rp);
    rp.kind = p.kind;
    std::memcpy(rp.a, p.a, sizeof rp.a);
    std::memcpy(rp.b, p.b, sizeof rp.b);
    std::memcpy(rp.inverse, p.inverse.m, sizeof rp.inverse);
    return pattern_bits_id(rp);
  }
  int32_t pattern_bits_id(const rtc_pattern& rp) {
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
  std::vector<uint64_t> class_hash_;   // by class id

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
    const uint64_t h = class_hash(s);
    return class_insert(h, s, xf, mat);
  }
  // hash of what tells leaves apart WITHOUT the transform / material ids (same_class compares those): a hash taken by
  // a worker that only knows its own local ids stays valid when the classes are merged under the global ids
  static uint64_t class_hash(const Shape& s) {
    const int32_t head[2] = {s.kind, s.shadow ? 1 : 0};
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
    return h;
  }
  int32_t class_insert(uint64_t h, const Shape& s, int32_t xf, int32_t mat) {
    if ((class_reps_.size() + 1) * 2 > class_slots_.size()) class_reserve(class_reps_.size() * 2 + 512);
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
    class_hash_.push_back(h);
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

// RTH_FLATTEN_THREADS (default: the host's threads, at most 16; 1 = the serial walk), RTH_FLATTEN_MIN_NODES (worlds
// below this many nodes are walked serially; default 65536), RTH_FLATTEN_GRAIN (siblings per job; tests)
inline FlatScene flatten(const World& w) {
  Flattener f;
  unsigned hc = std::thread::hardware_concurrency();
  int threads = hc == 0 ? 1 : (hc > 16 ? 16 : (int)hc);
  if (const char* v = std::getenv("RTH_FLATTEN_THREADS")) threads = std::max(1, std::min(64, std::atoi(v)));
  size_t min_nodes = 65536;
  if (const char* v = std::getenv("RTH_FLATTEN_MIN_NODES")) min_nodes = (size_t)std::max(0, std::atoi(v));
  if (threads > 1 && Flattener::at_least(w, min_nodes)) {
    size_t range = 4096;   // siblings per job
    if (const char* v = std::getenv("RTH_FLATTEN_GRAIN")) range = (size_t)std::max(1, std::atoi(v));
    return f.flatten_parallel(w, threads, range);
  }
  return f.flatten(w);
}

}  // namespace rth
