// rt_accel_build.hpp — HOST-side construction of the traversal accelerator from a flattened world
// (rtc_scene_desc). Pure C++ (no CUDA), used by rtc_scene_create and by the CPU-only test build of
// the kernel source.
//
// The reference intersects a ray with EVERY top-level object (world.rs:31-38) and prunes only with
// the boxes of its own Group tree (group.rs:49-58), whose quality collapses for rotated meshes
// (the AABB of every re-parented child is re-derived from its AABB, geometry/mod.rs:74-85, so
// Group::divide stops separating anything after a few levels). The accelerator keeps the
// reference's RESULT and replaces its SEARCH:
//   * prims    every leaf that is not inside a CSG, and every outermost CSG node (evaluated by the
//              in-order subtree walk, because its filter needs the reference's order);
//   * always   prims that are tested for every ray: unbounded ones (planes, open cylinders/cones),
//              kinds whose reference intersector can report points off the exact surface (cube,
//              cylinder, cone: EPSILON-parallel slabs, cone.rs:106) and CSGs;
//   * bvh      a binned-SAH binary BVH over the remaining prims (triangles, spheres) with our own
//              tight world-space boxes, traversed front to back with t-range culling;
//   * parent   for every node the innermost enclosing reference Group, so that a candidate hit is
//              accepted only if the ray passes BoundingBox::intersects of every enclosing Group —
//              exactly the condition under which the reference would have reached that leaf.
// Ties in t resolve to the lower pre-order node index (= first in the reference's traversal order).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <limits>
#include <thread>
#include <vector>

#include "../../include/rtc.h"

namespace rtk {

// Our own boxes only have to be CONSERVATIVE (the reference never sees them), so they are stored in single precision,
// rounded outward: a node is 64 bytes — both children's boxes and references in four 16-byte loads, two nodes to a
// 128-byte line — and the slab test runs on the FP32 pipe (rt_accel.cuh: search). -DRTK_BVH_F32=0 keeps the 128-byte
// double-precision node of rounds 1-2 for A/B runs.
#if !defined(RTK_BVH_F32)
#define RTK_BVH_F32 1
#endif
#if RTK_BVH_F32
typedef float bvh_real;
#else
typedef double bvh_real;
#endif

struct BvhNode {
  bvh_real lo0[3], hi0[3];
  bvh_real lo1[3], hi1[3];
  int32_t c0, n0;    // n > 0: leaf, prims [c, c+n) of bvh_prims; n == 0: inner node index c; n < 0: empty
  int32_t c1, n1;
#if !RTK_BVH_F32
  int32_t _pad[4];
#endif
};

// What the kernels traverse (-DRTK_BVH_WIDE=0: the binary node itself): the binary tree collapsed to FOUR children per
// node — half the dependent node fetches per ray, which is what a ray's search waits on. One node is one 128-byte line:
// the children's boxes axis by axis (a 16-byte load per plane family) and one reference each.
#if !defined(RTK_BVH_WIDE)
#define RTK_BVH_WIDE 0
#endif
#if RTK_BVH_WIDE && !RTK_BVH_F32
#error "RTK_BVH_WIDE needs RTK_BVH_F32"
#endif
constexpr int32_t BVH_EMPTY = -0x7fffffff - 1;
constexpr int WIDE_STACK_MAX = 64;   // = BVH_STACK (rt_accel.cuh)
struct BvhNode4 {
  float lox[4], loy[4], loz[4], hix[4], hiy[4], hiz[4];
  int32_t ref[4];    // >= 0: inner node; < 0: leaf ~(first << 3 | count), prims [first, first + count); BVH_EMPTY: no child
  int32_t _pad[4];
};
static_assert(sizeof(BvhNode4) == 128, "record layout");

// the nearest bvh_real at or below / at or above x
#if defined(__CUDACC__)
__host__ __device__
#endif
inline bvh_real bvh_below(double x) {
#if RTK_BVH_F32
  float f = (float)x;
  if ((double)f > x) f = nextafterf(f, -HUGE_VALF);
  return f;
#else
  return x;
#endif
}
#if defined(__CUDACC__)
__host__ __device__
#endif
inline bvh_real bvh_above(double x) {
#if RTK_BVH_F32
  float f = (float)x;
  if ((double)f < x) f = nextafterf(f, HUGE_VALF);
  return f;
#else
  return x;
#endif
}

struct PrimRec {     // 128 bytes: everything the every-ray loop needs about one prim, in one place
  int32_t node, aux, parent;
  uint32_t kind_flags;   // kind << 24 | flags — the whole header is one 16-byte load
  int32_t _pad[4];
  double inv[12];    // rows 0..2 of the leaf's transform_inverse (unused for a CSG prim)
};

struct LeafRec {     // 96 bytes: one BVH prim (any bounded leaf shape), in BVH leaf order
  int32_t node, xform, parent;
  uint32_t kind_flags;   // kind << 24 | flags
  double g[9];           // triangles: p1, e1, e2 (triangle.rs:14-21); other kinds: unused
  int32_t aux, _pad;     // rtc_node.aux (cylinder: index into limits)
};

#if RTK_BVH_WIDE
typedef BvhNode4 TravNode;
#else
typedef BvhNode TravNode;
#endif

// largest |coordinate| per axis over the children of the root = over the whole tree (inf for an all-covering box)
inline void bvh_extent(const BvhNode& root, double ext[3]) {
  for (int a = 0; a < 3; a++) {
    ext[a] = 0.0;
    if (root.n0 >= 0) ext[a] = std::fmax(ext[a], std::fmax(std::fabs((double)root.lo0[a]), std::fabs((double)root.hi0[a])));
    if (root.n1 >= 0) ext[a] = std::fmax(ext[a], std::fmax(std::fabs((double)root.lo1[a]), std::fabs((double)root.hi1[a])));
  }
}
inline void bvh_extent(const BvhNode4& root, double ext[3]) {
  const float* lo[3] = {root.lox, root.loy, root.loz};
  const float* hi[3] = {root.hix, root.hiy, root.hiz};
  for (int a = 0; a < 3; a++) {
    ext[a] = 0.0;
    for (int k = 0; k < 4; k++)
      if (root.ref[k] != BVH_EMPTY) ext[a] = std::fmax(ext[a], std::fmax(std::fabs((double)lo[a][k]), std::fabs((double)hi[a][k])));
  }
}

static_assert(sizeof(PrimRec) == 128 && sizeof(LeafRec) == 96 && sizeof(BvhNode) == (RTK_BVH_F32 ? 64 : 128), "record layout");

// World-space bounding sphere of an every-ray sphere prim (the first SPH_CULL_MAX of them; there are at
// most BVH_MIN_PRIMS - 1 unless a transform is degenerate). Travels inside SceneView, i.e. in the
// kernel's parameter (constant) bank. r2 = +inf disables the cull for that prim.
constexpr int SPH_CULL_MAX = 8;
struct SphCull {
  double cx, cy, cz;  // centre = forward transform of the object-space origin
  double r2;          // ||M^T M||_inf * (1 + 1e-4) >= sigma_max(M)^2 of the forward 3x3 M, padded
  double far2;        // the cull is only trusted for |o - c|^2 < far2 (rounding of either side stays far below the pad)
  double _pad;
};

// Cube::check_axis (cube.rs:30-48) treats an object-space direction component below EPSILON as parallel: the slab
// becomes "all of the line or none of it" by the ORIGIN's coordinate alone, so the reference reports hits whose point
// lies outside the cube by up to EPSILON * t. No box bounds those, so a ray that is near-parallel to a face family of
// some cube in the BVH must not be culled away from cubes. The face families (unit normals u, and the largest
// theta = EPSILON / |row| of the cubes sharing u: |row . d| < EPSILON  <=>  |u . d| < theta) travel in SceneView; a ray
// with (u . d)^2 < th2 + 1e-24 d.d for some family walks the tree without any culling (rt_accel.cuh). More than
// CUBE_DIR_MAX families, or more than CUBE_BVH_MAX_PRIMS prims in the tree (an unculled walk of a big mesh is a cliff):
// cubes stay every-ray prims.
constexpr int CUBE_DIR_MAX = 6;
constexpr size_t CUBE_BVH_MAX_PRIMS = 65536;
struct CubeDir {
  double ux, uy, uz;
  double th2;  // 2 * theta^2 * (1 + 1e-6): (a + b)^2 <= 2 a^2 + 2 b^2 covers the clustering tolerance of u
};

struct AccelBuild {
  std::vector<LeafRec> leaf_recs;  // one per BVH prim: prims[n_always + k] ↔ leaf_recs[k]
  std::vector<PrimRec> recs;       // one per always-prim, same order as prims[0, n_always)
  std::vector<int32_t> parent;     // per node: innermost enclosing Group node (outside CSGs), or -1
  std::vector<int32_t> prims;      // node indices: [0, n_always) tested for every ray (ascending), then BVH leaf order
  int32_t n_always = 0;
  int32_t n_sph = 0, n_pl = 0;     // the always-prefix is ordered spheres, planes, everything else
  std::vector<BvhNode> bvh;        // the binary tree; node 0 = root holder (empty vector: no BVH)
  std::vector<BvhNode4> wide;      // RTK_BVH_WIDE: `bvh` collapsed to four children per node (collapse_wide), node 0 = root
  int32_t max_depth = 0;
  int32_t wide_stack_need = 0;
#if RTK_BVH_WIDE
  const TravNode* trav_data() const { return wide.data(); }
  size_t trav_count() const { return wide.size(); }
#else
  const TravNode* trav_data() const { return bvh.data(); }
  size_t trav_count() const { return bvh.size(); }
#endif
  SphCull sph_cull[SPH_CULL_MAX];
  CubeDir cube_dirs[CUBE_DIR_MAX] = {};
  int32_t n_cube_dirs = 0;
  // device build (rt_accel_device.cuh): the host did the walk (every-ray prims, group parents) and left the tree to
  // the GPU — `cand` are the BVH prims' node indices, `fwd` the forward matrices (12 per transform, NaN when singular)
  bool deferred = false;
  bool cand_generic = false;  // some BVH prim is not a triangle or a sphere
  std::vector<int32_t> cand;
  std::vector<double> fwd;
};

namespace accel_detail {

struct Box {
  double lo[3], hi[3];
  void reset() {
    for (int a = 0; a < 3; a++) { lo[a] = std::numeric_limits<double>::infinity(); hi[a] = -lo[a]; }
  }
  void add(const double p[3]) {
    for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); }
  }
  void add(const Box& b) {
    for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); }
  }
  bool finite() const {
    for (int a = 0; a < 3; a++)
      if (!std::isfinite(lo[a]) || !std::isfinite(hi[a]) || lo[a] > hi[a]) return false;
    return true;
  }
  double area() const {
    double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return 2.0 * (dx * dy + dy * dz + dz * dx);
  }
};

// Forward 3x4 affine from the stored inverse (rows 0..2). Returns false if singular/non-finite.
inline bool invert_affine(const double* inv, double fwd[12]) {
  const double a = inv[0], b = inv[1], c = inv[2], d = inv[4], e = inv[5], f = inv[6], g = inv[8], h = inv[9], i = inv[10];
  const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
  if (!(std::fabs(det) > 0.0) || !std::isfinite(det)) return false;
  const double r = 1.0 / det;
  double m[9] = {(e * i - f * h) * r, (c * h - b * i) * r, (b * f - c * e) * r,
                 (f * g - d * i) * r, (a * i - c * g) * r, (c * d - a * f) * r,
                 (d * h - e * g) * r, (b * g - a * h) * r, (a * e - b * d) * r};
  for (int row = 0; row < 3; row++) {
    for (int col = 0; col < 3; col++) fwd[4 * row + col] = m[3 * row + col];
    fwd[4 * row + 3] = -(m[3 * row] * inv[3] + m[3 * row + 1] * inv[7] + m[3 * row + 2] * inv[11]);
  }
  for (int k = 0; k < 12; k++)
    if (!std::isfinite(fwd[k])) return false;
  return true;
}

inline void xf(const double* m, const double p[3], double out[3]) {
  for (int r = 0; r < 3; r++) out[r] = m[4 * r] * p[0] + m[4 * r + 1] * p[1] + m[4 * r + 2] * p[2] + m[4 * r + 3];
}

// Our own world-space bound of a BVH-eligible leaf (triangle, smooth triangle, sphere).
// Forward matrices of the scene's (de-duplicated) inverse transforms, inverted on first use: a mesh's
// million triangles share one.
struct FwdCache {
  std::vector<double> m;
  std::vector<signed char> state;  // 0 unknown, 1 ok, -1 singular
  explicit FwdCache(size_t n_xforms) : m(12 * n_xforms), state(n_xforms, 0) {}
  const double* get(const rtc_scene_desc* d, int32_t xf) {
    signed char& st = state[(size_t)xf];
    if (st == 0) st = invert_affine(d->xforms + 12 * (size_t)xf, &m[12 * (size_t)xf]) ? 1 : -1;
    return st > 0 ? &m[12 * (size_t)xf] : nullptr;
  }
};

inline bool leaf_world_box(const rtc_scene_desc* d, const rtc_node& nd, Box& out, FwdCache& cache, bool cubes_ok = true) {
  if (nd.kind == RTC_CUBE && !cubes_ok) return false;
  const double* fwd = cache.get(d, nd.xform);
  if (!fwd) return false;
  out.reset();
  double w[3];
  if (nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE) {
    const double* g = d->tri_p1e1e2 + 9 * (size_t)nd.aux;
    double p1[3] = {g[0], g[1], g[2]}, p2[3] = {g[0] + g[3], g[1] + g[4], g[2] + g[5]},
           p3[3] = {g[0] + g[6], g[1] + g[7], g[2] + g[8]};
    xf(fwd, p1, w); out.add(w);
    xf(fwd, p2, w); out.add(w);
    xf(fwd, p3, w); out.add(w);
  } else if (nd.kind == RTC_SPHERE || nd.kind == RTC_CUBE || nd.kind == RTC_CYLINDER) {
    // object-space extent: the unit box; a cylinder is radius 1 between its limits (cylinder.rs:27-40)
    // — an unbounded one stays an every-ray prim. A cube's EPSILON-parallel hits lie outside its box: see CubeDir. Cones are never admitted: for a ray parallel to a
    // cone half (|a| < EPSILON) the reference emits t = -c/2·b without the y-limit test (cone.rs:100-108),
    // a hit that can lie anywhere along the ray, so no box bounds a cone's intersections.
    double r = 1.0, y0 = -1.0, y1 = 1.0;
    if (nd.kind == RTC_CYLINDER) {
      y0 = d->limits[2 * (size_t)nd.aux];
      y1 = d->limits[2 * (size_t)nd.aux + 1];
      if (!std::isfinite(y0) || !std::isfinite(y1) || !(y0 <= y1)) return false;
    }
    for (int k = 0; k < 8; k++) {
      double p[3] = {(k & 1) ? r : -r, (k & 2) ? y1 : y0, (k & 4) ? r : -r};
      xf(fwd, p, w); out.add(w);
    }
  } else {
    return false;
  }
  if (!out.finite()) return false;
  // guard band: the reference's t comes from object-space arithmetic; 1e-7 relative dwarfs its rounding
  for (int a = 0; a < 3; a++) {
    double pad = 1e-7 * (1.0 + std::max(std::fabs(out.lo[a]), std::fabs(out.hi[a])) + (out.hi[a] - out.lo[a]));
    out.lo[a] -= pad;
    out.hi[a] += pad;
  }
  return out.finite();
}

}  // namespace accel_detail

namespace accel_detail {

struct Prim {
  Box box;
  double c[3];
  int32_t node;
  int32_t generic;  // 1: goes through the generic leaf test (cube, cylinder, cone); 0: triangle / sphere
};

constexpr int SAH_BINS = 16;
constexpr int LEAF_TRI = 3;            // triangles / spheres per leaf
constexpr double SAH_W_GENERIC = 3.0;  // cost of a generic leaf test (cube / cylinder / cone), in triangle tests

// The binary tree `b2` (node 0: holder whose first child is the whole tree) as four-child nodes. A wide node starts as
// the two children of a binary node; while it has room, its inner child with the largest surface area is replaced by
// that child's own two children. `max_children` = 2 keeps the binary shape (one wide node per binary node).
// Returns the traversal stack the tree can need: along any root-to-leaf path, every node leaves at most (children - 1)
// siblings on the stack.
inline int collapse_wide(const std::vector<BvhNode>& b2, std::vector<BvhNode4>& out, int max_children) {
  struct Ch { float lo[3], hi[3]; int32_t c, n; };
  auto child = [](const BvhNode& nd, int which) {
    Ch c;
    for (int a = 0; a < 3; a++) { c.lo[a] = which ? nd.lo1[a] : nd.lo0[a]; c.hi[a] = which ? nd.hi1[a] : nd.hi0[a]; }
    c.c = which ? nd.c1 : nd.c0;
    c.n = which ? nd.n1 : nd.n0;
    return c;
  };
  auto empty4 = [] {
    BvhNode4 w;
    for (int k = 0; k < 4; k++) {
      w.lox[k] = w.loy[k] = w.loz[k] = HUGE_VALF;
      w.hix[k] = w.hiy[k] = w.hiz[k] = -HUGE_VALF;
      w.ref[k] = BVH_EMPTY;
      w._pad[k] = 0;
    }
    return w;
  };
  out.clear();
  if (b2.empty()) return 0;
  struct Todo { int32_t src, dst, above; };  // binary node, wide node, stack entries the path to it can hold
  std::vector<Todo> todo;
  out.push_back(empty4());
  todo.push_back({0, 0, 0});
  int need = 0;
  while (!todo.empty()) {
    const Todo t = todo.back();
    todo.pop_back();
    Ch list[4];
    int m = 0;
    for (int which = 0; which < 2; which++)
      if ((which ? b2[(size_t)t.src].n1 : b2[(size_t)t.src].n0) >= 0) list[m++] = child(b2[(size_t)t.src], which);
    // (a holder's single inner child is always opened: a one-child node would only cost a fetch)
    while (m < max_children || m == 1) {
      int best = -1;
      double ba = -1.0;
      for (int k = 0; k < m; k++) {
        if (list[k].n != 0) continue;
        const double dx = (double)list[k].hi[0] - list[k].lo[0], dy = (double)list[k].hi[1] - list[k].lo[1], dz = (double)list[k].hi[2] - list[k].lo[2];
        const double area = dx * dy + dy * dz + dz * dx;
        if (best < 0 || area > ba) { best = k; ba = area; }
      }
      if (best < 0) break;
      const BvhNode& nd = b2[(size_t)list[best].c];
      const Ch a = child(nd, 0), b = child(nd, 1);
      const bool ha = a.n >= 0, hb = b.n >= 0;
      if (ha && hb) { list[best] = a; list[m++] = b; }
      else if (ha || hb) list[best] = ha ? a : b;
      else { list[best] = list[m - 1]; m--; }
      if (m == 0) break;
    }
    BvhNode4 w = empty4();
    const int below = t.above + (m > 0 ? m - 1 : 0);
    need = std::max(need, below);
    for (int k = 0; k < m; k++) {
      w.lox[k] = list[k].lo[0]; w.loy[k] = list[k].lo[1]; w.loz[k] = list[k].lo[2];
      w.hix[k] = list[k].hi[0]; w.hiy[k] = list[k].hi[1]; w.hiz[k] = list[k].hi[2];
      if (list[k].n > 0) {
        w.ref[k] = ~((list[k].c << 3) | list[k].n);
      } else {
        const int32_t id = (int32_t)out.size();
        out.push_back(empty4());
        w.ref[k] = id;
        todo.push_back({list[k].c, id, below});
      }
    }
    out[(size_t)t.dst] = w;
  }
  return need;
}

struct Builder {
  std::vector<Prim>& prims;
  AccelBuild& out;
  const rtc_scene_desc* desc;
  int threads = 1;                         // worker threads for large ranges / independent subtrees
  int PAR_MIN = 1 << 16;                   // ranges below this are binned on the calling thread
  int PAR_SUBTREES_MIN = 8192;             // trees below this are built on the calling thread alone
  std::vector<Prim> scratch;               // the parallel partition's other buffer
  explicit Builder(std::vector<Prim>& p, AccelBuild& o, const rtc_scene_desc* d) : prims(p), out(o), desc(d) {
    unsigned hc = std::thread::hardware_concurrency();
    threads = hc == 0 ? 1 : (hc > 16 ? 16 : (int)hc);
    if (const char* v = std::getenv("RTC_BVH_THREADS")) threads = std::max(1, std::min(64, std::atoi(v)));
    if (const char* v = std::getenv("RTC_BVH_PAR_MIN")) {  // tests: push small meshes through the threaded path
      PAR_MIN = std::max(8, std::atoi(v));
      PAR_SUBTREES_MIN = 4 * PAR_MIN;
      if (threads < 4) threads = 4;
    }
  }

  // fn(chunk index, begin, end) over [b, e) cut into nchunk pieces; piece 0 runs on the caller
  template <class F>
  static void parallel_chunks(int b, int e, int nchunk, F&& fn) {
    if (nchunk <= 1) { fn(0, b, e); return; }
    std::vector<std::thread> pool;
    const long long n = e - b;
    for (int ci = 1; ci < nchunk; ci++)
      pool.emplace_back([&, ci] { fn(ci, b + (int)(n * ci / nchunk), b + (int)(n * (ci + 1) / nchunk)); });
    fn(0, b, b + (int)(n / nchunk));
    for (auto& t : pool) t.join();
  }

  // split() for at most eight prims, which is most of the calls: every cut of the centroid order on every axis, exactly
  // (sixteen bins over so few prims nearly always separate them all, so this is the binned answer without zeroing and
  // scanning 48 bins), on this frame's stack. Also where the tree stops.
  int split_small(int b, int e, const Box& bounds, int leaf_max, Box* child_box, bool* have_boxes) {
    const int n = e - b;
    int n_generic = 0;
    for (int k = b; k < e; k++) n_generic += prims[(size_t)k].generic;
    // When to stop (profiles/r02ag_leafmax.log, r02ah_sah_stop.log: leaf sizes 1-6 and a cost-based stop on the three BVH
    // workloads): up to LEAF_TRI triangles / spheres make a leaf — their tests are cheap next to the node visits a
    // further split adds — while the kinds that go through the generic leaf test (cube, cylinder, cone) get a leaf each
    // (cover.yml: 3.47 -> 2.88 ms). RTC_SAH_KT=<cost of a node visit in leaf tests> switches to the SAH's own stop rule
    // instead (experiments; it wins on the million-triangle mesh and loses a third on the teapot).
    static const double kt = [] { const char* e = std::getenv("RTC_SAH_KT"); return e ? std::atof(e) : -1.0; }();
    if (leaf_max == 0 && n <= 7 && kt < 0.0 && n_generic == 0 && n <= LEAF_TRI) return -1;
    int best_axis = -1, best_pos = -1;
    double best_cost = std::numeric_limits<double>::infinity();
    int order[3][8];
    for (int a = 0; a < 3; a++) {
      int* idx = order[a];
      for (int k = 0; k < n; k++) {  // insertion sort by centroid, stable
        int j = k;
        while (j > 0 && prims[(size_t)(b + idx[j - 1])].c[a] > prims[(size_t)(b + k)].c[a]) { idx[j] = idx[j - 1]; j--; }
        idx[j] = k;
      }
      if (!(prims[(size_t)(b + idx[n - 1])].c[a] > prims[(size_t)(b + idx[0])].c[a])) continue;  // no extent on this axis
      double right_area[8];
      Box acc;
      acc.reset();
      for (int k = n - 1; k > 0; k--) {
        acc.add(prims[(size_t)(b + idx[k])].box);
        right_area[k] = acc.area();
      }
      acc.reset();
      for (int k = 0; k + 1 < n; k++) {
        acc.add(prims[(size_t)(b + idx[k])].box);
        const double cost = acc.area() * (k + 1) + right_area[k + 1] * (n - k - 1);
        if (cost < best_cost) { best_cost = cost; best_axis = a; best_pos = k + 1; }
      }
    }
    if (leaf_max == 0 && n <= 7 && kt >= 0.0) {
      const double wsum = (n - n_generic) + SAH_W_GENERIC * n_generic, area = bounds.area();
      if (best_axis < 0 || !(kt * area + best_cost * (wsum / n) < wsum * area)) return -1;
    }
    if (best_axis < 0) return b + n / 2;  // all centroids coincide
    Prim tmp[8];
    for (int k = 0; k < n; k++) tmp[k] = prims[(size_t)(b + order[best_axis][k])];
    child_box[0].reset();
    child_box[1].reset();
    for (int k = 0; k < n; k++) {
      prims[(size_t)(b + k)] = tmp[k];
      child_box[k < best_pos ? 0 : 1].add(tmp[k].box);
    }
    *have_boxes = true;
    return b + best_pos;
  }

  // Split prims[b,e) — returns mid (b < mid < e), or -1 to make a leaf. When the cut came from the bins (or the small
  // sweep) the two sides' boxes are already known: child_box[0..1], *have_boxes (the same min/max over the same prim
  // boxes as range_box, bit for bit), which saves the caller a pass over each child.
  int split(int b, int e, const Box& bounds, int depth, int nthreads, Box* child_box, bool* have_boxes) {
    const int n = e - b;
    *have_boxes = false;
    if (n <= 1) return -1;
    static const int leaf_max = [] { const char* e = std::getenv("RTC_LEAF_MAX"); int v = e ? std::atoi(e) : 0; return v < 0 ? 0 : (v > 7 ? 7 : v); }();
    if (leaf_max > 0 && n <= leaf_max) return -1;  // experiments: a fixed leaf size instead of the SAH's own stop
    if (n <= 8) return split_small(b, e, bounds, leaf_max, child_box, have_boxes);
    if (depth >= 40) {  // bound the traversal stack: median split along the widest axis from here on
      int a = 0;
      for (int k = 1; k < 3; k++)
        if (bounds.hi[k] - bounds.lo[k] > bounds.hi[a] - bounds.lo[a]) a = k;
      std::nth_element(prims.begin() + b, prims.begin() + b + n / 2, prims.begin() + e,
                       [a](const Prim& x, const Prim& y) { return x.c[a] < y.c[a]; });
      return b + n / 2;
    }
    // centroid bounds, then ONE pass that bins all three axes; both passes are cut into chunks run on
    // worker threads when the range is large (the top levels of a million-triangle mesh)
    const int nchunk = (n >= PAR_MIN && nthreads > 1) ? nthreads : 1;
    // (single-chunk ranges — all the small ones, i.e. most calls — keep their scratch on this frame: no heap traffic)
    Box cb;
    cb.reset();
    if (nchunk == 1) {
      for (int k = b; k < e; k++) cb.add(prims[(size_t)k].c);
    } else {
      std::vector<Box> cbs((size_t)nchunk);
      parallel_chunks(b, e, nchunk, [&](int ci, int cb0, int ce0) {
        Box x;
        x.reset();
        for (int k = cb0; k < ce0; k++) x.add(prims[(size_t)k].c);
        cbs[(size_t)ci] = x;
      });
      for (const Box& x : cbs) cb.add(x);
    }
    struct Bins {
      Box bb[3][SAH_BINS];
      int cnt[3][SAH_BINS];
    };
    double scale3[3];
    bool use_axis[3];
    for (int a = 0; a < 3; a++) {
      const double ext = cb.hi[a] - cb.lo[a];
      use_axis[a] = ext > 0.0;
      scale3[a] = use_axis[a] ? SAH_BINS / ext : 0.0;
    }
    Bins part_local;
    std::vector<Bins> part_heap;
    if (nchunk > 1) part_heap.resize((size_t)nchunk);
    Bins* part = nchunk > 1 ? part_heap.data() : &part_local;
    parallel_chunks(b, e, nchunk, [&](int ci, int cb0, int ce0) {
      Bins& B = part[(size_t)ci];
      for (int a = 0; a < 3; a++)
        for (int k = 0; k < SAH_BINS; k++) { B.bb[a][k].reset(); B.cnt[a][k] = 0; }
      for (int k = cb0; k < ce0; k++) {
        const Prim& p = prims[(size_t)k];
        for (int a = 0; a < 3; a++) {
          if (!use_axis[a]) continue;
          int bi = (int)((p.c[a] - cb.lo[a]) * scale3[a]);
          bi = bi < 0 ? 0 : (bi >= SAH_BINS ? SAH_BINS - 1 : bi);
          B.bb[a][bi].add(p.box);
          B.cnt[a][bi]++;
        }
      }
    });
    for (int ci = 1; ci < nchunk; ci++)
      for (int a = 0; a < 3; a++)
        for (int k = 0; k < SAH_BINS; k++)
          if (part[(size_t)ci].cnt[a][k]) {
            part[0].bb[a][k].add(part[(size_t)ci].bb[a][k]);
            part[0].cnt[a][k] += part[(size_t)ci].cnt[a][k];
          }
    int best_axis = -1, best_bin = -1;
    double best_cost = std::numeric_limits<double>::infinity();
    for (int a = 0; a < 3; a++) {
      if (!use_axis[a]) continue;
      const Box* bb = part[0].bb[a];
      const int* cnt = part[0].cnt[a];
      double right_area[SAH_BINS];
      int right_cnt[SAH_BINS];
      Box acc;
      acc.reset();
      int c = 0;
      for (int k = SAH_BINS - 1; k > 0; k--) {
        if (cnt[k]) acc.add(bb[k]);
        c += cnt[k];
        right_area[k] = c ? acc.area() : 0.0;
        right_cnt[k] = c;
      }
      acc.reset();
      c = 0;
      for (int k = 0; k + 1 < SAH_BINS; k++) {
        if (cnt[k]) acc.add(bb[k]);
        c += cnt[k];
        if (c == 0 || right_cnt[k + 1] == 0) continue;
        double cost = acc.area() * c + right_area[k + 1] * right_cnt[k + 1];
        if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = k; }
      }
    }
    int mid;
    if (best_axis < 0) {  // all centroids coincide: split in the middle
      mid = b + n / 2;
    } else {
      (void)bounds;
      const double ext = cb.hi[best_axis] - cb.lo[best_axis];
      const double scale = SAH_BINS / ext, lo = cb.lo[best_axis];
      const int a = best_axis, bin = best_bin;
      auto goes_left = [&](const Prim& p) {
        int bi = (int)((p.c[a] - lo) * scale);
        bi = bi < 0 ? 0 : (bi >= SAH_BINS ? SAH_BINS - 1 : bi);
        return bi <= bin;
      };
      if (nchunk > 1) {
        // the top levels of a big mesh: every chunk knows from its bins how many of its prims go left, so all chunks
        // scatter into `scratch` at once (a serial std::partition of a million 80-byte prims is ~30 ms per level)
        std::vector<int> lc((size_t)nchunk, 0);
        int total_left = 0;
        for (int k = 0; k <= bin; k++) total_left += part[0].cnt[a][k];  // part[0] holds the merged counts
        int others = 0;
        for (int ci = 1; ci < nchunk; ci++) {
          for (int k = 0; k <= bin; k++) lc[(size_t)ci] += part[ci].cnt[a][k];
          others += lc[(size_t)ci];
        }
        lc[0] = total_left - others;
        std::vector<int> loff((size_t)nchunk), roff((size_t)nchunk);
        int lacc = 0, racc = total_left;
        for (int ci = 0; ci < nchunk; ci++) {
          const int cb0 = b + (int)((long long)n * ci / nchunk), ce0 = b + (int)((long long)n * (ci + 1) / nchunk);
          loff[(size_t)ci] = lacc;
          roff[(size_t)ci] = racc;
          lacc += lc[(size_t)ci];
          racc += (ce0 - cb0) - lc[(size_t)ci];
        }
        if (scratch.size() < (size_t)n) scratch.resize((size_t)n);
        parallel_chunks(b, e, nchunk, [&](int ci, int cb0, int ce0) {
          int l = loff[(size_t)ci], r = roff[(size_t)ci];
          for (int k = cb0; k < ce0; k++) {
            const Prim& p = prims[(size_t)k];
            scratch[(size_t)(goes_left(p) ? l++ : r++)] = p;
          }
        });
        parallel_chunks(b, e, nchunk, [&](int, int cb0, int ce0) {
          std::copy(scratch.begin() + (cb0 - b), scratch.begin() + (ce0 - b), prims.begin() + cb0);
        });
        mid = b + total_left;
      } else {
        auto it = std::partition(prims.begin() + b, prims.begin() + e, goes_left);
        mid = (int)(it - prims.begin());
      }
      if (mid == b || mid == e) {
        mid = b + n / 2;
      } else {
        child_box[0].reset();
        child_box[1].reset();
        for (int k = 0; k < SAH_BINS; k++)
          if (part[0].cnt[a][k]) child_box[k <= bin ? 0 : 1].add(part[0].bb[a][k]);
        *have_boxes = true;
      }
    }
    return mid;
  }

  static Box range_box(const std::vector<Prim>& p, int b, int e) {
    Box r;
    r.reset();
    for (int k = b; k < e; k++) r.add(p[(size_t)k].box);
    return r;
  }

  void set_child(BvhNode& nd, int which, const Box& bx, int32_t c, int32_t n) {
    bvh_real* lo = which ? nd.lo1 : nd.lo0;
    bvh_real* hi = which ? nd.hi1 : nd.hi0;
    for (int a = 0; a < 3; a++) { lo[a] = bvh_below(bx.lo[a]); hi[a] = bvh_above(bx.hi[a]); }
    if (which) { nd.c1 = c; nd.n1 = n; } else { nd.c0 = c; nd.n0 = n; }
  }

  struct Task { int b, e, node, which, depth; bool has_box; Box box; };

  // Build the subtrees on `stack` into `nodes`. With `deferred`, a range of fewer than defer_below prims
  // is not built but handed back (its parent slot stays empty until the caller splices the subtree in).
  void run(std::vector<BvhNode>& nodes, std::vector<Task>& stack, int defer_below, std::vector<Task>* deferred,
           int32_t& max_depth, int nthreads) {
    Box empty;
    empty.reset();
    while (!stack.empty()) {
      Task t = stack.back();
      stack.pop_back();
      if (deferred && t.e - t.b < defer_below) {
        deferred->push_back(t);
        continue;
      }
      max_depth = std::max(max_depth, (int32_t)t.depth);
      Box bx;
      if (t.has_box) {
        bx = t.box;
      } else if (t.e - t.b >= PAR_MIN && nthreads > 1) {
        std::vector<Box> parts((size_t)nthreads);
        parallel_chunks(t.b, t.e, nthreads, [&](int ci, int b0, int e0) { parts[(size_t)ci] = range_box(prims, b0, e0); });
        bx.reset();
        for (const Box& x : parts) bx.add(x);
      } else {
        bx = range_box(prims, t.b, t.e);
      }
      Box child_box[2];
      bool have_boxes = false;
      int mid = split(t.b, t.e, bx, t.depth, nthreads, child_box, &have_boxes);
      if (mid < 0) {
        std::sort(prims.begin() + t.b, prims.begin() + t.e, [](const Prim& x, const Prim& y) { return x.node < y.node; });
        set_child(nodes[(size_t)t.node], t.which, bx, t.b, t.e - t.b);
      } else {
        int32_t id = (int32_t)nodes.size();
        BvhNode nd{};
        set_child(nd, 0, empty, 0, -1);
        set_child(nd, 1, empty, 0, -1);
        nodes.push_back(nd);
        set_child(nodes[(size_t)t.node], t.which, bx, id, 0);
        stack.push_back({t.b, mid, id, 0, t.depth + 1, have_boxes, child_box[0]});
        stack.push_back({mid, t.e, id, 1, t.depth + 1, have_boxes, child_box[1]});
      }
    }
  }

  void build() {
    const int n = (int)prims.size();
    out.bvh.clear();
    out.max_depth = 0;
    if (n == 0) return;
    std::vector<Task> stack;
    BvhNode root{};
    Box empty;
    empty.reset();
    set_child(root, 0, empty, 0, -1);
    set_child(root, 1, empty, 0, -1);
    out.bvh.push_back(root);
    stack.push_back({0, n, 0, 0, 1, false, empty});
    const auto tb0 = std::chrono::steady_clock::now();
    if (n < PAR_SUBTREES_MIN || threads <= 1) {
      run(out.bvh, stack, 0, nullptr, out.max_depth, 1);
    } else {
      // top of the tree on this thread (its big ranges binned by all threads), every subtree below
      // n / (4·threads) prims built by a worker into its own node vector and spliced in afterwards in
      // task order, so the tree does not depend on thread timing
      std::vector<Task> deferred;
      // (a big mesh hands out everything the calling thread would otherwise bin alone, i.e. below PAR_MIN: million
      // triangles, top of the tree 100 -> 63 ms, profiles/r02av_defer.log)
      int defer_below = std::max(n / (4 * threads), std::min(512, PAR_MIN));
      if (n >= 4 * PAR_MIN) defer_below = std::max(defer_below, PAR_MIN);
      if (const char* v = std::getenv("RTC_BVH_DEFER_BELOW")) defer_below = std::max(8, std::atoi(v));
      run(out.bvh, stack, defer_below, &deferred, out.max_depth, threads);
      const auto tp1 = std::chrono::steady_clock::now();
      struct Sub { std::vector<BvhNode> nodes; int32_t depth = 0; };
      std::vector<Sub> subs(deferred.size());
      std::vector<size_t> order(deferred.size());
      for (size_t i = 0; i < order.size(); i++) order[i] = i;
      std::sort(order.begin(), order.end(), [&](size_t x, size_t y) { return deferred[x].e - deferred[x].b > deferred[y].e - deferred[y].b; });
      std::atomic<size_t> next{0};
      auto worker = [&]() {
        for (;;) {
          const size_t k = next.fetch_add(1);
          if (k >= order.size()) return;
          const size_t i = order[k];
          Sub& sb = subs[i];
          BvhNode holder{};
          set_child(holder, 0, empty, 0, -1);
          set_child(holder, 1, empty, 0, -1);
          sb.nodes.push_back(holder);
          std::vector<Task> st;
          st.push_back({deferred[i].b, deferred[i].e, 0, 0, deferred[i].depth, deferred[i].has_box, deferred[i].box});
          run(sb.nodes, st, 0, nullptr, sb.depth, 1);
        }
      };
      std::vector<std::thread> pool;
      for (int t = 1; t < threads; t++) pool.emplace_back(worker);
      worker();
      for (auto& t : pool) t.join();
      if (std::getenv("RTC_TIMING"))
        std::fprintf(stderr, "BVH build: top of the tree %.1f ms (%zu subtrees handed out), subtrees %.1f ms\n",
                     std::chrono::duration<double, std::milli>(tp1 - tb0).count(), deferred.size(),
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tp1).count());
      for (size_t i = 0; i < deferred.size(); i++) {
        const Sub& sb = subs[i];
        const int32_t base = (int32_t)out.bvh.size() - 1;  // local node j (>= 1) becomes base + j
        for (size_t j = 1; j < sb.nodes.size(); j++) {
          BvhNode nd = sb.nodes[j];
          if (nd.n0 == 0) nd.c0 += base;
          if (nd.n1 == 0) nd.c1 += base;
          out.bvh.push_back(nd);
        }
        const BvhNode& h = sb.nodes[0];
        BvhNode& parent = out.bvh[(size_t)deferred[i].node];
        Box bx;
        for (int a = 0; a < 3; a++) { bx.lo[a] = h.lo0[a]; bx.hi[a] = h.hi0[a]; }
        set_child(parent, deferred[i].which, bx, h.n0 == 0 ? h.c0 + base : h.c0, h.n0);
        out.max_depth = std::max(out.max_depth, sb.depth);
      }
    }
    const auto tb1 = std::chrono::steady_clock::now();
    // BVH leaves address the combined prim array: shift by the always-prefix
    const int32_t shift = (int32_t)out.prims.size();
    for (auto& nd : out.bvh) {
      if (nd.n0 > 0) nd.c0 += shift;
      if (nd.n1 > 0) nd.c1 += shift;
    }
#if RTK_BVH_WIDE
    {
      static const int width = [] { const char* e = std::getenv("RTC_BVH_WIDTH"); int v = e ? std::atoi(e) : 4; return v < 2 ? 2 : (v > 4 ? 4 : v); }();
      int need = collapse_wide(out.bvh, out.wide, width);
      // the traversal stack is WIDE_STACK_MAX entries (rt_accel.cuh: BVH_STACK): a tree that could need more keeps its
      // binary shape, whose need is its depth (bounded by the builder)
      if (need + 2 > WIDE_STACK_MAX) need = collapse_wide(out.bvh, out.wide, 2);
      out.wide_stack_need = need;
    }
#endif
    // the leaf records gather from the node and triangle tables in BVH order, i.e. at random: on all threads for a
    // big mesh (serially it was a third of the million-triangle build)
    const size_t base = out.prims.size();
    out.prims.resize(base + (size_t)n);
    out.leaf_recs.resize((size_t)n);
    parallel_chunks(0, n, n >= PAR_MIN ? threads : 1, [&](int, int k0, int k1) {
      for (int k = k0; k < k1; k++) {
        const int32_t i = prims[(size_t)k].node;
        const rtc_node& nd = desc->nodes[i];
        out.prims[base + (size_t)k] = i;
        LeafRec r{};
        r.node = i; r.xform = nd.xform; r.parent = out.parent[(size_t)i];
        r.kind_flags = ((uint32_t)nd.kind << 24) | (nd.flags & 0xffffffu);
        r.aux = nd.aux;
        if (nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE)
          for (int c = 0; c < 9; c++) r.g[c] = desc->tri_p1e1e2[9 * (size_t)nd.aux + c];
        out.leaf_recs[(size_t)k] = r;
      }
    });
    if (std::getenv("RTC_TIMING")) {
      const auto tb2 = std::chrono::steady_clock::now();
      std::fprintf(stderr, "BVH build: tree %.1f ms, leaf records %.1f ms\n", std::chrono::duration<double, std::milli>(tb1 - tb0).count(),
                   std::chrono::duration<double, std::milli>(tb2 - tb1).count());
    }
  }
};

}  // namespace accel_detail

#ifndef RTK_BVH_MIN_PRIMS
#define RTK_BVH_MIN_PRIMS 8
#endif
constexpr int BVH_MIN_PRIMS = RTK_BVH_MIN_PRIMS;  // fewer BVH-eligible prims than this: test them all, no BVH

inline void fill_recs(const rtc_scene_desc* d, AccelBuild& A) {
  // kind-sorted so the every-ray loop runs one tight, switch-free loop per common kind; the order
  // of evaluation is free because ties in t resolve by node index, not by arrival
  auto cls = [&](int32_t i) { int k = d->nodes[i].kind; return k == RTC_SPHERE ? 0 : (k == RTC_PLANE ? 1 : 2); };
  std::stable_sort(A.prims.begin(), A.prims.begin() + A.n_always, [&](int32_t x, int32_t y) { return cls(x) < cls(y); });
  A.n_sph = A.n_pl = 0;
  for (int32_t k = 0; k < A.n_always; k++) {
    int c = cls(A.prims[(size_t)k]);
    if (c == 0) A.n_sph++; else if (c == 1) A.n_pl++;
  }
  A.recs.assign((size_t)A.n_always, PrimRec{});
  for (int32_t k = 0; k < A.n_always; k++) {
    const int32_t i = A.prims[(size_t)k];
    const rtc_node& nd = d->nodes[i];
    PrimRec& r = A.recs[(size_t)k];
    r.node = i; r.aux = nd.aux; r.parent = A.parent[(size_t)i];
    r.kind_flags = ((uint32_t)nd.kind << 24) | (nd.flags & 0xffffffu);
    if (nd.kind != RTC_CSG)
      for (int c = 0; c < 12; c++) r.inv[c] = d->xforms[12 * (size_t)nd.xform + c];
  }
  // Bounding spheres for the sphere prefix. A ray whose line stays outside the padded bounding sphere
  // has, in the reference's object space, a discriminant below -4a * 1e-4 — six orders of magnitude
  // beyond what f64 rounding can flip while |o - c|^2 < far2 = 1e6 / ||inv||_F^2 (object-space distance
  // below 1e3 radii) and the centre is not astronomically far from the world origin in object units.
  for (int k = 0; k < SPH_CULL_MAX; k++) A.sph_cull[k] = SphCull{0.0, 0.0, 0.0, HUGE_VAL, 0.0, 0.0};
  for (int32_t k = 0; k < A.n_sph && k < SPH_CULL_MAX; k++) {
    const double* inv = A.recs[(size_t)k].inv;
    double fwd[12];
    if (!accel_detail::invert_affine(inv, fwd)) continue;
    // sigma_max(M)^2 = rho(M^T M) <= ||M^T M||_inf (largest absolute row sum): tight for rotations times a
    // uniform scale, where the Frobenius norm would overshoot by sqrt(3)
    double f2 = 0.0, i2 = 0.0;
    for (int r = 0; r < 3; r++) {
      double row = 0.0;
      for (int c = 0; c < 3; c++) {
        double g = 0.0;
        for (int m = 0; m < 3; m++) g += fwd[4 * m + r] * fwd[4 * m + c];
        row += std::fabs(g);
        i2 += inv[4 * r + c] * inv[4 * r + c];
      }
      f2 = std::max(f2, row);
    }
    const double cx = fwd[3], cy = fwd[7], cz = fwd[11];
    const double c2 = cx * cx + cy * cy + cz * cz;
    if (!std::isfinite(f2) || !std::isfinite(i2) || !std::isfinite(c2) || !(i2 > 0.0) || !(f2 > 0.0)) continue;
    if (i2 * c2 > 1e12) continue;  // ||inv||_F * |c| > 1e6: translation cancellation in the reference's own o_obj
    A.sph_cull[k] = SphCull{cx, cy, cz, f2 * (1.0 + 1e-4), 1e6 / i2, 0.0};
  }
}

// Face families of the scene's cubes (see CubeDir). Returns false when there are more than CUBE_DIR_MAX, or a
// transform is unusable: the cubes then stay every-ray prims.
inline bool cube_face_families(const rtc_scene_desc* d, AccelBuild& A) {
  A.n_cube_dirs = 0;
  std::vector<char> seen(d->n_xforms, 0);
  for (uint32_t i = 0; i < d->n_nodes; i++) {
    const rtc_node& nd = d->nodes[i];
    if (nd.kind != RTC_CUBE || seen[(size_t)nd.xform]) continue;
    seen[(size_t)nd.xform] = 1;
    const double* inv = d->xforms + 12 * (size_t)nd.xform;
    for (int r = 0; r < 3; r++) {
      double u[3] = {inv[4 * r], inv[4 * r + 1], inv[4 * r + 2]};
      const double len = std::sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
      if (!(len > 0.0) || !std::isfinite(len)) return false;
      for (int c = 0; c < 3; c++) u[c] /= len;
      // canonical sign: the first component of magnitude above 1e-9 is positive
      for (int c = 0; c < 3; c++) {
        if (std::fabs(u[c]) > 1e-9) {
          if (u[c] < 0.0) for (int k = 0; k < 3; k++) u[k] = -u[k];
          break;
        }
      }
      const double theta = 1e-5 / len;  // EPSILON, lib.rs:18
      const double th2 = 2.0 * theta * theta * (1.0 + 1e-6);
      if (!std::isfinite(th2)) return false;
      int k = 0;
      for (; k < A.n_cube_dirs; k++) {
        const CubeDir& c = A.cube_dirs[k];
        if (std::fabs(c.ux - u[0]) < 1e-13 && std::fabs(c.uy - u[1]) < 1e-13 && std::fabs(c.uz - u[2]) < 1e-13) break;
      }
      if (k < A.n_cube_dirs) {
        A.cube_dirs[k].th2 = std::max(A.cube_dirs[k].th2, th2);
      } else {
        if (A.n_cube_dirs >= CUBE_DIR_MAX) { A.n_cube_dirs = 0; return false; }
        A.cube_dirs[A.n_cube_dirs++] = CubeDir{u[0], u[1], u[2], th2};
      }
    }
  }
  return true;
}

// `d` must already be structurally valid (rtc_api.cu validate()). device_min > 0: with at least that many BVH-eligible
// leaves (and no cube among them) only the walk is done here and the tree is left to the GPU (A.deferred).
inline void build_accel(const rtc_scene_desc* d, AccelBuild& A, size_t device_min = 0) {
  using namespace accel_detail;
  const auto t0 = std::chrono::steady_clock::now();
  const int32_t n = (int32_t)d->n_nodes;
  A.deferred = false;
  A.cand_generic = false;
  A.cand.clear();
  A.fwd.clear();
  bool defer = false;
  if (device_min > 0) {
    size_t eligible = 0;
    bool any_cube = false;
    for (int32_t i = 0; i < n;) {
      const rtc_node& nd = d->nodes[i];
      if (nd.kind == RTC_CSG) { i = nd.skip; continue; }
      if (nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE || nd.kind == RTC_SPHERE || nd.kind == RTC_CYLINDER) eligible++;
      if (nd.kind == RTC_CUBE) any_cube = true;
      i++;
    }
    defer = eligible >= device_min && !any_cube;
  }
  A.parent.assign((size_t)n, -1);
  A.prims.clear();
  A.n_always = 0;
  A.bvh.clear();
  A.leaf_recs.clear();
  std::vector<Prim> cand;
  if (!defer && n >= 65536) cand.reserve((size_t)n);  // (growing 80 MB of candidates by doubling copies them twice over)
  FwdCache fwd_cache(d->n_xforms);
  bool cubes_ok = cube_face_families(d, A);
  if (cubes_ok && A.n_cube_dirs > 0) {
    size_t eligible = 0;
    for (int32_t k = 0; k < n;) {
      const rtc_node& nk = d->nodes[k];
      if (nk.kind == RTC_CSG) { k = nk.skip; continue; }
      if (nk.kind != RTC_GROUP && nk.kind != RTC_PLANE && nk.kind != RTC_CONE) eligible++;
      k++;
    }
    if (eligible > CUBE_BVH_MAX_PRIMS) { cubes_ok = false; A.n_cube_dirs = 0; }
  }
  if (std::getenv("RTC_TIMING")) std::fprintf(stderr, "build_accel: %d cube face families, cubes %s\n", A.n_cube_dirs, cubes_ok ? "may enter the BVH" : "stay every-ray prims");
  struct Open { int32_t end, node; };
  std::vector<Open> groups;
  int32_t i = 0;
  while (i < n) {
    while (!groups.empty() && groups.back().end <= i) groups.pop_back();
    const rtc_node& nd = d->nodes[i];
    A.parent[(size_t)i] = groups.empty() ? -1 : groups.back().node;
    if (nd.kind == RTC_GROUP) {
      if (nd.skip > i + 1) groups.push_back({nd.skip, i});
      i = i + 1;
    } else if (nd.kind == RTC_CSG) {
      A.prims.push_back(i);
      i = nd.skip;  // the subtree is evaluated by the in-order walk
    } else {
      Prim p;
      p.node = i;
      p.generic = (nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE || nd.kind == RTC_SPHERE) ? 0 : 1;
      if (defer) {
        // eligibility without the box (the GPU computes it; a non-finite one sends the whole build back here)
        bool ok = (nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE || nd.kind == RTC_SPHERE || nd.kind == RTC_CYLINDER) &&
                  fwd_cache.get(d, nd.xform) != nullptr;
        if (ok && nd.kind == RTC_CYLINDER) {
          const double y0 = d->limits[2 * (size_t)nd.aux], y1 = d->limits[2 * (size_t)nd.aux + 1];
          ok = std::isfinite(y0) && std::isfinite(y1) && y0 <= y1;
        }
        if (ok) {
          A.cand.push_back(i);
          if (nd.kind == RTC_CYLINDER) A.cand_generic = true;
        } else {
          A.prims.push_back(i);
        }
      } else if (nd.kind != RTC_PLANE) {
        cand.push_back(p);  // its box: the pass below
      } else {
        A.prims.push_back(i);
      }
      i = i + 1;
    }
  }
  if (!defer && !cand.empty()) {
    // world boxes of the candidates (three transformed vertices per triangle): on all host threads for a big scene.
    // The forward matrices are inverted up front then (the cache is filled lazily otherwise, which is not thread-safe).
    unsigned hc = std::thread::hardware_concurrency();
    int nthreads = hc == 0 ? 1 : (hc > 16 ? 16 : (int)hc);
    if (const char* v = std::getenv("RTC_BVH_THREADS")) nthreads = std::max(1, std::min(64, std::atoi(v)));
    if (cand.size() < 65536) nthreads = 1;
    if (nthreads > 1)
      for (uint32_t x = 0; x < d->n_xforms; x++) fwd_cache.get(d, (int32_t)x);
    std::vector<unsigned char> boxed(cand.size(), 0);
    Builder::parallel_chunks(0, (int)cand.size(), nthreads, [&](int, int b0, int e0) {
      for (int k = b0; k < e0; k++) {
        Prim& q = cand[(size_t)k];
        if (leaf_world_box(d, d->nodes[q.node], q.box, fwd_cache, cubes_ok)) {
          for (int a = 0; a < 3; a++) q.c[a] = 0.5 * (q.box.lo[a] + q.box.hi[a]);
          boxed[(size_t)k] = 1;
        }
      }
    });
    size_t kept = 0;
    bool dropped = false;
    for (size_t k = 0; k < cand.size(); k++) {
      if (boxed[k]) { if (kept != k) cand[kept] = cand[k]; kept++; }
      else { A.prims.push_back(cand[k].node); dropped = true; }
    }
    cand.resize(kept);
    if (dropped) std::sort(A.prims.begin(), A.prims.end());  // every-ray prims stay in node order
  }
  // Nested boxes let one test stand for a whole chain. For one ray, per-axis slab intervals are
  // monotone in the box bounds even in floating point (x - o and x / d are monotone, and the
  // EPSILON-parallel branch only asks "inside the slab?"), so if box(g) ⊆ box(a) component-wise
  // (all finite) then passing g's box implies passing a's. parent[g] of a GROUP therefore skips
  // every ancestor that contains g's box and names the first one that does not (normally none:
  // Group boxes are unions of their children's, group.rs:128-133). Leaves keep their innermost group.
  {
    auto contains = [&](int32_t a, int32_t c) {
      const double* A_ = d->boxes + 6 * (size_t)d->nodes[a].aux;
      const double* C_ = d->boxes + 6 * (size_t)d->nodes[c].aux;
      for (int k = 0; k < 6; k++)
        if (!std::isfinite(A_[k]) || !std::isfinite(C_[k])) return false;
      for (int k = 0; k < 3; k++)
        if (!(A_[k] <= C_[k] && C_[k + 3] <= A_[k + 3] && C_[k] <= C_[k + 3])) return false;
      return true;
    };
    std::vector<int32_t> real_parent = A.parent;
    for (int32_t g = 0; g < n; g++) {
      if (d->nodes[g].kind != RTC_GROUP) continue;
      int32_t a = real_parent[(size_t)g];
      while (a >= 0 && contains(a, g)) a = real_parent[(size_t)a];
      A.parent[(size_t)g] = a;
    }
  }
  if (defer) {
    A.n_always = (int32_t)A.prims.size();
    A.n_cube_dirs = 0;
    fill_recs(d, A);
    A.deferred = true;
    A.fwd.assign(12 * (size_t)d->n_xforms, std::numeric_limits<double>::quiet_NaN());
    for (uint32_t x = 0; x < d->n_xforms; x++)
      if (fwd_cache.state[x] > 0)
        for (int c = 0; c < 12; c++) A.fwd[12 * (size_t)x + c] = fwd_cache.m[12 * (size_t)x + c];
    if (std::getenv("RTC_TIMING"))
      std::fprintf(stderr, "build_accel: walk + parents %.1f ms, %zu BVH prims left to the GPU\n",
                   std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(), A.cand.size());
    return;
  }
  if ((int)cand.size() < BVH_MIN_PRIMS) {
    for (auto& p : cand) A.prims.push_back(p.node);
    std::sort(A.prims.begin(), A.prims.end());
    A.n_always = (int32_t)A.prims.size();
    A.n_cube_dirs = 0;
    fill_recs(d, A);
    return;
  }
  A.n_always = (int32_t)A.prims.size();
  fill_recs(d, A);
  const auto t1 = std::chrono::steady_clock::now();
  Builder b(cand, A, d);
  b.build();
  if (std::getenv("RTC_TIMING")) {
    const auto t2 = std::chrono::steady_clock::now();
    std::fprintf(stderr, "build_accel: walk+boxes+parents %.1f ms, BVH over %zu prims %.1f ms (%d threads)\n",
                 std::chrono::duration<double, std::milli>(t1 - t0).count(), cand.size(),
                 std::chrono::duration<double, std::milli>(t2 - t1).count(), b.threads);
  }
}

}  // namespace rtk
