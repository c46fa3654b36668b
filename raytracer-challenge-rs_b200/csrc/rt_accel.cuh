// rt_accel.cuh — the accelerated scene search (device side of rt_accel_build.hpp).
//
// Same answers as the in-order walk (rt_walk.cuh), different search: the `always` prims are
// tested for every ray, the rest through a front-to-back BVH over OUR OWN tight boxes with t-range
// culling; a candidate intersection the consumer wants is accepted only after the ray has passed
// BoundingBox::intersects (bounding_box.rs:95-116, no t-range) of every reference Group that
// encloses the leaf — the exact condition under which Group::intersect (group.rs:49-58) would have
// reached it. Validation is lazy (only for wanted candidates) and memoised on the innermost group.
#pragma once
#include "rt_walk.cuh"

namespace rtk {

#if defined(RTK_STATS) && !defined(__CUDA_ARCH__)
struct AccelStats { unsigned long long walks, bvh_nodes, prim_tests, chain_calls, chain_boxes, always_tests, sph_culled; };
inline AccelStats& accel_stats() { static AccelStats s; return s; }
#define RTK_COUNT(f, n) (accel_stats().f += (n))
#elif defined(RTK_DEVICE_STATS) && defined(__CUDACC__)
// -DRTK_DEVICE_STATS (experiments: the algorithmic byte count of a frame): the same counters on the device, one set per
// SM to keep the atomics apart; rtc_scene_finish prints the sums. Exact, slow.
enum { DST_walks, DST_bvh_nodes, DST_prim_tests, DST_chain_calls, DST_chain_boxes, DST_always_tests, DST_sph_culled, DST_N };
__device__ unsigned long long g_dev_stats[256][8];
__device__ __forceinline__ unsigned dev_smid() { unsigned r; asm volatile("mov.u32 %0, %%smid;" : "=r"(r)); return r & 255u; }
#if defined(__CUDA_ARCH__)
#define RTK_COUNT(f, n) atomicAdd(&g_dev_stats[dev_smid()][DST_##f], (unsigned long long)(n))
#else
#define RTK_COUNT(f, n) ((void)0)
#endif
#else
#define RTK_COUNT(f, n) ((void)0)
#endif

constexpr int BVH_STACK = 64;  // build bounds the depth (SAH to depth 40, then median splits)

struct BvhStack {
  int32_t node[BVH_STACK];
  bvh_real tmin[BVH_STACK];
};

#if RTK_BVH_F32
RTC_HD float rt_fmaf(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(a, b, c);
#else
  return fmaf(a, b, c);
#endif
}

// The single-precision slab test of one of OUR boxes. t = lo*inv - o*inv as ONE fused multiply-add per plane:
// `i*` = fl32(1/d), `c*` = fl32(-o/d). Against the exact entry/exit parameters of the (already outward-rounded) box the
// computed ones are off by at most (|lo| + |o| + |lo - o|)*|1/d|*2^-24 per plane (one rounding each in i, c and the
// fma); `e*` = (R + |o|)*|1/d|*5e-7 with R the tree's largest |coordinate| on that axis is four times that, so
// [tmin, tmax] CONTAINS the true interval: a box the exact test would enter is never missed and tmin never exceeds the
// true entry parameter. The caller guarantees |i|, |c|, R <= 1e15 (no overflow, hence no NaN).
struct Slab32 {
  float ix, iy, iz, cx, cy, cz, ex, ey, ez;
  RTC_HD bool test(float lx, float ly, float lz, float hx, float hy, float hz, float& tmin, float& tmax) const {
    const float ax = rt_fmaf(lx, ix, cx), bx = rt_fmaf(hx, ix, cx);
    const float ay = rt_fmaf(ly, iy, cy), by = rt_fmaf(hy, iy, cy);
    const float az = rt_fmaf(lz, iz, cz), bz = rt_fmaf(hz, iz, cz);
    const float x0 = fminf(ax, bx) - ex, x1 = fmaxf(ax, bx) + ex;
    const float y0 = fminf(ay, by) - ey, y1 = fmaxf(ay, by) + ey;
    const float z0 = fminf(az, bz) - ez, z1 = fmaxf(az, bz) + ez;
    tmin = fmaxf(fmaxf(x0, y0), z0);
    tmax = fminf(fminf(x1, y1), z1);
    return !(tmin > tmax);
  }
  // The same with the error bound folded into the addend: the plane the ray meets first on an axis (the low one for a
  // positive direction) gets c - e, the other c + e — one fma per plane and no min/max per axis. ((c -+ e) is rounded
  // once more: (|c| + e)*2^-24, which the fourfold margin of e covers.)
  RTC_HD bool test_nf(float nx, float ny, float nz, float fx, float fy, float fz, float& tmin, float& tmax) const {
    const float x0 = rt_fmaf(nx, ix, cx - ex), x1 = rt_fmaf(fx, ix, cx + ex);
    const float y0 = rt_fmaf(ny, iy, cy - ey), y1 = rt_fmaf(fy, iy, cy + ey);
    const float z0 = rt_fmaf(nz, iz, cz - ez), z1 = rt_fmaf(fz, iz, cz + ez);
    tmin = fmaxf(fmaxf(x0, y0), z0);
    tmax = fminf(fminf(x1, y1), z1);
    return !(tmin > tmax);
  }
};
#endif

// Has `ray` passed the box of every reference Group enclosing `node`? `ok_group`: innermost group
// already validated for this ray (its ancestors are then valid too), -1 initially (= top level).
// The slow part is rare (only for candidates the query wants) and kept out of line: it is six
// IEEE divisions per box, and inlining it at every candidate site bloats the hot loop.
// BoundingBox::intersects (bounding_box.rs:95-136) as a FILTERED predicate: the answer is only the
// boolean tmin <= tmax, so first evaluate the slabs with one reciprocal per axis (6 multiplies
// instead of 6 IEEE divisions); only if tmin and tmax are within 1e-9 (relative) of each other —
// where the rounding of a/b versus a*(1/b) could flip the comparison — fall back to the reference's
// own divisions. The EPSILON-parallel branch (num * INFINITY) involves no division and is taken
// exactly as the reference takes it.
RTC_HD bool ref_box_passes(const double* bx, const RayD& r, V3 inv) {
  double tmin = -rt_inf(), tmax = rt_inf();
  const double o[3] = {r.o.x, r.o.y, r.o.z}, d[3] = {r.d.x, r.d.y, r.d.z}, iv[3] = {inv.x, inv.y, inv.z};
#if defined(__CUDACC__)
#pragma unroll
#endif
  for (int a = 0; a < 3; a++) {
    const double n0 = RTC_LDG(bx + a) - o[a], n1 = RTC_LDG(bx + 3 + a) - o[a];
    const double m = fabs(d[a]) >= EPS ? iv[a] : rt_inf();
    const double ta = n0 * m, tb = n1 * m;
    tmin = fmax(tmin, ta > tb ? tb : ta);   // f64::max/min ignore NaN like fmax/fmin
    tmax = fmin(tmax, ta > tb ? ta : tb);
  }
  const double slack = 1e-9 * (fabs(tmin) + fabs(tmax));
  if (tmin <= tmax - slack) return true;    // ±inf and clear passes
  if (tmin > tmax + slack) return false;    // clear misses
  double e0, e1;                            // ambiguous (or NaN): the reference's exact arithmetic
  return box_intersects(bx, r, e0, e1);
}

RTC_HD_NOINLINE bool chain_walk(const int32_t* parent, const rtc_node* nodes, const double* boxes, int32_t g, int32_t ok_group,
                                RayD ray) {
  const V3 inv = v3(1.0 / ray.d.x, 1.0 / ray.d.y, 1.0 / ray.d.z);
  while (g >= 0) {
    // already validated? ok_group and all its ancestors are; in pre-order, g is an ancestor-or-self
    // of ok_group iff g <= ok_group < skip(g) — so a sibling subtree costs one box, not the whole chain
    const int32_t skip = RTC_LDG(&nodes[g].skip);
    if (g <= ok_group && ok_group < skip) break;
    RTK_COUNT(chain_boxes, 1);
    int32_t bx = RTC_LDG(&nodes[g].aux);
    if (!ref_box_passes(boxes + 6 * (size_t)bx, ray, inv)) return false;
    g = RTC_LDG(parent + g);
  }
  return true;
}
RTC_HD bool chain_ok(const SceneView& S, int32_t g0, const RayD& ray, int32_t& ok_group) {
  RTK_COUNT(chain_calls, 1);
  if (g0 == ok_group) return true;
  if (!chain_walk(S.parent, S.nodes, S.boxes, g0, ok_group, ray)) return false;
  ok_group = g0;
  return true;
}

// Slab test against one of OUR boxes with the reciprocal direction. NaN (0*inf) drops out of
// fmax/fmin, which is the conservative outcome; the boxes carry a guard band (build).
// `plain`: no direction component is zero, so no 0*inf NaN can arise and min/max are single
// compare+select pairs instead of the NaN-aware fmin/fmax sequences.
RTC_HD bool own_box(double lx, double ly, double lz, double hx, double hy, double hz, const RayD& r, V3 inv, bool plain,
                    double& tmin, double& tmax) {
  double ax = (lx - r.o.x) * inv.x, bx = (hx - r.o.x) * inv.x;
  double ay = (ly - r.o.y) * inv.y, by = (hy - r.o.y) * inv.y;
  double az = (lz - r.o.z) * inv.z, bz = (hz - r.o.z) * inv.z;
  if (plain) {
    const double x0 = ax < bx ? ax : bx, x1 = ax < bx ? bx : ax;
    const double y0 = ay < by ? ay : by, y1 = ay < by ? by : ay;
    const double z0 = az < bz ? az : bz, z1 = az < bz ? bz : az;
    const double m0 = x0 > y0 ? x0 : y0, m1 = x1 < y1 ? x1 : y1;
    tmin = m0 > z0 ? m0 : z0;
    tmax = m1 < z1 ? m1 : z1;
  } else {
    tmin = fmax(fmax(fmin(ax, bx), fmin(ay, by)), fmin(az, bz));
    tmax = fmin(fmin(fmax(ax, bx), fmax(ay, by)), fmax(az, bz));
  }
  // relative slack for the reciprocal's rounding
  return tmin <= tmax + 1e-9 * (fabs(tmin) + fabs(tmax));
}

// Lazy division. t = num/den is only worth an IEEE division (≈ a dozen FP64-pipe instructions) if
// the query could want it. These pre-filters never reject a t the query would take: the sign of the
// quotient is the sign of num·den (magnitudes kept away from underflow, where -tiny/den could round
// to -0.0, which `t >= 0.0` accepts), and |num| > limit·(1+1e-9)·|den| puts the correctly rounded
// positive quotient strictly beyond the limit (the tie t == limit is never filtered). The range and
// sign tests read the upper word of the doubles as integers — no 64-bit constants in the hot loops.
// NaN and ±inf numerators are never filtered (a NaN t has to reach the caller's NaN check).
RTC_HD bool quot_unwanted(const Query& C, double num, double den) {
#if defined(RTK_QUERY_BRANCHY)
  if (den < 0.0) { num = -num; den = -den; }
  if (!(den > 0.0) || !(den < 1e150)) return false;
  if (C.mode == Q_NEG) return num > 1e-150;
  if (num < -1e-150) return true;
  return num > C.limit * den * (1.0 + 1e-9);
#else
  const uint32_t dh = hi_word(den), nh = hi_word(num);
  // |den| in [2^-1022, ~1e150): 0x00100000 <= upper word < 0x5F138D35 (upper word of 1e150)
  const bool den_ok = ((dh & 0x7fffffffu) - 0x00100000u) < (0x5F138D35u - 0x00100000u);
  const uint32_t qh = nh ^ (dh & 0x80000000u);                       // upper word of num carrying the quotient's sign
  const uint32_t th = qh ^ (C.mode == Q_NEG ? 0x80000000u : 0u);     // Q_NEG wants t < 0: flip
  // wrong side: sign set and 1e-150 < |num| < inf (upper word of 1e-150 is 0x20CA2FE7)
  const bool wrong_side = (th - 0xA0CA2FE8u) < (0xFFF00000u - 0xA0CA2FE8u);
  const bool beyond = ((qh >> 31) == 0u) & (fabs(num) > C.limit_hi * fabs(den));
  return den_ok & (wrong_side | beyond);
#endif
}

// World::intersect → query. FAITHFUL: the reference's own search order, node by node (walk_range
// over the whole array). Otherwise: the prims tested for every ray, then the BVH; ONE leaf-test
// site and ONE candidate site serve both, so the hot loop stays small.
// LEAN (FAITHFUL == 2): the scene is every-ray spheres and planes only (no BVH, no other kinds, no CSG) —
// the host picks this instantiation so that the generic leaf tests and the traversal are not even in
// the kernel: the hot code of the glass and chapter-7 scenes then fits the instruction caches better.
template <bool HAS_CSG, int FAITHFUL>
RTC_HD void search(const SceneView& S, const RayD& ray, Query& C, CsgScratch* cs, BvhStack* st, uint32_t* err) {
  RTK_COUNT(walks, 1);
  if constexpr (FAITHFUL == 1) {
    (void)st;
    walk_range<HAS_CSG>(S, ray, C, cs, err, 0, S.n_nodes);
    return;
  } else {
    int32_t ok_group = -1;
    RayD lr;
    int32_t lr_xf = -1;
#if !RTK_BVH_F32
    V3 inv = v3(0.0, 0.0, 0.0);
    bool plain = false;
#endif
    bool have_inv = false;
    int sp = 0;
#if RTK_BVH_F32
    // what a query carries through the traversal: nine floats and two flags (the double-precision reciprocal direction
    // is recomputed per node on the rare slow path instead of living in twelve registers)
    Slab32 F;
    F.ix = F.iy = F.iz = F.cx = F.cy = F.cz = F.ex = F.ey = F.ez = 0.0f;
    bool fast32 = false, zero_inv = false;
    auto plain_dir = [&]() {
      return ray.d.x != 0.0 && ray.d.y != 0.0 && ray.d.z != 0.0 && ray.d.x == ray.d.x && ray.d.y == ray.d.y && ray.d.z == ray.d.z;
    };
    // Query::cull in single precision. tmin / tmax are outer bounds of the box's parameter interval; the query's own
    // slack (1e-6 relative) is doubled to absorb the rounding of the limit and of these few float operations.
    // `lim32` follows C.limit, which only moves while leaves are tested: refreshed after every leaf phase.
    float lim32 = HUGE_VALF;
    auto cull32 = [&](float tmin, float tmax) -> bool {
      if (C.mode == Q_NEG) return tmin > 2e-6f * (1.0f + fabsf(tmin));
      return tmax < -2e-6f * (1.0f + fabsf(tmax)) || tmin > lim32;
    };
#define RTK_BVH_CULL(tmin, tmax) cull32((tmin), (tmax))
#define RTK_BVH_TINF HUGE_VALF
#else
#define RTK_BVH_CULL(tmin, tmax) C.cull((tmin), (tmax))
#define RTK_BVH_TINF rt_inf()
#endif
    // -DRTK_SHARED_STACK=n (A/B of the "per-warp shared-memory stack" of the brief): the first n entries of the traversal
    // stack live in shared memory ([entry][thread]: conflict-free), the rest in the thread's local BvhStack.
#if defined(RTK_SHARED_STACK) && defined(__CUDA_ARCH__)
    constexpr int SH_N = RTK_SHARED_STACK;
    __shared__ int32_t sh_node[SH_N][128];
    __shared__ bvh_real sh_tmin[SH_N][128];
    const int sh_t = threadIdx.x & 127;
#define RTK_ST_NODE(k) ((k) < SH_N ? sh_node[(k)][sh_t] : st->node[(k) - SH_N])
#define RTK_ST_TMIN(k) ((k) < SH_N ? sh_tmin[(k)][sh_t] : st->tmin[(k) - SH_N])
#define RTK_ST_PUSH(k, n, t) do { if ((k) < SH_N) { sh_node[(k)][sh_t] = (n); sh_tmin[(k)][sh_t] = (t); } else { st->node[(k) - SH_N] = (n); st->tmin[(k) - SH_N] = (t); } } while (0)
#else
#define RTK_ST_NODE(k) (st->node[(k)])
#define RTK_ST_TMIN(k) (st->tmin[(k)])
#define RTK_ST_PUSH(k, n, t) do { st->node[(k)] = (n); st->tmin[(k)] = (t); } while (0)
#endif
    // one candidate intersection: want → validate against the reference's group chain → take
    auto cand = [&](double t, int32_t i, uint32_t flags, int32_t par) {
      const bool nan = t != t;
      if (!nan && !C.want(t, i, flags)) return;
      if (par != ok_group && !chain_ok(S, par, ray, ok_group)) return;
      if (nan) *err |= ERR_NAN; else C.take(t, i);
    };
    // ---- every-ray spheres (sphere.rs:47-62): switch-free, divisions only for wanted roots ---------
    int32_t k0 = 0;
    // the roots of one sphere whose discriminant is not negative
    auto sphere_roots = [&](const PrimRec* rec, double a, double b, double disc) {
#if defined(__CUDA_ARCH__)
      const int4 hd = __ldg(reinterpret_cast<const int4*>(rec));
      const int32_t i = hd.x, par = hd.z;
      const uint32_t flags = (uint32_t)hd.w & 0xffffffu;
#else
      const int32_t i = rec->node, par = rec->parent;
      const uint32_t flags = rec->kind_flags & 0xffffffu;
#endif
      const double sq = sqrt(disc);
      const double a2 = 2.0 * a, num1 = -b - sq, num2 = -b + sq;
      if (!quot_unwanted(C, num1, a2)) cand(num1 / a2, i, flags, par);
      if (!quot_unwanted(C, num2, a2)) cand(num2 / a2, i, flags, par);
    };
    auto sphere_disc = [&](const PrimRec* rec, double& a, double& b) {
      const RayD r = ray_transform_v2(rec->inv, ray);
      a = dot(r.d, r.d);
      b = 2.0 * dot(r.d, r.o);
      const double c = dot(r.o, r.o) - 1.0;
      return b * b - 4.0 * a * c;
    };
#if !defined(RTK_NO_SPHCULL)
    // World-space bounding-sphere cull (SphCull, rt_accel_build.hpp): 17 flops on parameter-bank data
    // against the 57-flop object-space test. A sphere is skipped when the ray's line misses its padded
    // bounding sphere, or when the half-line the query asks about (t >= 0; t < 0 for Q_NEG) points
    // away from a bounding sphere the origin is outside of — both with a 1e-4 relative pad, orders of
    // magnitude beyond the rounding of either formulation while |o - c|^2 < far2.
    const double dd = dot(ray.d, ray.d);
    const bool want_neg = C.mode == Q_NEG;
#endif
#if !defined(RTK_NO_SPHCULL) && defined(RTK_SPHCULL_ILP)
    // (measured and rejected, kept for the record: glass frame 1.98 -> 2.19 ms — seven predicates at once cost more
    // registers and more work than the early exits they replace save in latency)
    // Phase 1: the cull predicate of ALL culled spheres before any object-space test. The predicates are independent
    // 17-flop chains on parameter-bank data; one after the other (a branch per sphere) each waits out its own FP64
    // latencies, together they fill each other's (stall_wait was 3 cycles per issued instruction on these lines).
    uint32_t sph_keep = 0xffffffffu;  // bit k clear: sphere k cannot be wanted by this query
    {
      const int nc = S.n_sph < SPH_CULL_MAX ? S.n_sph : SPH_CULL_MAX;
#if defined(__CUDACC__)
#pragma unroll
#endif
      for (int k = 0; k < SPH_CULL_MAX; k++) {
        const SphCull& sc = S.sph_cull[k];
        const V3 oc = v3(sc.cx - ray.o.x, sc.cy - ray.o.y, sc.cz - ray.o.z);
        const double bq = dot(oc, ray.d), ocsq = dot(oc, oc);
        const double cq = ocsq - sc.r2;
        const bool culled = ocsq < sc.far2 && cq > 0.0 && (bq * bq < dd * cq || (want_neg ? bq > 0.0 : bq < 0.0));
        if (k < nc && culled) sph_keep &= ~(1u << k);
      }
    }
#endif
    for (; k0 < S.n_sph && !C.done(); k0++) {
      const PrimRec* rec = S.recs + k0;
      RTK_COUNT(prim_tests, 1);
#if !defined(RTK_NO_SPHCULL)
#if defined(RTK_SPHCULL_ILP)
      if (!((sph_keep >> (k0 & 31)) & 1u) && k0 < SPH_CULL_MAX) {
        RTK_COUNT(sph_culled, 1);
        continue;
      }
#else
      if (k0 < SPH_CULL_MAX) {
        const SphCull& sc = S.sph_cull[k0];
        const V3 oc = v3(sc.cx - ray.o.x, sc.cy - ray.o.y, sc.cz - ray.o.z);
        const double bq = dot(oc, ray.d), ocsq = dot(oc, oc);
        const double cq = ocsq - sc.r2;
        if (ocsq < sc.far2 && cq > 0.0 && (bq * bq < dd * cq || (want_neg ? bq > 0.0 : bq < 0.0))) {
          RTK_COUNT(sph_culled, 1);
          continue;
        }
      }
#endif
#endif
      double a, b;
      const double disc = sphere_disc(rec, a, b);
      if (disc < 0.0) continue;
      sphere_roots(rec, a, b, disc);
    }
    // ---- every-ray planes (plane.rs:53-60): only row 1 of the inverse is needed --------------------
    for (; k0 < S.n_sph + S.n_pl && !C.done(); k0++) {
      const PrimRec* rec = S.recs + k0;
      RTK_COUNT(prim_tests, 1);
#if defined(__CUDA_ARCH__)
      const int4 hd = __ldg(reinterpret_cast<const int4*>(rec));
      const int32_t i = hd.x, par = hd.z;
      const uint32_t flags = (uint32_t)hd.w & 0xffffffu;
      const double2 m45 = __ldg(reinterpret_cast<const double2*>(rec->inv + 4));
      const double2 m67 = __ldg(reinterpret_cast<const double2*>(rec->inv + 6));
      const double m4 = m45.x, m5 = m45.y, m6 = m67.x, m7 = m67.y;
#else
      const int32_t i = rec->node, par = rec->parent;
      const uint32_t flags = rec->kind_flags & 0xffffffu;
      const double m4 = rec->inv[4], m5 = rec->inv[5], m6 = rec->inv[6], m7 = rec->inv[7];
#endif
      const double oy = m4 * ray.o.x + m5 * ray.o.y + m6 * ray.o.z + m7;  // matrix.rs:240, row 1
      const double dy = m4 * ray.d.x + m5 * ray.d.y + m6 * ray.d.z;       // matrix.rs:255, row 1
      if (fabs(dy) < EPS) continue;
      if (quot_unwanted(C, -oy, dy)) continue;
      cand(-oy / dy, i, flags, par);
    }
    if constexpr (FAITHFUL == 2) { (void)st; (void)cs; return; }
    // FAITHFUL == 4: no cube / cylinder / cone / CSG anywhere (every-ray prims are spheres and planes, BVH prims
    // are triangles and spheres): the generic leaf test is not in the kernel
    constexpr bool NO_GENERIC = FAITHFUL == 4;
    // ---- the remaining every-ray prims, then the BVH ---------------------------------------------------
    // "while-while" traversal: every lane first walks inner nodes until it holds a leaf (or runs out),
    // and only then does the warp test leaf prims — so the (long) leaf code runs with as many lanes as
    // have a leaf, instead of interleaving with other lanes' node tests. Items: a run of prims
    // [first, first+count) or a BVH node `cur`; refs on the stack are >= 0 for inner nodes and
    // ~(first << 3 | count) for leaves.
    constexpr int32_t NONE = -0x7fffffff - 1;
    int32_t first = k0, count = S.n_always - k0, cur = NONE;
    if (C.done()) return;
    for (;;) {
      // -- leaf phase
      for (int32_t k = first; k < first + count && !C.done(); k++) {
        int32_t i, par;
        rtc_node nd;
        LeafHits hits;
        bool generic = false;
        RTK_COUNT(prim_tests, 1);
        if (!NO_GENERIC && k < S.n_always) {  // (uniform per run) every-ray prim: one 128-byte record
          const PrimRec* rec = S.recs + k;
#if defined(__CUDA_ARCH__)
          const int4 hd = __ldg(reinterpret_cast<const int4*>(rec));
          i = hd.x; nd.aux = hd.y; par = hd.z;
          const uint32_t kf = (uint32_t)hd.w;
#else
          i = rec->node; nd.aux = rec->aux; par = rec->parent;
          const uint32_t kf = rec->kind_flags;
#endif
          nd.kind = (int32_t)(kf >> 24);
          nd.flags = kf & 0xffffffu;
          if (nd.kind == RTC_CSG) {
            if constexpr (HAS_CSG) {
              // the reference reaches this CSG only through its enclosing groups: validate first
              if (chain_ok(S, par, ray, ok_group)) walk_range<true>(S, ray, C, cs, err, i, RTC_LDG(&S.nodes[i].skip));
            }
            continue;
          }
          lr = ray_transform_v2(rec->inv, ray);
          lr_xf = -1;
          generic = true;
        } else {  // BVH prim: one 96-byte record, geometry inline (no node → triangle indirection)
          const LeafRec* lrec = S.leaf_recs + (k - S.n_always);
#if defined(__CUDA_ARCH__)
          const int4 hd = __ldg(reinterpret_cast<const int4*>(lrec));
          i = hd.x; par = hd.z;
          const int32_t xf = hd.y;
          const uint32_t kf = (uint32_t)hd.w;
#else
          i = lrec->node; par = lrec->parent;
          const int32_t xf = lrec->xform;
          const uint32_t kf = lrec->kind_flags;
#endif
          nd.kind = (int32_t)(kf >> 24);
          nd.flags = kf & 0xffffffu;
          if (xf != lr_xf) {
            lr = ray_transform(S.xforms + 12 * (size_t)xf, ray);
            lr_xf = xf;
          }
          hits.n = 0;
          hits.t0 = hits.t1 = hits.t2 = hits.t3 = 0.0;
          if (nd.kind == RTC_SPHERE) {
            isect_sphere(lr, [&](double t) { hits.push(t); });
          } else if (nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE) {
            double t, u, v;
            if (isect_triangle(lr, lrec->g, t, u, v)) hits.push(t);
          } else if (!NO_GENERIC) {  // cube, bounded cylinder
            nd.aux = RTC_LDG(&lrec->aux);
            generic = true;
          }
        }
        if (!NO_GENERIC && generic) {  // ONE isect_leaf site for every-ray prims and the BVH's non-triangle kinds
          LeafHits h;
          isect_leaf(S, nd, lr, h);
          hits = h;
        }
        const LeafHits& h = hits;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (int j = 0; j < h.n; j++) {
          const double t = j == 0 ? h.t0 : (j == 1 ? h.t1 : (j == 2 ? h.t2 : h.t3));
          cand(t, i, nd.flags, par);
        }
      }
      if (C.done() || S.n_bvh <= 0) return;
#if RTK_BVH_F32
      lim32 = (float)(C.limit + 2e-6 * (1.0 + fabs(C.limit)));
#endif
      if (!have_inv) {  // the always-run is done: enter the BVH at its root
#if RTK_BVH_F32
        V3 inv;
        bool plain;
        plain = plain_dir();
#else
        plain = ray.d.x != 0.0 && ray.d.y != 0.0 && ray.d.z != 0.0 && ray.d.x == ray.d.x && ray.d.y == ray.d.y && ray.d.z == ray.d.z;
#endif
        inv = v3(1.0 / ray.d.x, 1.0 / ray.d.y, 1.0 / ray.d.z);
        have_inv = true;
        if constexpr (!NO_GENERIC) {
          // a ray EPSILON-parallel to a face family of a cube in the BVH (CubeDir): no box bounds that cube's hits for
          // it, so it gets NO culling at all — a zero reciprocal direction makes every slab interval [0, 0]: every box
          // test passes and no t-range rejects (the walk visits the whole tree; such rays are a 1e-5 sliver of
          // directions, and scenes with more than 65 536 BVH prims keep their cubes out of the tree)
          bool cube_near = false;
          if (S.n_cube_dirs > 0) {
          const double slack = 1e-24 * dot(ray.d, ray.d);
          // (unrolled over the parameter-bank table: a loop index would make the compiler address SceneView dynamically)
#if defined(__CUDACC__)
#pragma unroll
#endif
          for (int k = 0; k < CUBE_DIR_MAX; k++) {
            const CubeDir& cd = S.cube_dirs[k];
            const double x = cd.ux * ray.d.x + cd.uy * ray.d.y + cd.uz * ray.d.z;
            cube_near = cube_near || (k < S.n_cube_dirs && x * x < cd.th2 + slack);
          }
          }
          if (cube_near) { inv = v3(0.0, 0.0, 0.0); plain = false; }
#if RTK_BVH_F32
          zero_inv = cube_near;
#endif
        }
#if RTK_BVH_F32
        {
          const double rx = S.bvh_ext[0], ry = S.bvh_ext[1], rz = S.bvh_ext[2];  // every box's |coordinate| bound
          const double cx = -(ray.o.x * inv.x), cy = -(ray.o.y * inv.y), cz = -(ray.o.z * inv.z);
          const double big = 1e15;
          fast32 = plain && fabs(inv.x) <= big && fabs(inv.y) <= big && fabs(inv.z) <= big && fabs(cx) <= big && fabs(cy) <= big &&
                   fabs(cz) <= big && rx <= big && ry <= big && rz <= big;
          F.ix = (float)inv.x; F.iy = (float)inv.y; F.iz = (float)inv.z;
          F.cx = (float)cx; F.cy = (float)cy; F.cz = (float)cz;
          F.ex = (float)((rx + fabs(ray.o.x)) * fabs(inv.x) * 5e-7);
          F.ey = (float)((ry + fabs(ray.o.y)) * fabs(inv.y) * 5e-7);
          F.ez = (float)((rz + fabs(ray.o.z)) * fabs(inv.z) * 5e-7);
        }
#endif
        cur = 0;
      } else {
        cur = NONE;
        while (sp > 0) {
          sp--;
          if (!RTK_BVH_CULL(RTK_ST_TMIN(sp), RTK_BVH_TINF)) { cur = RTK_ST_NODE(sp); break; }
        }
        if (cur == NONE) return;
      }
      // -- node phase: descend until this lane holds a leaf ref (cur < 0) or nothing (NONE)
      while (cur >= 0) {
        const TravNode* bn = S.bvh + cur;
        RTK_COUNT(bvh_nodes, 1);
#if RTK_BVH_WIDE
        // ---- four children per node: one 128-byte line in seven 16-byte loads
        float lx[4], ly[4], lz[4], hx[4], hy[4], hz[4];
        int32_t rf[4];
#if defined(__CUDA_ARCH__)
        {
          const float4* bq = reinterpret_cast<const float4*>(bn);
          const float4 q0 = __ldg(bq), q1 = __ldg(bq + 1), q2 = __ldg(bq + 2), q3 = __ldg(bq + 3), q4 = __ldg(bq + 4), q5 = __ldg(bq + 5);
          const int4 ri = __ldg(reinterpret_cast<const int4*>(bq + 6));
          lx[0] = q0.x; lx[1] = q0.y; lx[2] = q0.z; lx[3] = q0.w;
          ly[0] = q1.x; ly[1] = q1.y; ly[2] = q1.z; ly[3] = q1.w;
          lz[0] = q2.x; lz[1] = q2.y; lz[2] = q2.z; lz[3] = q2.w;
          hx[0] = q3.x; hx[1] = q3.y; hx[2] = q3.z; hx[3] = q3.w;
          hy[0] = q4.x; hy[1] = q4.y; hy[2] = q4.z; hy[3] = q4.w;
          hz[0] = q5.x; hz[1] = q5.y; hz[2] = q5.z; hz[3] = q5.w;
          rf[0] = ri.x; rf[1] = ri.y; rf[2] = ri.z; rf[3] = ri.w;
        }
#else
        for (int k = 0; k < 4; k++) {
          lx[k] = bn->lox[k]; ly[k] = bn->loy[k]; lz[k] = bn->loz[k];
          hx[k] = bn->hix[k]; hy[k] = bn->hiy[k]; hz[k] = bn->hiz[k];
          rf[k] = bn->ref[k];
        }
#endif
        // key: the child's entry parameter if it has to be visited, +inf if not
        float key[4];
        if (fast32) {
          const bool sx = F.ix < 0.0f, sy = F.iy < 0.0f, sz = F.iz < 0.0f;
#if defined(__CUDACC__)
#pragma unroll
#endif
          for (int k = 0; k < 4; k++) {
            float tmin, tmax;
            const bool in = F.test_nf(sx ? hx[k] : lx[k], sy ? hy[k] : ly[k], sz ? hz[k] : lz[k],
                                      sx ? lx[k] : hx[k], sy ? ly[k] : hy[k], sz ? lz[k] : hz[k], tmin, tmax);
            key[k] = (rf[k] != BVH_EMPTY && in && !cull32(tmin, tmax)) ? tmin : HUGE_VALF;
          }
        } else {
          // a zero / NaN / extreme direction component, or a cube-parallel ray (inv == 0): the double-precision test
          const bool plain = !zero_inv && plain_dir();
          const V3 inv = zero_inv ? v3(0.0, 0.0, 0.0) : v3(1.0 / ray.d.x, 1.0 / ray.d.y, 1.0 / ray.d.z);
#if defined(__CUDACC__)
#pragma unroll 1
#endif
          for (int k = 0; k < 4; k++) {
            double dmin = 0.0, dmax = 0.0;
            const bool in = rf[k] != BVH_EMPTY && own_box(lx[k], ly[k], lz[k], hx[k], hy[k], hz[k], ray, inv, plain, dmin, dmax) && !C.cull(dmin, dmax);
            // at or below the double; a NaN entry parameter orders first (the child must not be dropped)
            key[k] = in ? (dmin == dmin ? fminf(bvh_below(dmin), 3.0e38f) : -HUGE_VALF) : HUGE_VALF;
          }
        }
        // ascending by key (five compare-exchanges): the nearest child is next, the other wanted ones go on the stack
        // farthest first
#define RTK_CSWAP(a, b) do { const bool sw = key[b] < key[a]; const float tk = sw ? key[b] : key[a]; key[b] = sw ? key[a] : key[b]; key[a] = tk; \
                             const int32_t tr = sw ? rf[b] : rf[a]; rf[b] = sw ? rf[a] : rf[b]; rf[a] = tr; } while (0)
        RTK_CSWAP(0, 1); RTK_CSWAP(2, 3); RTK_CSWAP(0, 2); RTK_CSWAP(1, 3); RTK_CSWAP(1, 2);
#undef RTK_CSWAP
        cur = NONE;
        if (key[0] < HUGE_VALF) {
          cur = rf[0];
#if defined(__CUDACC__)
#pragma unroll
#endif
          for (int k = 3; k >= 1; k--) {
            if (key[k] < HUGE_VALF) {
              if (sp < BVH_STACK) {
                RTK_ST_PUSH(sp, rf[k], key[k]);
                sp++;
              } else {
                *err |= ERR_OVERFLOW;
              }
            }
          }
        } else {
          while (sp > 0) {
            sp--;
            if (!RTK_BVH_CULL(RTK_ST_TMIN(sp), RTK_BVH_TINF)) { cur = RTK_ST_NODE(sp); break; }
          }
        }
#else  // ---- two children per node
#if RTK_BVH_F32
#if defined(__CUDA_ARCH__)
        // the whole 64-byte node in four 16-byte loads
        const float4* bq = reinterpret_cast<const float4*>(bn);
        const float4 q0 = __ldg(bq), q1 = __ldg(bq + 1), q2 = __ldg(bq + 2);
        const int4 ci = __ldg(reinterpret_cast<const int4*>(bq + 3));
        const int32_t c0 = ci.x, n0 = ci.y, c1 = ci.z, n1 = ci.w;
        const float l0x = q0.x, l0y = q0.y, l0z = q0.z, h0x = q0.w, h0y = q1.x, h0z = q1.y;
        const float l1x = q1.z, l1y = q1.w, l1z = q2.x, h1x = q2.y, h1y = q2.z, h1z = q2.w;
#else
        const int32_t c0 = bn->c0, n0 = bn->n0, c1 = bn->c1, n1 = bn->n1;
        const float l0x = bn->lo0[0], l0y = bn->lo0[1], l0z = bn->lo0[2], h0x = bn->hi0[0], h0y = bn->hi0[1], h0z = bn->hi0[2];
        const float l1x = bn->lo1[0], l1y = bn->lo1[1], l1z = bn->lo1[2], h1x = bn->hi1[0], h1y = bn->hi1[1], h1z = bn->hi1[2];
#endif
        float t0min = 0.0f, t0max = 0.0f, t1min = 0.0f, t1max = 0.0f;
        bool h0, h1;
        if (fast32) {
          const bool b0 = F.test(l0x, l0y, l0z, h0x, h0y, h0z, t0min, t0max);
          const bool b1 = F.test(l1x, l1y, l1z, h1x, h1y, h1z, t1min, t1max);
          h0 = n0 >= 0 && b0 && !cull32(t0min, t0max);
          h1 = n1 >= 0 && b1 && !cull32(t1min, t1max);
        } else {
          // a zero / NaN / extreme direction component, or a cube-parallel ray (inv == 0): the double-precision test
          const bool plain = !zero_inv && plain_dir();
          const V3 inv = zero_inv ? v3(0.0, 0.0, 0.0) : v3(1.0 / ray.d.x, 1.0 / ray.d.y, 1.0 / ray.d.z);
          double d0min = 0.0, d0max = 0.0, d1min = 0.0, d1max = 0.0;
          h0 = n0 >= 0 && own_box(l0x, l0y, l0z, h0x, h0y, h0z, ray, inv, plain, d0min, d0max) && !C.cull(d0min, d0max);
          h1 = n1 >= 0 && own_box(l1x, l1y, l1z, h1x, h1y, h1z, ray, inv, plain, d1min, d1max) && !C.cull(d1min, d1max);
          // (for the stack and the near/far order: at or below the double; a NaN or inf stays what it is)
          t0min = bvh_below(d0min);
          t1min = bvh_below(d1min);
        }
#else
#if defined(__CUDA_ARCH__)
        // the whole 128-byte node in seven 16-byte loads
        const double2* bq = reinterpret_cast<const double2*>(bn);
        const double2 q0 = __ldg(bq), q1 = __ldg(bq + 1), q2 = __ldg(bq + 2), q3 = __ldg(bq + 3), q4 = __ldg(bq + 4), q5 = __ldg(bq + 5);
        const int4 ci = __ldg(reinterpret_cast<const int4*>(&bn->c0));
        const int32_t c0 = ci.x, n0 = ci.y, c1 = ci.z, n1 = ci.w;
        const double l0x = q0.x, l0y = q0.y, l0z = q1.x, h0x = q1.y, h0y = q2.x, h0z = q2.y;
        const double l1x = q3.x, l1y = q3.y, l1z = q4.x, h1x = q4.y, h1y = q5.x, h1z = q5.y;
#else
        const int32_t c0 = bn->c0, n0 = bn->n0, c1 = bn->c1, n1 = bn->n1;
        const double l0x = bn->lo0[0], l0y = bn->lo0[1], l0z = bn->lo0[2], h0x = bn->hi0[0], h0y = bn->hi0[1], h0z = bn->hi0[2];
        const double l1x = bn->lo1[0], l1y = bn->lo1[1], l1z = bn->lo1[2], h1x = bn->hi1[0], h1y = bn->hi1[1], h1z = bn->hi1[2];
#endif
        double t0min = 0.0, t0max = 0.0, t1min = 0.0, t1max = 0.0;
        const bool h0 = n0 >= 0 && own_box(l0x, l0y, l0z, h0x, h0y, h0z, ray, inv, plain, t0min, t0max) && !C.cull(t0min, t0max);
        const bool h1 = n1 >= 0 && own_box(l1x, l1y, l1z, h1x, h1y, h1z, ray, inv, plain, t1min, t1max) && !C.cull(t1min, t1max);
#endif
        // refs: >= 0 inner node, < 0 leaf ~(first << 3 | count)
        const int32_t r0 = n0 > 0 ? ~((c0 << 3) | n0) : c0;
        const int32_t r1 = n1 > 0 ? ~((c1 << 3) | n1) : c1;
        cur = NONE;
        if (h0 && h1) {
          const bool swap = t1min < t0min;  // near child first, far child on the stack with its tmin
          if (sp < BVH_STACK) {
            RTK_ST_PUSH(sp, swap ? r0 : r1, swap ? t0min : t1min);
            sp++;
          } else {
            *err |= ERR_OVERFLOW;
          }
          cur = swap ? r1 : r0;
        } else if (h0 || h1) {
          cur = h0 ? r0 : r1;
        } else {
          while (sp > 0) {
            sp--;
            if (!RTK_BVH_CULL(RTK_ST_TMIN(sp), RTK_BVH_TINF)) { cur = RTK_ST_NODE(sp); break; }
          }
        }
#endif  // RTK_BVH_WIDE
      }
      if (cur == NONE) return;
      {
        const int32_t ref = ~cur;
        first = ref >> 3;
        count = ref & 7;
      }
    }
  }
}

}  // namespace rtk
