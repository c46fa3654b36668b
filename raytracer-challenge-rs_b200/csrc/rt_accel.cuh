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
struct AccelStats { unsigned long long walks, bvh_nodes, prim_tests, chain_calls, chain_boxes, always_tests; };
inline AccelStats& accel_stats() { static AccelStats s; return s; }
#define RTK_COUNT(f, n) (accel_stats().f += (n))
#else
#define RTK_COUNT(f, n) ((void)0)
#endif

constexpr int BVH_STACK = 64;  // build bounds the depth (SAH to depth 40, then median splits)

struct BvhStack {
  int32_t node[BVH_STACK];
  double tmin[BVH_STACK];
};

// Has `ray` passed the box of every reference Group enclosing `node`? `ok_group`: innermost group
// already validated for this ray (its ancestors are then valid too), -1 initially (= top level).
RTC_HD bool chain_ok(const SceneView& S, int32_t node, const RayD& ray, int32_t& ok_group) {
  int32_t g0 = RTC_LDG(S.parent + node);
  int32_t g = g0;
  RTK_COUNT(chain_calls, 1);
  while (g >= 0 && g != ok_group) {
    double tmin, tmax;
    RTK_COUNT(chain_boxes, 1);
    int32_t bx = RTC_LDG(&S.nodes[g].aux);
    if (!box_intersects(S.boxes + 6 * (size_t)bx, ray, tmin, tmax)) return false;
    g = RTC_LDG(S.parent + g);
  }
  ok_group = g0;
  return true;
}

// Slab test against one of OUR boxes with the reciprocal direction. NaN (0*inf) drops out of
// fmax/fmin, which is the conservative outcome; the boxes carry a guard band (build).
RTC_HD bool own_box(const double* lo, const double* hi, const RayD& r, V3 inv, double& tmin, double& tmax) {
  double ax = (RTC_LDG(lo + 0) - r.o.x) * inv.x, bx = (RTC_LDG(hi + 0) - r.o.x) * inv.x;
  double ay = (RTC_LDG(lo + 1) - r.o.y) * inv.y, by = (RTC_LDG(hi + 1) - r.o.y) * inv.y;
  double az = (RTC_LDG(lo + 2) - r.o.z) * inv.z, bz = (RTC_LDG(hi + 2) - r.o.z) * inv.z;
  tmin = fmax(fmax(fmin(ax, bx), fmin(ay, by)), fmin(az, bz));
  tmax = fmin(fmin(fmax(ax, bx), fmax(ay, by)), fmax(az, bz));
  // relative slack for the reciprocal's rounding
  return tmin <= tmax + 1e-9 * (fabs(tmin) + fabs(tmax));
}

// One candidate intersection of leaf `i` (chain node `i`): NaN bookkeeping, want → validate → take.
RTC_HD void offer(const SceneView& S, const RayD& ray, Query& C, double t, int32_t i, uint32_t flags, int32_t& ok_group,
                  uint32_t* err) {
  if (t != t) { if (chain_ok(S, i, ray, ok_group)) *err |= ERR_NAN; return; }
  if (!C.want(t, i, flags)) return;
  if (!chain_ok(S, i, ray, ok_group)) return;
  C.take(t, i);
}

RTC_HD void test_leaf(const SceneView& S, const RayD& ray, Query& C, int32_t i, const rtc_node& nd, RayD& lr, int32_t& lr_xf,
                      int32_t& ok_group, uint32_t* err) {
  if (nd.xform != lr_xf) {
    lr = ray_transform(S.xforms + 12 * (size_t)nd.xform, ray);
    lr_xf = nd.xform;
  }
  LeafHits h;
  isect_leaf(S, nd, lr, h);
  if (h.n > 0) offer(S, ray, C, h.t0, i, nd.flags, ok_group, err);
  if (h.n > 1) offer(S, ray, C, h.t1, i, nd.flags, ok_group, err);
  if (h.n > 2) offer(S, ray, C, h.t2, i, nd.flags, ok_group, err);
  if (h.n > 3) offer(S, ray, C, h.t3, i, nd.flags, ok_group, err);
}

// World::intersect → query. faithful: the reference's own search order, node by node.
template <bool HAS_CSG>
RTC_HD void search(const SceneView& S, const RayD& ray, Query& C, CsgScratch* cs, BvhStack* st, bool faithful, uint32_t* err) {
  int32_t ok_group = -1;
  RayD lr;
  int32_t lr_xf = -1;
  RTK_COUNT(walks, 1);

  // ---- prims tested for every ray (faithful: one "prim" = the whole node array) ------------------
  const int32_t n_always = faithful ? 1 : S.n_always;
  for (int32_t k = 0; k < n_always && !C.done(); k++) {
    int32_t i = 0, end = S.n_nodes;
    bool range = true;
    rtc_node nd;
    if (!faithful) {
      i = RTC_LDG(S.always + k);
      nd = load_node(S.nodes + i);
      RTK_COUNT(always_tests, 1);
      range = nd.kind == RTC_CSG;
      end = nd.skip;
      // the reference reaches this CSG only through its enclosing groups: validate first
      if (range && !chain_ok(S, i, ray, ok_group)) continue;
    }
    if (range) {
      walk_range<HAS_CSG>(S, ray, C, cs, err, i, end);
      continue;
    }
    test_leaf(S, ray, C, i, nd, lr, lr_xf, ok_group, err);
  }
  if (faithful || S.n_bvh <= 0 || C.done()) return;

  // ---- BVH over triangles and spheres: refs >= 0 inner node, < 0 leaf ~(first << 3 | count) --------
  const V3 inv = v3(1.0 / ray.d.x, 1.0 / ray.d.y, 1.0 / ray.d.z);
  int sp = 0;
  int32_t cur = 0;
  for (;;) {
    if (cur < 0) {
      const int32_t ref = ~cur, first = ref >> 3, count = ref & 7;
      for (int32_t k = first; k < first + count; k++) {
        const int32_t i = RTC_LDG(S.bvh_prims + k);
        rtc_node nd = load_node(S.nodes + i);
        RTK_COUNT(prim_tests, 1);
        test_leaf(S, ray, C, i, nd, lr, lr_xf, ok_group, err);
      }
      if (C.done()) return;
      cur = 0;
      bool got = false;
      while (sp > 0) {
        sp--;
        if (!C.cull(st->tmin[sp], rt_inf())) { cur = st->node[sp]; got = true; break; }
      }
      if (!got) return;
      continue;
    }
    const BvhNode* bn = S.bvh + cur;
    RTK_COUNT(bvh_nodes, 1);
#if defined(__CUDA_ARCH__)
    const int4 ci = __ldg(reinterpret_cast<const int4*>(&bn->c0));
    const int32_t c0 = ci.x, n0 = ci.y, c1 = ci.z, n1 = ci.w;
#else
    const int32_t c0 = bn->c0, n0 = bn->n0, c1 = bn->c1, n1 = bn->n1;
#endif
    double t0min = 0.0, t0max = 0.0, t1min = 0.0, t1max = 0.0;
    const bool h0 = n0 >= 0 && own_box(bn->lo0, bn->hi0, ray, inv, t0min, t0max) && !C.cull(t0min, t0max);
    const bool h1 = n1 >= 0 && own_box(bn->lo1, bn->hi1, ray, inv, t1min, t1max) && !C.cull(t1min, t1max);
    const int32_t r0 = n0 > 0 ? ~((c0 << 3) | n0) : c0;
    const int32_t r1 = n1 > 0 ? ~((c1 << 3) | n1) : c1;
    if (h0 && h1) {
      const bool swap = t1min < t0min;   // near child first, far child on the stack with its tmin
      if (sp < BVH_STACK) {
        st->node[sp] = swap ? r0 : r1;
        st->tmin[sp] = swap ? t0min : t1min;
        sp++;
      } else {
        *err |= ERR_OVERFLOW;
      }
      cur = swap ? r1 : r0;
    } else if (h0 || h1) {
      cur = h0 ? r0 : r1;
    } else {
      bool got = false;
      while (sp > 0) {
        sp--;
        if (!C.cull(st->tmin[sp], rt_inf())) { cur = st->node[sp]; got = true; break; }
      }
      if (!got) return;
    }
  }
}

}  // namespace rtk
