// rt_walk.cuh — the query a ray puts to the world, and the reference-order walk.
//
// The reference collects every intersection of every object into a Vec, sorts it by t and then
// looks at the front (world.rs:31-38, intersection.rs:108-125). Nothing is collected here: a
// `Query` says what the caller would read off that sorted list, and the searches stream
// intersections into it:
//   Q_CLOSEST  hit(): first t >= 0 of the sorted list (intersection.rs:118-120). Equal t resolves to
//              the lower pre-order node index = the entry a stable sort of the reference's
//              traversal-ordered list puts first.
//   Q_SHADOW   shadow_hit() + `t < distance` (intersection.rs:122-125, world.rs:101-104): any
//              has_shadow intersection with 0 <= t < distance; early exit.
//   Q_NEG      the sorted prefix that precedes the hit, i.e. every t < 0 — all the containers loop
//              of prepare_computations ever looks at (intersection.rs:63-90).
// walk_range() is the reference's own search: depth-first, in order, over a range of the flattened
// node array (World::intersect for the whole array, Group::intersect with its box test
// group.rs:49-58, Csg::local_intersect + filter_intersections csg.rs:71-126 with a bounded
// per-thread buffer that is stable-sorted and filtered when the walk leaves the CSG node).
#pragma once
#include "rt_isect.cuh"

namespace rtk {

constexpr int CSG_MAX_DEPTH = 4;    // nested CSG levels per ray
constexpr int HIT_CAP = 64;         // bounded list capacity (CSG hit buffer, negative-t list)

enum : uint32_t { ERR_OVERFLOW = 1u, ERR_NAN = 2u };
enum : int32_t { Q_CLOSEST = 0, Q_SHADOW = 1, Q_NEG = 2, Q_DONE = 3 };  // Q_DONE: a Q_SHADOW that found its occluder

RTC_HD rtc_node load_node(const rtc_node* p) {
#if defined(__CUDA_ARCH__)
  const int4* q = reinterpret_cast<const int4*>(p);
  int4 a = __ldg(q), b = __ldg(q + 1);
  rtc_node n;
  n.kind = a.x; n.skip = a.y; n.xform = a.z; n.material = a.w;
  n.aux = b.x; n.flags = (uint32_t)b.y; n.right = b.z; n.eq_class = b.w;
  return n;
#else
  return *p;
#endif
}

// Sorted bounded list of (t, node).
struct HitList {
  double t[HIT_CAP];
  int32_t node[HIT_CAP];
  int32_t n;
};

struct Query {
  double limit;      // Q_CLOSEST: best t so far (inf); Q_SHADOW: distance to the light; Q_NEG: inf
  double limit_hi;   // limit * (1 + 1e-9): what a not-yet-divided quotient is compared with (quot_unwanted)
  int32_t mode;
  int32_t node;      // Q_CLOSEST: best node (-1); Q_SHADOW: 1 when found
  HitList* list;     // Q_NEG
  uint32_t* err;

  RTC_HD void init_closest(uint32_t* e) { mode = Q_CLOSEST; limit = limit_hi = rt_inf(); node = -1; list = nullptr; err = e; }
  RTC_HD void init_shadow(double distance, uint32_t* e) {
    mode = Q_SHADOW; limit = distance; limit_hi = distance * (1.0 + 1e-9); node = 0; list = nullptr; err = e;
  }
  RTC_HD void init_neg(HitList* l, uint32_t* e) { mode = Q_NEG; limit = limit_hi = rt_inf(); node = -1; list = l; l->n = 0; err = e; }

  // would the reference's sorted list hand this intersection to the caller? One branch-free form for
  // the three modes (lanes of a warp are in different modes): the sign test flips for Q_NEG, whose
  // limit is +inf; the tie `t == limit` can only win in Q_CLOSEST (node is 0/1 in Q_SHADOW, -1 in
  // Q_NEG, never above a node index); only Q_SHADOW asks for the has_shadow flag.
  RTC_HD bool want(double t, int32_t nd, uint32_t flags) const {
#if defined(RTK_QUERY_BRANCHY)
    if (mode == Q_CLOSEST) return t >= 0.0 && (t < limit || (t == limit && nd < node));
    if (mode == Q_SHADOW) return t >= 0.0 && t < limit && (flags & RTC_FLAG_SHADOW);
    return t < 0.0;
#else
    const bool side = (t >= 0.0) != (mode == Q_NEG);
    const bool within = (t < limit) | ((t == limit) & (nd < node));
    const uint32_t need = mode == Q_SHADOW ? (uint32_t)RTC_FLAG_SHADOW : 0u;
    return side & within & ((flags & need) == need) & (t == t);
#endif
  }
  RTC_HD void take(double t, int32_t nd) {
    if (mode == Q_CLOSEST) { limit = t; limit_hi = t * (1.0 + 1e-9); node = nd; return; }
    if (mode & 1) { node = 1; mode = Q_DONE; return; }  // Q_SHADOW, or Q_DONE again (further hits of the same leaf)
    int n = list->n;
    if (n >= HIT_CAP) { *err |= ERR_OVERFLOW; return; }
    int k = n;
    // sorted by (t, node): the order of the reference's stable-sorted list
    while (k > 0 && (list->t[k - 1] > t || (list->t[k - 1] == t && list->node[k - 1] > nd))) {
      list->t[k] = list->t[k - 1];
      list->node[k] = list->node[k - 1];
      k--;
    }
    list->t[k] = t;
    list->node[k] = nd;
    list->n = n + 1;
  }
  RTC_HD void hit(double t, int32_t nd, uint32_t flags) { if (want(t, nd, flags)) take(t, nd); }
  RTC_HD bool done() const { return mode == Q_DONE; }
  // conservative t-range culling of one of OUR boxes [tmin,tmax] (never of a reference box)
  RTC_HD bool cull(double tmin, double tmax) const {
    if (mode == Q_NEG) return tmin > 1e-6 * (1.0 + fabs(tmin));
    return tmax < -1e-6 * (1.0 + fabs(tmax)) || tmin > limit + 1e-6 * (1.0 + fabs(limit));
  }
};

// csg.rs:22-40
RTC_HD bool csg_allowed(uint32_t op, bool lhit, bool inl, bool inr) {
  if (op == RTC_CSG_UNION) return (lhit && !inr) || (!lhit && !inl);
  if (op == RTC_CSG_INTERSECTION) return (lhit && inr) || (!lhit && inl);
  return (lhit && !inr) || (!lhit && inl);
}

struct CsgLevel {
  RayD saved;      // the ray outside this CSG
  int32_t node;    // CSG node index
  int32_t skip;    // first node after the CSG subtree
  int32_t start;   // first buffer slot of this CSG's intersections
};

// Scratch for the walk; only instantiated with HAS_CSG (no-CSG scenes carry none of it).
struct CsgScratch {
  HitList buf;
  CsgLevel level[CSG_MAX_DEPTH];
};

// Sort (stable insertion) + filter buf[start, n) for CSG node `ci`; returns the new n.
RTC_HD int csg_finalize(const SceneView& S, HitList& buf, int start, int32_t ci, uint32_t* err) {
  int n = buf.n;
  for (int a = start + 1; a < n; a++) {
    double t = buf.t[a];
    int32_t nd = buf.node[a];
    int k = a;
    while (k > start && buf.t[k - 1] > t) {
      buf.t[k] = buf.t[k - 1];
      buf.node[k] = buf.node[k - 1];
      k--;
    }
    buf.t[k] = t;
    buf.node[k] = nd;
  }
  rtc_node cn = load_node(S.nodes + ci);
  uint32_t op = (cn.flags >> 8) & 3u;
  bool inl = false, inr = false;
  int w = start;
  for (int a = start; a < n; a++) {
    // left.includes(object): pointer identity over the left subtree (geometry/mod.rs:87-89,
    // group.rs:104-106, csg.rs:132-134) == node index before the right subtree's root
    bool lhit = buf.node[a] < cn.right;
    if (csg_allowed(op, lhit, inl, inr)) {
      buf.t[w] = buf.t[a];
      buf.node[w] = buf.node[a];
      w++;
    }
    if (lhit) inl = !inl; else inr = !inr;
  }
  (void)err;
  return w;
}

// The in-order walk over nodes [begin, end): visits exactly the nodes the reference visits
// (World::intersect for the whole array; one CSG subtree when called by the accelerated search).
template <bool HAS_CSG>
RTC_HD void walk_range(const SceneView& S, const RayD& world_ray, Query& C, CsgScratch* cs, uint32_t* err,
                       int32_t begin, int32_t end) {
  RayD cur = world_ray;
  RayD lr;               // object-space ray cache, keyed by transform id
  int32_t lr_xf = -1;
  int depth = 0;
  if (HAS_CSG) cs->buf.n = 0;
  const int32_t n_nodes = end;
  int32_t i = begin;
  for (;;) {
    if (HAS_CSG) {
      while (depth > 0 && i == cs->level[depth - 1].skip) {
        CsgLevel& L = cs->level[depth - 1];
        int w = csg_finalize(S, cs->buf, L.start, L.node, err);
        cur = L.saved;
        lr_xf = -1;
        depth--;
        if (depth == 0) {
          for (int a = 0; a < w; a++) {
            int32_t hn = cs->buf.node[a];
            C.hit(cs->buf.t[a], hn, RTC_LDG(&S.nodes[hn].flags));
          }
          cs->buf.n = 0;
        } else {
          cs->buf.n = w;
        }
      }
    }
    if (i >= n_nodes || C.done()) break;
    rtc_node nd = load_node(S.nodes + i);
    if (nd.kind == RTC_GROUP) {
      double tmin, tmax;
      bool ok = box_intersects(S.boxes + 6 * (size_t)nd.aux, cur, tmin, tmax);
      i = ok ? i + 1 : nd.skip;
    } else if (nd.kind == RTC_CSG) {
      if (HAS_CSG) {
        // Shape::intersect default (geometry/mod.rs:46-49): ray → CSG space, then the CSG's box
        // (kept in PARENT space by set_transform — reference quirk) against that LOCAL ray.
        RayD local = ray_transform(S.xforms + 12 * (size_t)nd.xform, cur);
        double tmin, tmax;
        if (!box_intersects(S.boxes + 6 * (size_t)nd.aux, local, tmin, tmax)) {
          i = nd.skip;
        } else if (depth >= CSG_MAX_DEPTH) {
          *err |= ERR_OVERFLOW;
          i = nd.skip;
        } else {
          CsgLevel& L = cs->level[depth];
          L.saved = cur;
          L.node = i;
          L.skip = nd.skip;
          L.start = cs->buf.n;
          depth++;
          cur = local;
          lr_xf = -1;
          i = i + 1;
        }
      } else {
        i = nd.skip;
      }
    } else {
      if (nd.xform != lr_xf) {
        lr = ray_transform(S.xforms + 12 * (size_t)nd.xform, cur);
        lr_xf = nd.xform;
      }
      LeafHits h;
      isect_leaf(S, nd, lr, h);
#if defined(__CUDACC__)
#pragma unroll
#endif
      for (int k = 0; k < 4; k++) {
        if (k < h.n) {
          const double t = k == 0 ? h.t0 : (k == 1 ? h.t1 : (k == 2 ? h.t2 : h.t3));
          if (t != t) *err |= ERR_NAN;
          if (HAS_CSG && depth > 0) {
            HitList& buf = cs->buf;
            if (buf.n >= HIT_CAP) { *err |= ERR_OVERFLOW; }
            else { buf.t[buf.n] = t; buf.node[buf.n] = i; buf.n++; }
          } else {
            C.hit(t, i, nd.flags);
          }
        }
      }
      i = i + 1;
    }
  }
}

// Object-space ray of leaf `node` for a world ray: the chain of enclosing CSG inverses
// (outermost first), then the leaf's own inverse. Used to recompute u,v of a smooth-triangle hit.
RTC_HD RayD leaf_local_ray(const SceneView& S, int32_t node, const RayD& world_ray) {
  int32_t chain[CSG_MAX_DEPTH];
  int n = 0;
  if (S.has_csg) {
    int32_t p = RTC_LDG(S.csg_parent + node);
    while (p >= 0 && n < CSG_MAX_DEPTH) {
      chain[n++] = p;
      p = RTC_LDG(S.csg_parent + p);
    }
  }
  RayD r = world_ray;
  for (int k = n - 1; k >= 0; k--) r = ray_transform(S.xforms + 12 * (size_t)RTC_LDG(&S.nodes[chain[k]].xform), r);
  return ray_transform(S.xforms + 12 * (size_t)RTC_LDG(&S.nodes[node].xform), r);
}

}  // namespace rtk
