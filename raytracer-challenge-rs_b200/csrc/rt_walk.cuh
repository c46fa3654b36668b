// rt_walk.cuh — scene traversal: World::intersect (world.rs:31-38), Group::intersect
// (group.rs:49-58) and Csg::local_intersect + filter_intersections (csg.rs:71-91,115-126) as ONE
// stackless pre-order walk over the flattened node array (skip pointers), feeding a consumer.
//
// The reference collects every intersection of every object into a Vec, sorts it by t and then
// looks at the front; the consumers below compute the same answers from the stream:
//   ClosestHit  — hit(): first t >= 0 of the sorted list (intersection.rs:118-120). Ties in t are
//                 resolved like a stable sort: first in traversal order wins (strict <).
//   ShadowHit   — shadow_hit() + `t < distance` (intersection.rs:122-125, world.rs:101-104):
//                 any has_shadow intersection with 0 <= t < distance, early exit.
//   NegCollect  — the sorted prefix of the list that precedes the hit, i.e. every t < 0, which is
//                 all prepare_computations' containers loop ever looks at (intersection.rs:63-90).
// Inside a CSG subtree intersections go to a bounded per-thread buffer which is sorted (stable) and
// filtered when the walk leaves that CSG node — exactly the reference's per-CSG sort and filter,
// nested CSGs included; survivors of the outermost CSG are then streamed to the consumer at the
// CSG's position in the traversal order.
#pragma once
#include "rt_isect.cuh"

namespace rtk {

constexpr int CSG_MAX_DEPTH = 4;    // nested CSG levels per ray
constexpr int HIT_CAP = 64;         // bounded list capacity (CSG hit buffer, negative-t list)

enum : uint32_t { ERR_OVERFLOW = 1u, ERR_NAN = 2u };

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

// Sorted bounded list of (t, node); insertion keeps equal t in arrival order (stable).
struct HitList {
  double t[HIT_CAP];
  int32_t node[HIT_CAP];
  int32_t n;
};

// csg.rs:22-40
RTC_HD bool csg_allowed(uint32_t op, bool lhit, bool inl, bool inr) {
  if (op == RTC_CSG_UNION) return (lhit && !inr) || (!lhit && !inl);
  if (op == RTC_CSG_INTERSECTION) return (lhit && inr) || (!lhit && inl);
  return (lhit && !inr) || (!lhit && inl);
}

struct ClosestHit {
  double best_t;
  int32_t best_node;
  RTC_HD void init() { best_t = rt_inf(); best_node = -1; }
  RTC_HD void hit(double t, int32_t node, uint32_t) {
    if (t >= 0.0 && t < best_t) { best_t = t; best_node = node; }
  }
  RTC_HD bool done() const { return false; }
  // conservative t-range culling of a group box [tmin,tmax] (not in the reference; never removes
  // an intersection the reference would accept — see DESIGN.md "t-range culling")
  RTC_HD bool cull(double tmin, double tmax) const {
    return tmax < -1e-6 * (1.0 + fabs(tmax)) || tmin > best_t + 1e-6 * (1.0 + fabs(best_t));
  }
};

struct ShadowHit {
  double distance;
  bool found;
  RTC_HD void init(double d) { distance = d; found = false; }
  RTC_HD void hit(double t, int32_t, uint32_t flags) {
    if (t >= 0.0 && t < distance && (flags & RTC_FLAG_SHADOW)) found = true;
  }
  RTC_HD bool done() const { return found; }
  RTC_HD bool cull(double tmin, double tmax) const {
    return tmax < -1e-6 * (1.0 + fabs(tmax)) || tmin > distance + 1e-6 * (1.0 + fabs(distance));
  }
};

struct NegCollect {
  HitList* list;
  uint32_t* err;
  RTC_HD void init(HitList* l, uint32_t* e) { list = l; err = e; l->n = 0; }
  RTC_HD void hit(double t, int32_t node, uint32_t) {
    if (!(t < 0.0)) return;
    int n = list->n;
    if (n >= HIT_CAP) { *err |= ERR_OVERFLOW; return; }
    int k = n;
    while (k > 0 && list->t[k - 1] > t) {  // stable: stop at equal
      list->t[k] = list->t[k - 1];
      list->node[k] = list->node[k - 1];
      k--;
    }
    list->t[k] = t;
    list->node[k] = node;
    list->n = n + 1;
  }
  RTC_HD bool done() const { return false; }
  RTC_HD bool cull(double tmin, double) const { return tmin > 1e-6 * (1.0 + fabs(tmin)); }
};

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

// The walk. FAITHFUL=true visits exactly the nodes the reference visits (no t-range culling).
template <bool HAS_CSG, bool FAITHFUL, class Consumer>
RTC_HD void walk_scene(const SceneView& S, const RayD& world_ray, Consumer& C, CsgScratch* cs, uint32_t* err) {
  RayD cur = world_ray;
  RayD lr;               // object-space ray cache, keyed by transform id
  int32_t lr_xf = -1;
  int depth = 0;
  if (HAS_CSG) cs->buf.n = 0;
  const int32_t n_nodes = S.n_nodes;
  int32_t i = 0;
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
      if (ok && !FAITHFUL && (!HAS_CSG || depth == 0) && C.cull(tmin, tmax)) ok = false;
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
      if (HAS_CSG && depth > 0) {
        HitList& buf = cs->buf;
        isect_leaf(S, nd, lr, [&](double t) {
          if (t != t) *err |= ERR_NAN;
          if (buf.n >= HIT_CAP) { *err |= ERR_OVERFLOW; return; }
          buf.t[buf.n] = t;
          buf.node[buf.n] = i;
          buf.n++;
        });
      } else {
        isect_leaf(S, nd, lr, [&](double t) {
          if (t != t) *err |= ERR_NAN;
          C.hit(t, i, nd.flags);
        });
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
