// rt_shade.cuh — hit-point precompute, Phong lighting, shadows and the Whitted recursion rewritten
// as a bounded iterative loop with an explicit per-thread frame stack that keeps the reference's
// exact evaluation ORDER and ASSOCIATION:  (surface + reflected·k) + refracted·k'.
//
// Reference (relative to raytracer/src): world.rs:40-134 (shade_hit, color_at, is_shadowed,
// reflected_color, refracted_color), geometry/intersection.rs:53-105,147-162 (prepare_computations,
// schlick), geometry/mod.rs:51-56 (normal_at), material.rs:38-82 (lighting), pattern/*.rs,
// camera.rs:57-126 (ray_for_pixel, rays_for_pixel, get_offsets), color.rs:26-33 (average).
#pragma once
#include "rt_accel.cuh"

namespace rtk {

constexpr int MAX_FRAMES = 8;  // max_depth + 1 <= MAX_FRAMES

struct RayCounters {
  uint32_t primary, shadow, reflect, refract;
};

struct Comps {  // intersection.rs:132-144
  V3 point, over_point, under_point, eyev, normalv, reflectv;
  double t, n1, n2;
  int32_t node;
  bool inside;
};

// Per-thread scratch that lives in local memory.
template <bool HAS_CSG>
struct Scratch;
template <>
struct Scratch<false> {
  HitList neg;
  BvhStack bvh;
  RTC_HD CsgScratch* csg() { return nullptr; }
};
template <>
struct Scratch<true> {
  HitList neg;
  BvhStack bvh;
  CsgScratch cs;
  RTC_HD CsgScratch* csg() { return &cs; }
};

// *_local_normal_at for every leaf kind (sphere.rs:64, plane.rs:62, cube.rs:95-105,
// cylinder.rs:121-130, cone.rs:136-149, triangle.rs:92, smooth_triangle.rs:98-102).
RTC_HD V3 local_normal(const SceneView& S, const rtc_node& nd, V3 p, int32_t node, const RayD& world_ray) {
  switch (nd.kind) {
    case RTC_SPHERE: return v3(p.x - 0.0, p.y - 0.0, p.z - 0.0);
    case RTC_PLANE: return v3(0.0, 1.0, 0.0);
    case RTC_CUBE: {
      double ax = fabs(p.x), ay = fabs(p.y), az = fabs(p.z);
      double maxc = fmax(fmax(ax, ay), az);
      if (feq(maxc, ax)) return v3(p.x, 0.0, 0.0);
      if (feq(maxc, ay)) return v3(0.0, p.y, 0.0);
      return v3(0.0, 0.0, p.z);
    }
    case RTC_CYLINDER: {
      const double* lim = S.limits + 2 * nd.aux;
      double dist = p.x * p.x + p.z * p.z;
      if (dist < 1.0 && p.y >= RTC_LDG(lim + 1) - EPS) return v3(0.0, 1.0, 0.0);
      if (dist < 1.0 && p.y <= RTC_LDG(lim) + EPS) return v3(0.0, -1.0, 0.0);
      return v3(p.x, 0.0, p.z);
    }
    case RTC_CONE: {
      const double* lim = S.limits + 2 * nd.aux;
      double dist = p.x * p.x + p.z * p.z;
      if (dist < 1.0 && p.y >= RTC_LDG(lim + 1) - EPS) return v3(0.0, 1.0, 0.0);  // dist<1.0: cone.rs:138
      if (dist < 1.0 && p.y <= RTC_LDG(lim) + EPS) return v3(0.0, -1.0, 0.0);
      double y = sqrt(p.x * p.x + p.z * p.z);
      if (p.y > 0.0) y = -y;
      return v3(p.x, y, p.z);
    }
    case RTC_TRIANGLE: {
      const double* nn = S.tri_normals + 9 * (size_t)nd.aux;
      return v3(RTC_LDG(nn), RTC_LDG(nn + 1), RTC_LDG(nn + 2));
    }
    case RTC_SMOOTH_TRIANGLE: {
      // u,v of the hit: re-run the (deterministic) triangle test on the same object-space ray
      RayD lr = leaf_local_ray(S, node, world_ray);
      double t, u = 0.0, v = 0.0;
      isect_triangle(lr, S.tri_geom + 9 * (size_t)nd.aux, t, u, v);
      const double* nn = S.tri_normals + 9 * (size_t)nd.aux;
      V3 n1 = v3(RTC_LDG(nn), RTC_LDG(nn + 1), RTC_LDG(nn + 2));
      V3 n2 = v3(RTC_LDG(nn + 3), RTC_LDG(nn + 4), RTC_LDG(nn + 5));
      V3 n3 = v3(RTC_LDG(nn + 6), RTC_LDG(nn + 7), RTC_LDG(nn + 8));
      return n2 * u + n3 * v + n1 * (1.0 - u - v);
    }
    default: return v3(0.0, 0.0, 0.0);
  }
}

// Containers walk for n1/n2 (intersection.rs:61-90). Every list entry before the hit has t < 0;
// `neg` holds them sorted. Objects toggle by VALUE equality → eq_class.
RTC_HD void containers_n1n2(const SceneView& S, const HitList& neg, int32_t hit_node, double& n1, double& n2,
                            uint32_t* err) {
  int32_t cont[HIT_CAP];
  int nc = 0;
  auto toggle = [&](int32_t node) {
    int32_t cls = RTC_LDG(&S.nodes[node].eq_class);
    int k = 0;
    for (; k < nc; k++)
      if (RTC_LDG(&S.nodes[cont[k]].eq_class) == cls) break;
    if (k < nc) {
      for (int m = k; m + 1 < nc; m++) cont[m] = cont[m + 1];
      nc--;
    } else if (nc < HIT_CAP) {
      cont[nc++] = node;
    } else {
      *err |= ERR_OVERFLOW;
    }
  };
  auto last_index = [&]() -> double {
    if (nc == 0) return 1.0;
    int32_t m = RTC_LDG(&S.nodes[cont[nc - 1]].material);
    return RTC_LDG(&S.materials[m].refractive_index);
  };
  for (int a = 0; a < neg.n; a++) toggle(neg.node[a]);
  n1 = last_index();
  toggle(hit_node);
  n2 = last_index();
}

// Intersection::prepare_computations (intersection.rs:53-105) without the containers part.
RTC_HD void prepare(const SceneView& S, const RayD& ray, double t, int32_t node, const rtc_node& nd, Comps& c) {
  c.t = t;
  c.node = node;
  c.point = ray_position(ray, t);
  c.eyev = -ray.d;
  const M34 inv = load_m34(S.xforms + 12 * (size_t)nd.xform);
  V3 local_point = xf_point(inv, c.point);                       // geometry/mod.rs:52
  V3 ln = local_normal(S, nd, local_point, node, ray);           // :53
  V3 n = normalize(xf_vector_transposed(inv, ln));               // :54-55
  c.inside = false;
  if (dot(n, c.eyev) < 0.0) {
    c.inside = true;
    n = -n;
  }
  c.normalv = n;
  c.over_point = c.point + n * EPS;
  c.under_point = c.point - n * EPS;
  c.reflectv = reflect(ray.d, n);
  c.n1 = -1.0;
  c.n2 = -1.0;
}

RTC_HD long long as_isize(double v) {  // Rust `as isize`: saturating, NaN → 0
  if (v != v) return 0;
  if (v >= 9223372036854775807.0) return 9223372036854775807LL;
  if (v <= -9223372036854775808.0) return (-9223372036854775807LL - 1LL);
  return (long long)v;
}

// Pattern::color_at_shape (pattern/mod.rs:39-49)
RTC_HD V3 pattern_color(const SceneView& S, int32_t pat, const double* leaf_inv, V3 world_point) {
  const rtc_pattern* P = S.patterns + pat;
  V3 op = xf_point(load_m34(leaf_inv), world_point);
  V3 p = xf_point(P->inverse, op);
  V3 a = v3(RTC_LDG(&P->a[0]), RTC_LDG(&P->a[1]), RTC_LDG(&P->a[2]));
  V3 b = v3(RTC_LDG(&P->b[0]), RTC_LDG(&P->b[1]), RTC_LDG(&P->b[2]));
  switch (RTC_LDG(&P->kind)) {
    case RTC_PAT_TEST: return p;                                                  // test_pattern.rs:7-9
    case RTC_PAT_STRIPE: return fmod(floor(p.x), 2.0) == 0.0 ? a : b;             // stripe.rs:14-20
    case RTC_PAT_GRADIENT: {                                                      // gradient.rs:14-18
      V3 distance = b - a;
      double fraction = p.x - floor(p.x);
      return a + distance * fraction;
    }
    case RTC_PAT_RING: {                                                          // ring.rs:14-21
      double distance = floor(sqrt(p.x * p.x + p.z * p.z));
      return fmod(distance, 2.0) == 0.0 ? a : b;
    }
    default: {                                                                    // checkers.rs:14-21
      double distance = floor(p.x) + floor(p.y) + floor(p.z);
      return as_isize(distance) % 2 == 0 ? a : b;
    }
  }
}

struct MatD {  // material scalars pulled into registers
  V3 color;
  double ambient, diffuse, specular, shininess, reflective, transparency;
  int32_t pattern;
};
RTC_HD MatD load_material(const SceneView& S, int32_t m) {
  const rtc_material* M = S.materials + m;
  MatD r;
  r.color = v3(RTC_LDG(&M->color[0]), RTC_LDG(&M->color[1]), RTC_LDG(&M->color[2]));
  r.ambient = RTC_LDG(&M->ambient);
  r.diffuse = RTC_LDG(&M->diffuse);
  r.specular = RTC_LDG(&M->specular);
  r.shininess = RTC_LDG(&M->shininess);
  r.reflective = RTC_LDG(&M->reflective);
  r.transparency = RTC_LDG(&M->transparency);
  r.pattern = RTC_LDG(&M->pattern);
  return r;
}

// Material::lighting (material.rs:38-82); `color` is the pattern/material colour at over_point.
// `lightv` = (light.position - point).normalize() (material.rs:52) is handed in: it is, bit for bit,
// the direction of the shadow ray is_shadowed built for the same point and light (world.rs:96-99).
RTC_HD V3 lighting(const MatD& m, V3 color, V3 lightv, V3 light_int, V3 eyev, V3 normalv, bool in_shadow) {
  V3 effective = color * light_int;
  V3 ambient = effective * m.ambient;
  if (in_shadow) return ambient;
  double ldn = dot(lightv, normalv);
  V3 diffuse = v3(0.0, 0.0, 0.0), specular = v3(0.0, 0.0, 0.0);
  if (!(ldn < 0.0)) {
    diffuse = effective * m.diffuse * ldn;
    V3 reflectv = reflect(-lightv, normalv);
    double rde = dot(reflectv, eyev);
    if (!(rde <= 0.0)) {
      // specular 0 (matte walls and floors): light_int * 0 * factor is the signed zero light_int * 0 for any
      // finite factor >= 0, and 0 < rde <= 2, 0 <= shininess <= 1000 bounds pow to (0, 2^1000]
#if !defined(RTK_NO_POWSKIP)
      if (m.specular == 0.0 && rde <= 2.0 && m.shininess >= 0.0 && m.shininess <= 1000.0) {
        specular = light_int * m.specular;
      } else
#endif
      {
        double factor = pow(rde, m.shininess);
        specular = light_int * m.specular * factor;
      }
    }
  }
  return ambient + diffuse + specular;
}

// Computations::schlick (intersection.rs:147-162)
RTC_HD double schlick(const Comps& c) {
  double cs = dot(c.eyev, c.normalv);
  if (c.n1 > c.n2) {
    double n = c.n1 / c.n2;
    double sin2_t = n * n * (1.0 - cs * cs);
    if (sin2_t > 1.0) return 1.0;
    cs = sqrt(1.0 - sin2_t);
  }
  double q = (c.n1 - c.n2) / (c.n1 + c.n2);
  double r0 = q * q;
  double x = 1.0 - cs;
  double x2 = x * x;
  double x4 = x2 * x2;
  return r0 + (1.0 - r0) * (x * x4);
}

// A suspended shade_hit: what a level of the color_at recursion needs once the colour of its
// reflection / refraction child comes back. Lives in LOCAL memory (dynamically indexed array) and is
// written once, when the level descends — 128 bytes per level, so the levels a thread really uses
// stay in L1. Everything else about the hit being shaded is transient and lives in Tracer::cur.
struct Frame {
  V3 surface;      // Σ lighting(...) over the lights (shade_hit's `surface`, world.rs:41-58)
  V3 reflected;    // reflected_color once known
  RayD refr_ray;   // the refraction ray, valid if need_refract
  double reflective, transparency, reflectance;
  int32_t state;   // 1: waiting for reflection child, 2: waiting for refraction child
  uint32_t bits;   // FR_*
};
enum : uint32_t { FR_NEED_REFRACT = 1u, FR_USE_SCHLICK = 2u };

// world.rs:62-67
RTC_HD V3 combine(V3 surface, V3 reflected, V3 refracted, double reflectance, bool use_schlick) {
  return use_schlick ? surface + reflected * reflectance + refracted * (1.0 - reflectance)
                     : surface + reflected + refracted;
}

// The hit being shaded at the current level, from the closest-hit answer until the level descends or
// returns. One copy serves every level: a parent's copy is dead once its child starts.
struct Shading {
  RayD ray;        // the ray color_at is being evaluated for at this level
  V3 point, over_point, normalv;  // eyev = -ray.d, under_point and reflectv are re-derived
  V3 color;        // pattern / material colour at over_point
  V3 surface;
  int32_t node, mat;
};

// World::color_at(ray, remaining = max_depth) (world.rs:70-81) with shade_hit, is_shadowed,
// reflected_color and refracted_color (world.rs:40-134) as ONE state machine around ONE search
// site: every closest-hit, shadow (world.rs:95-105) and containers query of the whole recursion
// tree is put to the same search, in the reference's evaluation order. The caller loops
//     tr.start(ray, depth);  for (;;) { search(S, tr.qray, tr.q, ...);  if (tr.advance(S, ...)) break; }
// so the (large) search code exists once per kernel and a warp stays converged on it whatever
// pixel, recursion depth or phase each lane is in.
enum : int { PH_CLOSEST = 0, PH_SHADOW = 1, PH_NEG = 2 };
enum : int { NX_RETURN = 0, NX_SHADOW = 1, NX_LIT = 2, NX_N1N2 = 3 };

struct Tracer {
  RayD qray;       // the ray of the pending query
  Query q;
  Shading cur;
  V3 result;
  double t_primary;
  int sp, remaining, li, phase;
  bool w_emit;  // wavefront mode (rt_wave.cuh): this call needs children and has handed them to `emit`

  RTC_HD void start(const RayD& r, int max_depth, Frame* frames, uint32_t* err) {
    (void)frames;
    w_emit = false;
    cur.ray = r;
    qray = r;
    sp = 0;
    remaining = max_depth;
    li = 0;
    phase = PH_CLOSEST;
    t_primary = -1.0;
    q.init_closest(err);
  }

  // Consume the answer of the pending query; set up the next one. Returns true when color_at of the
  // primary ray is complete (`result`).
  // WAVE (rt_wave.cuh): a call that needs children does not descend; it hands the suspended shade_hit
  // and the reflection ray to `emit` (the refraction ray is f.refr_ray) and reports completion.
  struct NoEmit { RTC_HD void operator()(const Frame&, const RayD&, bool) const {} };
  template <bool WAVE = false, class EmitFn = NoEmit>
  RTC_HD bool advance(const SceneView& S, Frame* frames, HitList& neg, RayCounters& rc, uint32_t* err, EmitFn emit = EmitFn()) {
    int next;
    V3 res = v3(0.0, 0.0, 0.0);
    double n1 = -1.0, n2 = -1.0;
    if (phase == PH_CLOSEST) {
      if (sp == 0) t_primary = q.node >= 0 ? q.limit : -1.0;
      if (q.node < 0) {
        res = v3(0.0, 0.0, 0.0);  // Color::black()
        next = NX_RETURN;
      } else {
        rtc_node nd = load_node(S.nodes + q.node);
        Comps c;
        prepare(S, cur.ray, q.limit, q.node, nd, c);
        cur.point = c.point;
        cur.over_point = c.over_point;
        cur.normalv = c.normalv;
        cur.node = q.node;
        cur.mat = nd.material;
        const rtc_material* M = S.materials + nd.material;
        const int32_t pat = RTC_LDG(&M->pattern);
        const double* leaf_inv = S.xforms + 12 * (size_t)nd.xform;
        cur.color = pat >= 0 ? pattern_color(S, pat, leaf_inv, c.over_point)
                             : v3(RTC_LDG(&M->color[0]), RTC_LDG(&M->color[1]), RTC_LDG(&M->color[2]));
        cur.surface = v3(0.0, 0.0, 0.0);  // shade_hit: Σ lighting in light order from (0,0,0), world.rs:41-58
        li = 0;
        next = S.n_lights > 0 ? NX_SHADOW : NX_LIT;
      }
    } else if (phase == PH_SHADOW) {
      const rtc_light* L = S.lights + li;
      V3 lint = v3(RTC_LDG(&L->intensity[0]), RTC_LDG(&L->intensity[1]), RTC_LDG(&L->intensity[2]));
      const MatD m = load_material(S, cur.mat);
      cur.surface = cur.surface + lighting(m, cur.color, qray.d, lint, -cur.ray.d, cur.normalv, q.done());
      li++;
      next = li < S.n_lights ? NX_SHADOW : NX_LIT;
    } else {
      containers_n1n2(S, neg, cur.node, n1, n2, err);
      next = NX_N1N2;
    }
    if (next == NX_SHADOW) {  // World::is_shadowed(over_point, light li), world.rs:95-105
      const rtc_light* L = S.lights + li;
      V3 lp = v3(RTC_LDG(&L->position[0]), RTC_LDG(&L->position[1]), RTC_LDG(&L->position[2]));
      const V3 op = cur.over_point;
      V3 v = lp - op;
      double distance = magnitude(v);
      qray.o = op;
      qray.d = normalize(v);
      rc.shadow++;
      q.init_shadow(distance, err);
      phase = PH_SHADOW;
#if !defined(RTK_NO_BACKFACE_SHADOW_SKIP)
      // The light is behind the shaded side (material.rs:60, light_dot_normal < 0): lighting() is `ambient`
      // in shadow and `ambient + 0 + 0` out of it, and `surface` — a sum that starts from +0 and therefore
      // can never be -0 — takes the same bits from either (x + -0 == x + +0 for every x but -0). The
      // occlusion query cannot change the frame: the ray is counted (the reference casts it) but the query
      // is born answered (Q_DONE), so the search returns at once. The lane still spends the shadow phase
      // with its warp: letting it run ahead desynchronises the warp's phases and costs more in this
      // function than the skipped search saves (2.61 -> 2.77 ms on the glass scene; skipping ahead only
      // to a further light helps two-light scenes by 8 % and costs the one-light headline as much, in
      // both of the forms tried — profiles/r01b_experiment_persistent_refill.md).
      if (dot(qray.d, cur.normalv) < 0.0) {
        q.mode = Q_DONE;
        q.node = 1;
      }
#endif
      return false;
    }
    bool descend = false;
    RayD child;
    child.o = child.d = v3(0.0, 0.0, 0.0);
    if (next == NX_LIT || next == NX_N1N2) {
      // the material's reflective / transparency decide what this level still needs (re-read after a
      // containers query rather than carried through it)
      const rtc_material* M = S.materials + cur.mat;
      const double reflective = RTC_LDG(&M->reflective), transparency = RTC_LDG(&M->transparency);
      const bool want_reflect = !(feq(reflective, 0.0) || remaining == 0);    // world.rs:108
      const bool want_refract = !(feq(transparency, 0.0) || remaining == 0);  // world.rs:117
      bool use_schlick = reflective > 0.0 && transparency > 0.0;              // world.rs:62
      // At remaining == 0 both children are black (world.rs:108,117), so world.rs:64 is
      // surface + 0·R + 0·(1−R): for any FINITE reflectance R >= 0 that is surface + (+0) + (±0), bit for bit what
      // the non-Schlick line surface + 0 + 0 gives — and R = schlick() is finite whenever every refractive
      // index in the scene is finite and positive. Then neither the containers query nor schlick() can
      // change a bit of the result, and the deepest level (the most populated one on glass) skips both.
#if !defined(RTK_NO_DEAD_SCHLICK)
      if (remaining == 0 && S.sane_indices) use_schlick = false;
#endif
      if (next == NX_LIT && (want_refract || use_schlick)) {
        // n1/n2 are only ever consumed by refracted_color and schlick
        qray = cur.ray;
        q.init_neg(&neg, err);
        phase = PH_NEG;
        return false;
      }
      const V3 rd = cur.ray.d, normalv = cur.normalv;
      const V3 eyev = -rd;
      double reflectance = 0.0;
      bool need_refract = false;
      RayD refr;
      refr.o = refr.d = v3(0.0, 0.0, 0.0);
      if (use_schlick) {
        Comps c;
        c.eyev = eyev; c.normalv = normalv; c.n1 = n1; c.n2 = n2;
        reflectance = schlick(c);
      }
      if (want_refract) {  // world.rs:121-131
        double n_ratio = n1 / n2;
        double cos_i = dot(eyev, normalv);
        double sin2_t = n_ratio * n_ratio * (1.0 - cos_i * cos_i);
        if (!(sin2_t > 1.0)) {
          double cos_t = sqrt(1.0 - sin2_t);
          refr.o = cur.point - normalv * EPS;  // under_point, intersection.rs:99
          refr.d = normalv * (n_ratio * cos_i - cos_t) - eyev * n_ratio;  // not normalised
          need_refract = true;
        }
      }
      if (want_reflect || need_refract) {
        Frame wave_f;
        Frame& f = WAVE ? wave_f : frames[sp];
        f.surface = cur.surface;
        f.reflected = v3(0.0, 0.0, 0.0);
        f.refr_ray = refr;
        f.reflective = reflective;
        f.transparency = transparency;
        f.reflectance = reflectance;
        f.bits = (need_refract ? FR_NEED_REFRACT : 0u) | (use_schlick ? FR_USE_SCHLICK : 0u);
        if (want_reflect) {
          rc.reflect++;
          f.state = 1;
          child.o = cur.over_point;
          child.d = reflect(rd, normalv);  // reflectv, intersection.rs:100
        } else {
          rc.refract++;
          f.state = 2;
          child = refr;
        }
        descend = true;
        if (WAVE) {  // both children are known now: emit them instead of descending
          if (want_reflect && need_refract) rc.refract++;
          emit(f, child, want_reflect);
          w_emit = true;
          return true;
        }
      } else {
        res = combine(cur.surface, v3(0.0, 0.0, 0.0), v3(0.0, 0.0, 0.0), reflectance, use_schlick);
      }
    }
    // ---- return `res` to the parents --------------------------------------------------------------
    while (!descend) {
      if (sp == 0) { result = res; return true; }
      sp--;
      remaining++;
      Frame& p = frames[sp];
      const uint32_t bits = p.bits;
      if (p.state == 1) {
        const V3 reflected = res * p.reflective;  // world.rs:113
        if (bits & FR_NEED_REFRACT) {
          p.reflected = reflected;
          rc.refract++;
          p.state = 2;
          child = p.refr_ray;
          descend = true;
        } else {
          res = combine(p.surface, reflected, v3(0.0, 0.0, 0.0), p.reflectance, (bits & FR_USE_SCHLICK) != 0);
        }
      } else {
        res = combine(p.surface, p.reflected, res * p.transparency, p.reflectance, (bits & FR_USE_SCHLICK) != 0);  // world.rs:133
      }
    }
    sp++;
    remaining--;
    cur.ray = child;
    qray = child;
    q.init_closest(err);
    phase = PH_CLOSEST;
    return false;
  }
};

template <bool HAS_CSG, int FAITHFUL>
RTC_HD V3 color_at(const SceneView& S, const RayD& ray, int max_depth, Scratch<HAS_CSG>& scratch, RayCounters& rc,
                   uint32_t* err, double* t_out = nullptr) {
  Tracer tr;
  Frame frames[MAX_FRAMES];  // kept out of Tracer: a dynamically indexed member would pin the whole struct in local memory
  tr.start(ray, max_depth, frames, err);
  for (;;) {
    search<HAS_CSG, FAITHFUL>(S, tr.qray, tr.q, scratch.csg(), &scratch.bvh, err);
    if (tr.advance(S, frames, scratch.neg, rc, err)) break;
  }
  if (t_out) *t_out = tr.t_primary;
  return tr.result;
}

struct CameraD {
  double pixel_size, half_width, half_height;
  double inv[12];
  V3 origin;  // inverse * Point::origin(), computed once on the host with the same expression
  uint32_t hsize, vsize;
};

// camera.rs:57-90
RTC_HD RayD ray_for_pixel(const CameraD& cam, uint32_t px, uint32_t py, double ox, double oy) {
  double xoffset = ((double)px + ox) * cam.pixel_size;
  double yoffset = ((double)py + oy) * cam.pixel_size;
  double world_x = cam.half_width - xoffset;
  double world_y = cam.half_height - yoffset;
  V3 w = v3(world_x, world_y, -1.0);
  V3 pixel = v3(cam.inv[0] * w.x + cam.inv[1] * w.y + cam.inv[2] * w.z + cam.inv[3],
                cam.inv[4] * w.x + cam.inv[5] * w.y + cam.inv[6] * w.z + cam.inv[7],
                cam.inv[8] * w.x + cam.inv[9] * w.y + cam.inv[10] * w.z + cam.inv[11]);
  RayD r;
  r.o = cam.origin;
  r.d = normalize(pixel - cam.origin);
  return r;
}

// camera.rs:92-126: sample s of an `aa`-sample pixel
RTC_HD void aa_offset(uint32_t aa, uint32_t s, double& ox, double& oy) {
  switch (aa) {
    case 2: ox = s == 0 ? 0.25 : 0.75; oy = 0.5; break;
    case 4: ox = (s & 1) ? 0.75 : 0.25; oy = (s & 2) ? 0.75 : 0.25; break;
    case 8: {
      const double xs[8] = {0.25, 0.5, 0.75, 0.25, 0.75, 0.25, 0.5, 0.75};
      const double ys[8] = {0.25, 0.25, 0.25, 0.5, 0.5, 0.75, 0.75, 0.75};
      ox = xs[s & 7]; oy = ys[s & 7];
      break;
    }
    case 16: ox = 0.125 + 0.25 * (double)(s & 3); oy = 0.125 + 0.25 * (double)((s >> 2) & 3); break;
    default: ox = 0.5; oy = 0.5; break;
  }
}

// One pixel of Camera::render (aa==1, camera.rs:133-148: colour stored as is) or of
// render_multithreaded (camera.rs:178-187: Color::average = sequential sum * (1.0/n)).
template <bool HAS_CSG, int FAITHFUL>
RTC_HD V3 render_pixel(const SceneView& S, const CameraD& cam, uint32_t px, uint32_t py, uint32_t aa, bool average,
                       int max_depth, Scratch<HAS_CSG>& scratch, RayCounters& rc, uint32_t* err) {
  // ONE color_at site for both (two would double the kernel's code and thrash the instruction cache)
  V3 acc = v3(0.0, 0.0, 0.0);
  const uint32_t n = average ? aa : 1u;
  for (uint32_t s = 0; s < n; s++) {
    double ox = 0.5, oy = 0.5;
    if (average) aa_offset(aa, s, ox, oy);
    rc.primary++;
    V3 c = color_at<HAS_CSG, FAITHFUL>(S, ray_for_pixel(cam, px, py, ox, oy), max_depth, scratch, rc, err);
    acc = average ? acc + c : c;
  }
  return average ? acc * (1.0 / (double)aa) : acc;
}

}  // namespace rtk
