// rt_isect.cuh — device scene view, AABB slab test and the seven leaf `local_intersect`s.
//
// Each leaf routine receives the OBJECT-space ray and calls emit(t) once per intersection in the
// exact order the reference pushes them into its Vec<Intersection>.
// Reference (relative to raytracer/src/geometry/shape): sphere.rs:47-62, plane.rs:53-60,
// cube.rs:30-48,71-93, cylinder.rs:42-65,88-119, cone.rs:48-72,94-134, triangle.rs:67-90,
// smooth_triangle.rs:73-96; ../../bounding_box.rs:95-136.
#pragma once
#include "../../include/rtc.h"
#include "rt_math.cuh"
#include "rt_accel_build.hpp"

namespace rtk {

// Device-resident copy of rtc_scene_desc (same array meanings), plus derived tables.
struct SceneView {
  const rtc_node* nodes;
  const double* xforms;      // 12 per transform
  const double* boxes;       // 6 per box
  const double* limits;      // 2 per cylinder/cone
  const double* tri_geom;    // 9 per triangle: p1, e1, e2
  const double* tri_normals; // 9 per triangle
  const rtc_material* materials;
  const rtc_pattern* patterns;
  const rtc_light* lights;
  const int32_t* csg_parent; // per node: nearest enclosing CSG node, or -1
  // traversal accelerator (rt_accel_build.hpp)
  const int32_t* parent;     // per node: innermost enclosing reference Group (outside CSGs), or -1
  const int32_t* prims;      // node indices: [0, n_always) tested for every ray, then BVH leaf order
  const TravNode* bvh;       // the BVH the kernels traverse (four children per node), node 0 = root
  const PrimRec* recs;       // one record per always-prim
  const LeafRec* leaf_recs;  // one record per BVH prim (prims[n_always + k])
  int32_t n_nodes;
  int32_t n_lights;
  int32_t has_csg;
  int32_t n_always;
  int32_t n_bvh;             // BVH node count (0: no BVH)
  int32_t n_sph, n_pl;       // always-prefix: [0,n_sph) spheres, [n_sph,n_sph+n_pl) planes, then the rest
  int32_t sane_indices;      // 1: every material's refractive_index is finite and > 0 (schlick() is then finite)
  SphCull sph_cull[SPH_CULL_MAX];  // bounding spheres of the first every-ray spheres (parameter bank)
  CubeDir cube_dirs[CUBE_DIR_MAX]; // face families of the cubes in the BVH (EPSILON-parallel rule of cube.rs:30-48)
  int32_t n_cube_dirs;
  int32_t _pad_cube;
  double bvh_ext[3];         // the BVH's largest |coordinate| per axis (error bound of the single-precision slab test)
};

RTC_HD void slab_axis(double origin, double direction, double mn, double mx, double& tmin, double& tmax) {
  // bounding_box.rs:118-136 and cube.rs:30-48 (mn=-1, mx=1)
  double tmin_numerator = mn - origin;
  double tmax_numerator = mx - origin;
  double a, b;
  if (fabs(direction) >= EPS) {
    divide2(tmin_numerator, tmax_numerator, direction, a, b);
  } else {
    a = tmin_numerator * rt_inf();
    b = tmax_numerator * rt_inf();
  }
  if (a > b) { tmin = b; tmax = a; } else { tmin = a; tmax = b; }
}

// bounding_box.rs:95-116. Returns tmin <= tmax and the interval itself (for the optional
// conservative t-range culling, which the reference does not have).
RTC_HD bool box_intersects(const double* bx, const RayD& r, double& tmin, double& tmax) {
  double xtmin, xtmax, ytmin, ytmax, ztmin, ztmax;
  slab_axis(r.o.x, r.d.x, RTC_LDG(bx + 0), RTC_LDG(bx + 3), xtmin, xtmax);
  slab_axis(r.o.y, r.d.y, RTC_LDG(bx + 1), RTC_LDG(bx + 4), ytmin, ytmax);
  slab_axis(r.o.z, r.d.z, RTC_LDG(bx + 2), RTC_LDG(bx + 5), ztmin, ztmax);
  tmin = fmax(fmax(xtmin, ytmin), ztmin);  // f64::max == fmax (NaN-ignoring)
  tmax = fmin(fmin(xtmax, ytmax), ztmax);
  return tmin <= tmax;
}

template <class Emit>
RTC_HD void isect_sphere(const RayD& r, Emit&& emit) {
  double a = dot(r.d, r.d);
  double b = 2.0 * dot(r.d, r.o);
  double c = dot(r.o, r.o) - 1.0;
  double disc = b * b - 4.0 * a * c;
  if (disc < 0.0) return;
  double sq = sqrt(disc);
  double t1, t2;
  divide2(-b - sq, -b + sq, 2.0 * a, t1, t2);
  emit(t1);
  emit(t2);
}

template <class Emit>
RTC_HD void isect_plane(const RayD& r, Emit&& emit) {
  if (fabs(r.d.y) < EPS) return;
  emit(-r.o.y / r.d.y);
}

template <class Emit>
RTC_HD void isect_cube(const RayD& r, Emit&& emit) {
  double xtmin, xtmax, ytmin, ytmax, ztmin, ztmax;
  slab_axis(r.o.x, r.d.x, -1.0, 1.0, xtmin, xtmax);
  slab_axis(r.o.y, r.d.y, -1.0, 1.0, ytmin, ytmax);
  slab_axis(r.o.z, r.d.z, -1.0, 1.0, ztmin, ztmax);
  double tmin = fmax(fmax(xtmin, ytmin), ztmin);
  double tmax = fmin(fmin(xtmax, ytmax), ztmax);
  if (tmin > tmax) return;
  emit(tmin);
  emit(tmax);
}

// cylinder.rs:48-65 / cone.rs:48-65 — note: no |d.y| < EPSILON guard; ±inf/NaN t fail check_cap.
template <class Emit>
RTC_HD void isect_caps(const RayD& r, double minimum, double maximum, bool cone, Emit&& emit) {
  double t, tmax_cap;
  divide2(minimum - r.o.y, maximum - r.o.y, r.d.y, t, tmax_cap);
  double x = r.o.x + t * r.d.x;
  double z = r.o.z + t * r.d.z;
  double rad2 = cone ? minimum * minimum : 1.0;
  if ((x * x + z * z) <= rad2) emit(t);
  t = tmax_cap;
  x = r.o.x + t * r.d.x;
  z = r.o.z + t * r.d.z;
  rad2 = cone ? maximum * maximum : 1.0;
  if ((x * x + z * z) <= rad2) emit(t);
}

template <class Emit>
RTC_HD void isect_cylinder(const RayD& r, double minimum, double maximum, bool closed, Emit&& emit) {
  double a = r.d.x * r.d.x + r.d.z * r.d.z;
  if (fabs(a) < EPS) {
    if (closed) isect_caps(r, minimum, maximum, false, emit);
    return;
  }
  double b = 2.0 * r.o.x * r.d.x + 2.0 * r.o.z * r.d.z;
  double c = r.o.x * r.o.x + r.o.z * r.o.z - 1.0;
  double disc = b * b - 4.0 * a * c;
  if (disc < 0.0) return;  // before the caps, as in the reference
  double sq = sqrt(disc);
  double t0, t1;
  divide2(-b - sq, -b + sq, 2.0 * a, t0, t1);
  double y0 = r.o.y + t0 * r.d.y;
  if (minimum < y0 && y0 < maximum) emit(t0);
  double y1 = r.o.y + t1 * r.d.y;
  if (minimum < y1 && y1 < maximum) emit(t1);
  if (closed) isect_caps(r, minimum, maximum, false, emit);
}

template <class Emit>
RTC_HD void isect_cone(const RayD& r, double minimum, double maximum, bool closed, Emit&& emit) {
  double a = r.d.x * r.d.x - r.d.y * r.d.y + r.d.z * r.d.z;
  double b = 2.0 * r.o.x * r.d.x - 2.0 * r.o.y * r.d.y + 2.0 * r.o.z * r.d.z;
  double c = r.o.x * r.o.x - r.o.y * r.o.y + r.o.z * r.o.z;
  if (fabs(a) < EPS) {
    if (fabs(b) < EPS) {
      if (closed) isect_caps(r, minimum, maximum, true, emit);
      return;
    }
    emit(-c / 2.0 * b);  // cone.rs:106: (−c/2)·b — reference behaviour, kept
    if (closed) isect_caps(r, minimum, maximum, true, emit);
    return;
  }
  double disc = b * b - 4.0 * a * c;
  if (disc < 0.0) return;
  double sq = sqrt(disc);
  double t0, t1;
  divide2(-b - sq, -b + sq, 2.0 * a, t0, t1);
  double y0 = r.o.y + t0 * r.d.y;
  if (minimum < y0 && y0 < maximum) emit(t0);
  double y1 = r.o.y + t1 * r.d.y;
  if (minimum < y1 && y1 < maximum) emit(t1);
  if (closed) isect_caps(r, minimum, maximum, true, emit);
}

// Möller–Trumbore in object space; g = p1(3), e1(3), e2(3).
RTC_HD bool isect_triangle(const RayD& r, const double* g, double& t, double& u, double& v) {
  V3 p1 = v3(RTC_LDG(g + 0), RTC_LDG(g + 1), RTC_LDG(g + 2));
  V3 e1 = v3(RTC_LDG(g + 3), RTC_LDG(g + 4), RTC_LDG(g + 5));
  V3 e2 = v3(RTC_LDG(g + 6), RTC_LDG(g + 7), RTC_LDG(g + 8));
  V3 dir_cross_e2 = cross(r.d, e2);
  double det = dot(e1, dir_cross_e2);
  if (fabs(det) < EPS) return false;
  double f = 1.0 / det;
  V3 p1_to_origin = r.o - p1;
  u = f * dot(p1_to_origin, dir_cross_e2);
  if (!(0.0 <= u && u <= 1.0)) return false;
  V3 origin_cross_e1 = cross(p1_to_origin, e1);
  v = f * dot(r.d, origin_cross_e1);
  if (v < 0.0 || (u + v) > 1.0) return false;
  t = f * dot(e2, origin_cross_e1);
  return true;
}

// Up to 4 intersections of one leaf, in the order the reference pushes them (cylinder/cone: two
// walls + two caps is the maximum).
struct LeafHits {
  double t0, t1, t2, t3;
  int n;
  RTC_HD void push(double t) {
    if (n == 0) t0 = t; else if (n == 1) t1 = t; else if (n == 2) t2 = t; else t3 = t;
    n++;
  }
};

// Dispatch on the leaf kind. `nd` is the node (already loaded), `lr` the object-space ray.
RTC_HD void isect_leaf(const SceneView& S, const rtc_node& nd, const RayD& lr, LeafHits& h) {
  h.n = 0;
  h.t0 = h.t1 = h.t2 = h.t3 = 0.0;
  auto emit = [&](double t) { h.push(t); };
  switch (nd.kind) {
    case RTC_SPHERE: isect_sphere(lr, emit); break;
    case RTC_PLANE: isect_plane(lr, emit); break;
    case RTC_CUBE: isect_cube(lr, emit); break;
    case RTC_CYLINDER: {
      const double* lim = S.limits + 2 * nd.aux;
      isect_cylinder(lr, RTC_LDG(lim), RTC_LDG(lim + 1), (nd.flags & RTC_FLAG_CLOSED) != 0, emit);
      break;
    }
    case RTC_CONE: {
      const double* lim = S.limits + 2 * nd.aux;
      isect_cone(lr, RTC_LDG(lim), RTC_LDG(lim + 1), (nd.flags & RTC_FLAG_CLOSED) != 0, emit);
      break;
    }
    case RTC_TRIANGLE:
    case RTC_SMOOTH_TRIANGLE: {
      double t, u, v;
      if (isect_triangle(lr, S.tri_geom + 9 * (size_t)nd.aux, t, u, v)) h.push(t);
      break;
    }
    default: break;
  }
}

}  // namespace rtk
