// TEST-ONLY host build of the kernel SOURCE (raytracer-challenge-rs_b200/csrc/rt_*.cuh compiled by
// g++ with -ffp-contract=off), so that the CPU-only test tier (`pytest -m "not gpu"`, no GPU in the
// build container) can check the device traversal/shading logic and the flattener against the
// oracle before GPU time is spent. It is NOT a fallback: it is never linked into librtc_cuda.so or
// librth_host.so, never imported by the package, and lives under tests/. The product path fails
// loudly without a CUDA device.
#include <cstring>
#include <thread>
#include <vector>

#include "../../raytracer-challenge-rs_b200/csrc/rt_shade.cuh"

using namespace rtk;

namespace {
struct Built {
  SceneView view;
  std::vector<int32_t> csg_parent;
  AccelBuild accel;
};
void build_view(const rtc_scene_desc* d, Built& b) {
  const int32_t n = (int32_t)d->n_nodes;
  b.csg_parent.assign((size_t)n, -1);
  bool has_csg = false;
  struct Open { int32_t end; bool csg; int32_t node; };
  std::vector<Open> open;
  for (int32_t i = 0; i < n; i++) {
    while (!open.empty() && open.back().end == i) open.pop_back();
    for (auto it = open.rbegin(); it != open.rend(); ++it)
      if (it->csg) { b.csg_parent[(size_t)i] = it->node; break; }
    const rtc_node& nd = d->nodes[i];
    if (nd.kind == RTC_CSG) has_csg = true;
    if ((nd.kind == RTC_GROUP || nd.kind == RTC_CSG) && nd.skip > i + 1) open.push_back({nd.skip, nd.kind == RTC_CSG, i});
  }
  SceneView& v = b.view;
  v.nodes = d->nodes; v.xforms = d->xforms; v.boxes = d->boxes; v.limits = d->limits;
  v.tri_geom = d->tri_p1e1e2; v.tri_normals = d->tri_normals; v.materials = d->materials;
  v.patterns = d->patterns; v.lights = d->lights; v.csg_parent = b.csg_parent.data();
  v.n_nodes = n; v.n_lights = (int32_t)d->n_lights; v.has_csg = has_csg ? 1 : 0; v.sane_indices = 1;
  for (uint32_t m = 0; m < d->n_materials; m++)
    if (!(d->materials[m].refractive_index > 0.0) || !std::isfinite(d->materials[m].refractive_index)) v.sane_indices = 0;
  build_accel(d, b.accel);
  v.parent = b.accel.parent.data(); v.prims = b.accel.prims.data(); v.recs = b.accel.recs.data(); v.leaf_recs = b.accel.leaf_recs.data();
  v.bvh = b.accel.trav_data(); v.n_always = b.accel.n_always; v.n_sph = b.accel.n_sph; v.n_pl = b.accel.n_pl; v.n_bvh = (int32_t)b.accel.trav_count();
  for (int k = 0; k < SPH_CULL_MAX; k++) v.sph_cull[k] = b.accel.sph_cull[k];
  for (int k = 0; k < CUBE_DIR_MAX; k++) v.cube_dirs[k] = b.accel.cube_dirs[k];
  v.n_cube_dirs = b.accel.n_cube_dirs;
  v.bvh_ext[0] = v.bvh_ext[1] = v.bvh_ext[2] = 0.0;
  if (!b.accel.bvh.empty()) bvh_extent(b.accel.bvh[0], v.bvh_ext);
}
CameraD make_cam(const rtc_camera* c) {
  CameraD cam;
  cam.pixel_size = c->pixel_size; cam.half_width = c->half_width; cam.half_height = c->half_height;
  for (int i = 0; i < 12; i++) cam.inv[i] = c->inverse[i];
  const double z = 0.0;
  cam.origin = v3(c->inverse[0] * z + c->inverse[1] * z + c->inverse[2] * z + c->inverse[3],
                  c->inverse[4] * z + c->inverse[5] * z + c->inverse[6] * z + c->inverse[7],
                  c->inverse[8] * z + c->inverse[9] * z + c->inverse[10] * z + c->inverse[11]);
  cam.hsize = c->hsize; cam.vsize = c->vsize;
  return cam;
}
template <bool CSG, bool FA>
void rows(const SceneView& S, const CameraD& cam, uint32_t y0, uint32_t y1, uint32_t aa, bool avg, int depth, double* out,
          uint64_t* counters) {
  Scratch<CSG>* scratch = new Scratch<CSG>();
  RayCounters rc = {0, 0, 0, 0};
  uint64_t ovf = 0, nan = 0;
  for (uint32_t y = y0; y < y1; y++)
    for (uint32_t x = 0; x < cam.hsize; x++) {
      uint32_t err = 0;
      V3 c = render_pixel<CSG, FA>(S, cam, x, y, aa, avg, depth, *scratch, rc, &err);
      double* o = out + ((size_t)y * cam.hsize + x) * 3;
      o[0] = c.x; o[1] = c.y; o[2] = c.z;
      if (err & ERR_OVERFLOW) ovf++;
      if (err & ERR_NAN) nan++;
      counters[0] += rc.primary; counters[1] += rc.shadow; counters[2] += rc.reflect; counters[3] += rc.refract;
      rc = {0, 0, 0, 0};
    }
  counters[4] += ovf; counters[5] += nan;
  delete scratch;
}
}  // namespace

#define EMUL_API __attribute__((visibility("default")))
extern "C" {

// Same meaning as rtc_render, on host threads. counters[6] = rays primary/shadow/reflect/refract,
// overflow pixels, NaN pixels.
EMUL_API int emul_render(const rtc_scene_desc* d, const rtc_camera* c, const rtc_opts* o, unsigned threads, double* rgb,
                uint64_t* counters) {
  Built b;
  build_view(d, b);
  CameraD cam = make_cam(c);
  uint32_t y0 = o->row_begin, y1 = o->row_end;
  if (y0 == 0 && y1 == 0) y1 = c->vsize;
  if (y1 > c->vsize) y1 = c->vsize;
  bool avg = o->aa_samples > 1 || (o->flags & RTC_OPT_AVERAGE);
  bool fa = (o->flags & RTC_OPT_FAITHFUL_TRAVERSAL) != 0;
  if (threads == 0) threads = 1;
  std::vector<std::thread> th;
  std::vector<std::vector<uint64_t>> per(threads, std::vector<uint64_t>(6, 0));
  for (unsigned t = 0; t < threads; t++) {
    th.emplace_back([&, t]() {
      // interleaved rows for load balance
      const uint32_t nranks = o->nranks ? o->nranks : 1, tile_rows = o->tile_rows ? o->tile_rows : 8;
      for (uint32_t y = y0 + t; y < y1; y += threads) {
        if ((y / tile_rows) % nranks != o->rank) continue;  // rtc_opts sharding: tile t belongs to rank t % nranks
        if (b.view.has_csg) { if (fa) rows<true, true>(b.view, cam, y, y + 1, o->aa_samples, avg, (int)o->max_depth, rgb, per[t].data());
                              else rows<true, false>(b.view, cam, y, y + 1, o->aa_samples, avg, (int)o->max_depth, rgb, per[t].data()); }
        else { if (fa) rows<false, true>(b.view, cam, y, y + 1, o->aa_samples, avg, (int)o->max_depth, rgb, per[t].data());
               else rows<false, false>(b.view, cam, y, y + 1, o->aa_samples, avg, (int)o->max_depth, rgb, per[t].data()); }
      }
    });
  }
  for (auto& t : th) t.join();
  for (int k = 0; k < 6; k++) { counters[k] = 0; for (auto& p : per) counters[k] += p[(size_t)k]; }
  return 0;
}

EMUL_API int emul_trace_rays(const rtc_scene_desc* d, size_t n, const double* origins, const double* dirs, uint32_t remaining,
                    double* rgb, double* t_out) {
  Built b;
  build_view(d, b);
  uint32_t err = 0;
  RayCounters rc = {0, 0, 0, 0};
  for (size_t i = 0; i < n; i++) {
    RayD r;
    r.o = v3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]);
    r.d = v3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
    double t = -1.0;
    V3 c;
    if (b.view.has_csg) { Scratch<true> s; c = color_at<true, true>(b.view, r, (int)remaining, s, rc, &err, &t); }
    else { Scratch<false> s; c = color_at<false, true>(b.view, r, (int)remaining, s, rc, &err, &t); }
    if (rgb) { rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z; }
    if (t_out) t_out[i] = t;
  }
  return (int)err;
}

EMUL_API int emul_probe_hit(const rtc_scene_desc* d, const double* o, const double* dir, double* out) {
  Built b;
  build_view(d, b);
  const SceneView& S = b.view;
  RayD ray;
  ray.o = v3(o[0], o[1], o[2]);
  ray.d = v3(dir[0], dir[1], dir[2]);
  for (int i = 0; i < 20; i++) out[i] = 0.0;
  uint32_t err = 0;
  Scratch<true> scratch;
  Query ch;
  ch.init_closest(&err);
  search<true, true>(S, ray, ch, scratch.csg(), &scratch.bvh, &err);
  if (ch.node < 0) return (int)err;
  rtc_node nd = load_node(S.nodes + ch.node);
  Comps c;
  prepare(S, ray, ch.limit, ch.node, nd, c);
  Query nc;
  nc.init_neg(&scratch.neg, &err);
  search<true, true>(S, ray, nc, scratch.csg(), &scratch.bvh, &err);
  containers_n1n2(S, scratch.neg, ch.node, c.n1, c.n2, &err);
  out[0] = 1.0; out[1] = c.t;
  out[2] = c.point.x; out[3] = c.point.y; out[4] = c.point.z;
  out[5] = c.normalv.x; out[6] = c.normalv.y; out[7] = c.normalv.z;
  out[8] = c.over_point.x; out[9] = c.over_point.y; out[10] = c.over_point.z;
  out[11] = c.under_point.x; out[12] = c.under_point.y; out[13] = c.under_point.z;
  out[14] = c.reflectv.x; out[15] = c.reflectv.y; out[16] = c.reflectv.z;
  out[17] = c.n1; out[18] = c.n2; out[19] = c.inside ? 1.0 : 0.0;
  return (int)err;
}

EMUL_API int emul_is_shadowed(const rtc_scene_desc* d, const double* p, uint32_t light) {
  Built b;
  build_view(d, b);
  uint32_t err = 0;
  Scratch<true> scratch;
  const rtc_light& L = d->lights[light];
  V3 lp = v3(L.position[0], L.position[1], L.position[2]), pt = v3(p[0], p[1], p[2]);
  V3 v = lp - pt;
  RayD r;
  r.o = pt;
  r.d = normalize(v);
  Query sh;
  sh.init_shadow(magnitude(v), &err);
  search<true, true>(b.view, r, sh, scratch.csg(), &scratch.bvh, &err);
  return sh.done() ? 1 : 0;
}

}  // extern "C"

#if defined(RTK_STATS)
// Diagnostic build only (-DRTK_STATS): per-pixel search cost (BVH nodes, prim tests, reference-box validations) of
// the accelerated renderer, single-threaded. cost[pixel*4 + {0,1,2,3}] = walks, bvh_nodes, prim_tests, chain_boxes.
extern "C" EMUL_API int emul_cost_map(const rtc_scene_desc* d, const rtc_camera* c, const rtc_opts* o, uint64_t* cost) {
  Built b;
  build_view(d, b);
  CameraD cam = make_cam(c);
  if (b.view.has_csg) return -1;
  Scratch<false>* scratch = new Scratch<false>();
  for (uint32_t y = 0; y < c->vsize; y++)
    for (uint32_t x = 0; x < c->hsize; x++) {
      accel_stats() = AccelStats();
      RayCounters rc = {0, 0, 0, 0};
      uint32_t err = 0;
      render_pixel<false, false>(b.view, cam, x, y, 1, false, (int)o->max_depth, *scratch, rc, &err);
      uint64_t* q = cost + ((size_t)y * c->hsize + x) * 4;
      q[0] = accel_stats().walks; q[1] = accel_stats().bvh_nodes; q[2] = accel_stats().prim_tests; q[3] = accel_stats().chain_boxes;
    }
  delete scratch;
  return 0;
}
#endif
