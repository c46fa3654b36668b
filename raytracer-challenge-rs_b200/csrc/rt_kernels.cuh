// rt_kernels.cuh — the __global__ entry points (sm_100a). One thread = one pixel (all of its AA
// samples, summed in the reference's order); a warp covers an 8x4 pixel patch and a 128-thread
// CTA a 16x8 tile so that the threads of a warp walk the same nodes and hit the same leaves.
// FP64 CUDA-core arithmetic only (no dense contraction on this path → no tensor cores).
#pragma once
#include <cuda_runtime.h>

#include "rt_shade.cuh"

#ifndef RTK_MIN_BLOCKS
#define RTK_MIN_BLOCKS 4
#endif
// resident CTAs per SM of the kernels with the BVH traversal and the generic leaf tests in them (no CSG): three, i.e.
// 168 registers — the search's state spills at 128 (profiles/r02bd_min_blocks.log: teapot 1.50 -> 1.34 ms, cover 2.87 ->
// 2.55 ms with three, 1.57 / 2.77 with two, 1.71 / 3.05 with five; the CSG kernels do not care, the lean kernel wants four)
#ifndef RTK_BVH_MINB
#define RTK_BVH_MINB 3
#endif
// resident CTAs per SM of the kernels launched for accelerators far beyond L2 (rtc_scene::big_accel)
#ifndef RTK_BIG_MINB
#define RTK_BIG_MINB 8
#endif

namespace rtk {

struct RenderParams {
  uint32_t row_begin, row_end;  // frame rows rendered by this call
  uint32_t tile_rows, rank, nranks;
  uint32_t owned_rows;          // rows of the launch domain (owned tiles * tile_rows)
  uint32_t oy0;                 // first owned row of THIS launch (row bands: render band k+1 while band k is copied out)
  uint32_t aa, average;
  int32_t max_depth;
  uint32_t faithful;             // RTC_OPT_FAITHFUL_TRAVERSAL: the reference's own search order
  uint32_t flags;                // rtc_opts.flags
  const uint32_t* order;         // CTA dispatch order: order[linear block id] = tile_y << 16 | tile_x (NULL: row-major)
};

// counters[0..3] rays primary/shadow/reflect/refract, [4] overflow pixels, [5] NaN pixels
__device__ __forceinline__ void flush_counters(const RayCounters& rc, uint32_t err, unsigned long long* counters) {
  unsigned a = rc.primary, b = rc.shadow, c = rc.reflect, d = rc.refract;
  unsigned e = (err & ERR_OVERFLOW) ? 1u : 0u, f = (err & ERR_NAN) ? 1u : 0u;
  unsigned mask = __activemask();
  a = __reduce_add_sync(mask, a);
  b = __reduce_add_sync(mask, b);
  c = __reduce_add_sync(mask, c);
  d = __reduce_add_sync(mask, d);
  e = __reduce_add_sync(mask, e);
  f = __reduce_add_sync(mask, f);
  int leader = __ffs(mask) - 1;
  if ((int)(threadIdx.x & 31) == leader) {
    if (a) atomicAdd(counters + 0, (unsigned long long)a);
    if (b) atomicAdd(counters + 1, (unsigned long long)b);
    if (c) atomicAdd(counters + 2, (unsigned long long)c);
    if (d) atomicAdd(counters + 3, (unsigned long long)d);
    if (e) atomicAdd(counters + 4, (unsigned long long)e);
    if (f) atomicAdd(counters + 5, (unsigned long long)f);
  }
}

// A warp's 8x4 patch is four row segments of 8 pixels = 24 doubles = 192 contiguous bytes. Storing c.x, c.y,
// c.z per lane is three instructions of 8-byte pieces at a 24-byte stride: every 32-byte sector is written three
// times in part, which a device canvas absorbs in L2 but a mapped HOST canvas (PCIe) or a PEER canvas (NVLink) pays
// for per piece — at 2.05 ms per frame the link's small-write rate stretched the kernel to 3.3 ms. Here the 8 lanes
// of a row transpose through shuffles so that one instruction writes 128 contiguous bytes (16 B per lane) and a
// second the remaining 64. Needs the whole warp (full patch) and a 16-byte aligned segment; otherwise plain stores.
// Called by all 32 lanes of the warp, converged (`alive`: this lane has a pixel).
__device__ __forceinline__ void store_patch_row(double* o, V3 c, uint32_t lane, bool alive) {
#if defined(RTK_PLAIN_STORES)
  if (alive) { o[0] = c.x; o[1] = c.y; o[2] = c.z; }
  return;
#endif
  const uint32_t j = lane & 7u;
  const bool ok = alive && ((reinterpret_cast<uintptr_t>(o - 3 * j)) & 15u) == 0;
  if (!__all_sync(0xffffffffu, ok)) {
    if (alive) { o[0] = c.x; o[1] = c.y; o[2] = c.z; }
    return;
  }
  // value k (0..23) of the segment lives in lane k/3, component k%3
  auto value = [&](uint32_t k) {
    const int src = (int)(k / 3u);
    const uint32_t m = k - 3u * (uint32_t)src;
    const double x = __shfl_sync(0xffffffffu, c.x, src, 8), y = __shfl_sync(0xffffffffu, c.y, src, 8),
                 z = __shfl_sync(0xffffffffu, c.z, src, 8);
    return m == 0u ? x : (m == 1u ? y : z);
  };
  double* seg = o - 3 * j;
  const double a0 = value(2u * j), a1 = value(2u * j + 1u);
  const uint32_t kb = 16u + 2u * (j & 3u);
  const double b0 = value(kb), b1 = value(kb + 1u);
  reinterpret_cast<double2*>(seg)[j] = make_double2(a0, a1);
  if (j < 4u) reinterpret_cast<double2*>(seg)[8u + j] = make_double2(b0, b1);
}

// Camera::render / render_multithreaded: one thread per pixel (all its AA samples, summed in the
// reference's order). 16x8 tile per CTA; warp w → 8x4 patch (w&1, w>>1); lane → (lane&7, lane>>3),
// so the lanes of a warp put near-identical queries to the scene and stay in the same phase of the
// Tracer state machine. (A persistent variant with lane-level re-fill was measured and rejected:
// profiles/r01b_experiment_persistent_refill.md.)
// MINB = CTAs per SM the register budget is cut for: 4 (128 registers; the FP64 state of search and
// shading just fits) for scenes whose tables live in L1/L2; 8 (64 registers, 32 warps per SM) for BVHs
// far larger than L2, where traversal waits on DRAM and only more warps hide it (1 M triangles at 4K:
// 91 → 69 ms; an 8 k-triangle mesh is 25 % slower that way, so the host picks by accelerator size).
template <bool HAS_CSG, int FAITHFUL, int MINB = RTK_MIN_BLOCKS>
__global__ void __launch_bounds__(128, MINB) render_kernel(SceneView S, CameraD cam, RenderParams P, double* __restrict__ out,
                                                                     unsigned long long* __restrict__ counters) {
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  // CTAs are dispatched in linear block order; `order` maps that order to tiles, centre of the image first: the
  // heaviest warps of a frame (glass, mesh silhouettes: 1 ms chains) usually sit near the centre, and a chain that
  // starts at t = 0 hides behind the rest of the frame instead of sticking out of its end
  uint32_t bx = blockIdx.x, by = blockIdx.y;
  if (P.order) {
    const uint32_t t = __ldg(P.order + blockIdx.y * gridDim.x + blockIdx.x);
    bx = t & 0xffffu;
    by = t >> 16;
  }
  const uint32_t px = bx * 16u + (warp & 1u) * 8u + (lane & 7u);
  const uint32_t oy = P.oy0 + by * 8u + (warp >> 1) * 4u + (lane >> 3);
  bool alive = px < cam.hsize && oy < P.owned_rows;
  // owned row index → frame row: tile (rank + k*nranks), row within tile
  const uint32_t k = oy / P.tile_rows, r = oy - k * P.tile_rows;
  const uint32_t py = (P.rank + k * P.nranks) * P.tile_rows + r;
  alive = alive && py >= P.row_begin && py < P.row_end;

  Scratch<HAS_CSG> scratch;
  RayCounters rc = {0u, 0u, 0u, 0u};
  uint32_t err = 0;
  V3 c = v3(0.0, 0.0, 0.0);
  if (alive) c = render_pixel<HAS_CSG, FAITHFUL>(S, cam, px, py, P.aa, P.average != 0, P.max_depth, scratch, rc, &err);
  __syncwarp();
  store_patch_row(out + ((size_t)py * cam.hsize + px) * 3, c, lane, alive);  // Canvas::pixel_idx, canvas.rs:44-48
  flush_counters(rc, err, counters);
}

// World::color_at for a batch of rays (debug / known-answer entry)
template <bool HAS_CSG>
__global__ void __launch_bounds__(128) trace_kernel(SceneView S, size_t n, const double* __restrict__ origins,
                                                    const double* __restrict__ dirs, int remaining,
                                                    double* __restrict__ rgb, double* __restrict__ t_out,
                                                    unsigned long long* __restrict__ counters) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Scratch<HAS_CSG> scratch;
  RayCounters rc = {0u, 0u, 0u, 0u};
  uint32_t err = 0;
  RayD r;
  r.o = v3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]);
  r.d = v3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
  double t = -1.0;
  rc.primary++;
  V3 c = color_at<HAS_CSG, true>(S, r, remaining, scratch, rc, &err, &t);
  if (rgb) { rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z; }
  if (t_out) t_out[i] = t;
  flush_counters(rc, err, counters);
}

// prepare_computations for one ray (debug / known-answer entry); out[20] as documented in rtc.h
template <bool HAS_CSG>
__global__ void probe_kernel(SceneView S, RayD ray, double* __restrict__ out, unsigned long long* __restrict__ counters) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  Scratch<HAS_CSG> scratch;
  uint32_t err = 0;
  for (int i = 0; i < 20; i++) out[i] = 0.0;
  Query ch;
  ch.init_closest(&err);
  search<HAS_CSG, true>(S, ray, ch, scratch.csg(), &scratch.bvh, &err);
  if (ch.node >= 0) {
    rtc_node nd = load_node(S.nodes + ch.node);
    Comps c;
    prepare(S, ray, ch.limit, ch.node, nd, c);
    Query nc;
    nc.init_neg(&scratch.neg, &err);
    search<HAS_CSG, true>(S, ray, nc, scratch.csg(), &scratch.bvh, &err);
    containers_n1n2(S, scratch.neg, ch.node, c.n1, c.n2, &err);
    out[0] = 1.0; out[1] = c.t;
    out[2] = c.point.x; out[3] = c.point.y; out[4] = c.point.z;
    out[5] = c.normalv.x; out[6] = c.normalv.y; out[7] = c.normalv.z;
    out[8] = c.over_point.x; out[9] = c.over_point.y; out[10] = c.over_point.z;
    out[11] = c.under_point.x; out[12] = c.under_point.y; out[13] = c.under_point.z;
    out[14] = c.reflectv.x; out[15] = c.reflectv.y; out[16] = c.reflectv.z;
    out[17] = c.n1; out[18] = c.n2; out[19] = c.inside ? 1.0 : 0.0;
  }
  RayCounters rc = {0u, 0u, 0u, 0u};
  flush_counters(rc, err, counters);
}

template <bool HAS_CSG>
__global__ void shadow_kernel(SceneView S, V3 point, int light, int* __restrict__ out,
                              unsigned long long* __restrict__ counters) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  Scratch<HAS_CSG> scratch;
  uint32_t err = 0;
  const rtc_light* L = S.lights + light;
  V3 lp = v3(L->position[0], L->position[1], L->position[2]);
  V3 v = lp - point;  // World::is_shadowed, world.rs:95-105
  RayD r;
  r.o = point;
  r.d = normalize(v);
  Query sh;
  sh.init_shadow(magnitude(v), &err);
  search<HAS_CSG, true>(S, r, sh, scratch.csg(), &scratch.bvh, &err);
  out[0] = sh.done() ? 1 : 0;
  RayCounters rc = {0u, 0u, 0u, 0u};
  flush_counters(rc, err, counters);
}

// Known-answer kernel for the device normalize() (shared-reciprocal division, rt_math.cuh)
__global__ void normalize_kernel(const double* __restrict__ xyz, size_t n, double* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    V3 r = normalize(v3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
    out[3 * i] = r.x; out[3 * i + 1] = r.y; out[3 * i + 2] = r.z;
  }
}

// image/ppm.rs:73-75: (v * 255.0).round() as u8 — round half away from zero, saturating, NaN → 0
__global__ void quantize_kernel(const double* __restrict__ rgb, size_t n, uint8_t* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    double r = round(rgb[i] * 255.0);
    uint8_t q;
    if (r != r) q = 0;
    else if (r <= 0.0) q = 0;
    else if (r >= 255.0) q = 255;
    else q = (uint8_t)r;
    out[i] = q;
  }
}

// FP64 roofline denominators. 8 independent chains per thread.
template <bool FUSED>
__global__ void __launch_bounds__(256) fp64_peak_kernel(double* __restrict__ out, int iters, double b, double c) {
  double a0 = threadIdx.x * 1e-3, a1 = a0 + 1.0, a2 = a0 + 2.0, a3 = a0 + 3.0;
  double a4 = a0 + 4.0, a5 = a0 + 5.0, a6 = a0 + 6.0, a7 = a0 + 7.0;
  for (int i = 0; i < iters; i++) {
    if (FUSED) {
      a0 = __fma_rn(a0, b, c); a1 = __fma_rn(a1, b, c); a2 = __fma_rn(a2, b, c); a3 = __fma_rn(a3, b, c);
      a4 = __fma_rn(a4, b, c); a5 = __fma_rn(a5, b, c); a6 = __fma_rn(a6, b, c); a7 = __fma_rn(a7, b, c);
    } else {
      a0 = __dadd_rn(__dmul_rn(a0, b), c); a1 = __dadd_rn(__dmul_rn(a1, b), c);
      a2 = __dadd_rn(__dmul_rn(a2, b), c); a3 = __dadd_rn(__dmul_rn(a3, b), c);
      a4 = __dadd_rn(__dmul_rn(a4, b), c); a5 = __dadd_rn(__dmul_rn(a5, b), c);
      a6 = __dadd_rn(__dmul_rn(a6, b), c); a7 = __dadd_rn(__dmul_rn(a7, b), c);
    }
  }
  double s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (s == 123.456) out[0] = s;  // keep the chains alive
}

}  // namespace rtk
