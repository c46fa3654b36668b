// rt_accel_device.cuh — the acceleration structure of a LARGE mesh built on the GPU (sm_100a), SURVEY §8-f1.
//
// The reference's only hierarchy is Group::divide (group.rs:108-157), which the flattener reproduces because its
// boxes are part of the parity contract; OUR search tree is free (the search result never depends on it: ties resolve by
// node index and every candidate is validated against the reference's group chain). The host builder
// (rt_accel_build.hpp: binned SAH on up to 16 threads) takes ≈ 0.4 s for the million-triangle scene; here the same
// BvhNode / 96-byte LeafRec tables are produced by seven kernels and one radix sort, in milliseconds:
//
//   prim_box_kernel     world-space box of every BVH-eligible leaf from the tables already in the slab (triangles: three
//                       transformed vertices; sphere / cube / bounded cylinder: the eight transformed corners of the
//                       object-space extent), with the host builder's 1e-7 guard band
//   bounds_kernel       centroid bounds (ordered-integer atomics on the doubles)
//   morton_kernel       63-bit Morton key of every centroid
//   cub::DeviceRadixSort  (key, prim) pairs                                                  [library code: the sort]
//   cluster_box_kernel  four consecutive prims in Morton order form one leaf; its box
//   sah_level_kernel    top-down, one launch per level, one warp per node: the node's run of the Morton order is cut
//                       where the surface-area heuristic is cheapest among 31 equal-count cuts (two warp scans); child
//                       boxes (rounded outward to the node's precision) and references go straight into the BvhNode; node 0 is the root
//   emit_kernel         LeafRec (node, transform, group parent, kind, inline triangle geometry) in leaf order
//
// (The plain radix tree over the Morton keys — Karras 2012 — was built first: 0.26 ms for 26 k prims, but 2.3x slower
// to traverse than the host's SAH tree on both mesh scenes. Morton order for locality, SAH for the cuts.) The host
// builder stays in charge below RTC_DEVICE_BVH_MIN prims (default 65 536), for scenes with cubes in the tree (their
// EPSILON-parallel marking walks the tree on the host) and whenever this build reports a non-finite box or a tree
// deeper than the traversal stack.
#pragma once
#include <cub/cub.cuh>

#include "rt_accel.cuh"

namespace rtk {
namespace devbuild {

__device__ __forceinline__ long long enc_ordered(double d) {  // monotone map double -> int64 (no NaN expected)
  const long long b = __double_as_longlong(d);
  return b >= 0 ? b : b ^ 0x7fffffffffffffffLL;
}
__device__ __forceinline__ double dec_ordered(long long e) {
  return __longlong_as_double(e >= 0 ? e : e ^ 0x7fffffffffffffffLL);
}

struct In {
  const rtc_node* nodes;
  const double* fwd;       // 12 per transform (rows 0..2 of the FORWARD matrix), NaN-filled when singular
  const double* limits;
  const double* tri_geom;
  const int32_t* parent;   // per node: the reference Group whose box chain validates it
  const int32_t* cand;     // node index of every BVH prim
  int32_t n;
};

__device__ __forceinline__ void xf_point(const double* m, double x, double y, double z, double out[3]) {
  for (int r = 0; r < 3; r++) out[r] = m[4 * r] * x + m[4 * r + 1] * y + m[4 * r + 2] * z + m[4 * r + 3];
}

// box[6k..6k+5] = lo xyz, hi xyz (padded); bad[0] |= 1 on a non-finite box
__global__ void prim_box_kernel(In in, double* __restrict__ box, int* __restrict__ bad) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= in.n) return;
  const rtc_node nd = in.nodes[in.cand[k]];
  const double* m = in.fwd + 12 * (size_t)nd.xform;
  double lo[3] = {rt_inf(), rt_inf(), rt_inf()}, hi[3] = {-rt_inf(), -rt_inf(), -rt_inf()}, w[3];
  auto add = [&]() {
    for (int a = 0; a < 3; a++) { lo[a] = fmin(lo[a], w[a]); hi[a] = fmax(hi[a], w[a]); }
  };
  if (nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE) {
    const double* g = in.tri_geom + 9 * (size_t)nd.aux;
    xf_point(m, g[0], g[1], g[2], w); add();
    xf_point(m, g[0] + g[3], g[1] + g[4], g[2] + g[5], w); add();
    xf_point(m, g[0] + g[6], g[1] + g[7], g[2] + g[8], w); add();
  } else {
    double y0 = -1.0, y1 = 1.0;
    if (nd.kind == RTC_CYLINDER) { y0 = in.limits[2 * (size_t)nd.aux]; y1 = in.limits[2 * (size_t)nd.aux + 1]; }
    for (int c = 0; c < 8; c++) {
      xf_point(m, (c & 1) ? 1.0 : -1.0, (c & 2) ? y1 : y0, (c & 4) ? 1.0 : -1.0, w);
      add();
    }
  }
  bool ok = true;
  for (int a = 0; a < 3; a++) {
    // guard band: the reference's t comes from object-space arithmetic; 1e-7 relative dwarfs its rounding
    const double pad = 1e-7 * (1.0 + fmax(fabs(lo[a]), fabs(hi[a])) + (hi[a] - lo[a]));
    lo[a] -= pad;
    hi[a] += pad;
    ok = ok && isfinite(lo[a]) && isfinite(hi[a]) && lo[a] <= hi[a];
    box[6 * (size_t)k + a] = lo[a];
    box[6 * (size_t)k + 3 + a] = hi[a];
  }
  if (!ok) atomicOr(bad, 1);
}

// bounds[0..2] = min centroid, [3..5] = max centroid, as ordered integers
__global__ void bounds_kernel(const double* __restrict__ box, int n, long long* __restrict__ bounds) {
  double lo[3] = {rt_inf(), rt_inf(), rt_inf()}, hi[3] = {-rt_inf(), -rt_inf(), -rt_inf()};
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x)
    for (int a = 0; a < 3; a++) {
      const double c = 0.5 * (box[6 * (size_t)k + a] + box[6 * (size_t)k + 3 + a]);
      lo[a] = fmin(lo[a], c);
      hi[a] = fmax(hi[a], c);
    }
  for (int a = 0; a < 3; a++) {
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fmin(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmax(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(bounds + a, enc_ordered(lo[a]));
      atomicMax(bounds + 3 + a, enc_ordered(hi[a]));
    }
  }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long x) {  // 21 bits -> every third bit
  x &= 0x1fffffull;
  x = (x | x << 32) & 0x1f00000000ffffull;
  x = (x | x << 16) & 0x1f0000ff0000ffull;
  x = (x | x << 8) & 0x100f00f00f00f00full;
  x = (x | x << 4) & 0x10c30c30c30c30c3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}

__global__ void morton_kernel(const double* __restrict__ box, int n, const long long* __restrict__ bounds,
                              unsigned long long* __restrict__ keys, uint32_t* __restrict__ idx) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  unsigned long long code = 0;
  for (int a = 0; a < 3; a++) {
    const double lo = dec_ordered(bounds[a]), hi = dec_ordered(bounds[3 + a]);
    const double c = 0.5 * (box[6 * (size_t)k + a] + box[6 * (size_t)k + 3 + a]);
    const double ext = hi - lo;
    double u = ext > 0.0 ? (c - lo) / ext : 0.0;
    u = fmin(fmax(u, 0.0), 1.0);
    const unsigned long long q = (unsigned long long)fmin(u * 2097152.0, 2097151.0);
    code |= spread21(q) << a;
  }
  keys[k] = code;
  idx[k] = (uint32_t)k;
}

#ifndef RTK_DEV_LEAF
#define RTK_DEV_LEAF 4
#endif
constexpr int LEAF_PRIMS = RTK_DEV_LEAF;

// Box of every leaf (LEAF_PRIMS consecutive prims of the Morton order).
__global__ void cluster_box_kernel(const double* __restrict__ box, const uint32_t* __restrict__ sorted_idx, int n_prims, int nc,
                                   double* __restrict__ cbox) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nc) return;
  double lo[3] = {rt_inf(), rt_inf(), rt_inf()}, hi[3] = {-rt_inf(), -rt_inf(), -rt_inf()};
  for (int k = c * LEAF_PRIMS; k < n_prims && k < (c + 1) * LEAF_PRIMS; k++) {
    const double* b = box + 6 * (size_t)sorted_idx[k];
    for (int a = 0; a < 3; a++) { lo[a] = fmin(lo[a], b[a]); hi[a] = fmax(hi[a], b[a + 3]); }
  }
  for (int a = 0; a < 3; a++) { cbox[6 * (size_t)c + a] = lo[a]; cbox[6 * (size_t)c + 3 + a] = hi[a]; }
}

struct Range { int32_t first, count, node, depth; };  // leaves [first, first + count) of the Morton order become node `node`

struct Box3 {
  double lo[3], hi[3];
  __device__ void reset() { for (int a = 0; a < 3; a++) { lo[a] = rt_inf(); hi[a] = -rt_inf(); } }
  __device__ void add(const Box3& o) { for (int a = 0; a < 3; a++) { lo[a] = fmin(lo[a], o.lo[a]); hi[a] = fmax(hi[a], o.hi[a]); } }
  __device__ double area() const {
    const double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx >= 0.0 ? 2.0 * (dx * dy + dy * dz + dz * dx) : 0.0;
  }
  __device__ Box3 shfl(int src) const {
    Box3 r;
    for (int a = 0; a < 3; a++) { r.lo[a] = __shfl_sync(0xffffffffu, lo[a], src); r.hi[a] = __shfl_sync(0xffffffffu, hi[a], src); }
    return r;
  }
};

// One level of a top-down build over the Morton order: one WARP per range. A radix tree over the Morton keys (Karras)
// was built first and measured: 129 ms per frame on the million-triangle scene against 55 ms with the host's SAH tree,
// 4.4 against 2.0 ms on the teapot. Here every node still covers a contiguous run of the Morton order (so nothing is
// ever moved), but WHERE the run is cut is chosen by the surface-area heuristic: the run is cut into 32 equal-count
// chunks (one per lane), prefix and suffix boxes come from two warp scans, and the cheapest of the 31 cuts wins.
__global__ void sah_level_kernel(const Range* __restrict__ in, int n_in, Range* __restrict__ out, int* __restrict__ n_out,
                                 const double* __restrict__ cbox, int n_prims, int32_t shift, TravNode* __restrict__ nodes,
                                 int* __restrict__ n_nodes, int* __restrict__ max_depth) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= n_in) return;
  const Range r = in[w];
  // chunk of this lane: leaves [b0, b1)
  const int chunks = r.count < 32 ? r.count : 32;
  const int b0 = r.first + (int)((long long)r.count * lane / chunks), b1 = r.first + (int)((long long)r.count * (lane + 1) / chunks);
  Box3 mine;
  mine.reset();
  if (lane < chunks)
    for (int c = b0; c < b1; c++) {
      Box3 b;
      for (int a = 0; a < 3; a++) { b.lo[a] = cbox[6 * (size_t)c + a]; b.hi[a] = cbox[6 * (size_t)c + 3 + a]; }
      mine.add(b);
    }
  // inclusive prefix (chunks 0..lane) and suffix (chunks lane..31) unions
  Box3 pre = mine, suf = mine;
  for (int o = 1; o < 32; o <<= 1) {
    Box3 p = pre.shfl(lane - o < 0 ? lane : lane - o), q = suf.shfl(lane + o > 31 ? lane : lane + o);
    if (lane - o >= 0) pre.add(p);
    if (lane + o <= 31) suf.add(q);
  }
  // cut after chunk `lane` (lane = 0 .. chunks-2): left = prefix(lane), right = suffix(lane + 1)
  const Box3 right = suf.shfl(lane + 1 > 31 ? 31 : lane + 1);
  const int n_left = b1 - r.first, n_right = r.count - n_left;
  double cost = rt_inf();
  if (lane < chunks - 1) cost = pre.area() * (double)n_left + right.area() * (double)n_right;
  if (r.depth >= 48) cost = lane == (chunks - 1) / 2 && lane < chunks - 1 ? 0.0 : rt_inf();  // bound the depth: median cut
  if (chunks == 2) cost = lane == 0 ? 0.0 : rt_inf();
  int best = lane;
  double bc = cost;
  for (int o = 16; o > 0; o >>= 1) {
    const double oc = __shfl_xor_sync(0xffffffffu, bc, o);
    const int ob = __shfl_xor_sync(0xffffffffu, best, o);
    if (oc < bc || (oc == bc && ob < best)) { bc = oc; best = ob; }
  }
  if (!(bc < rt_inf())) best = (chunks - 1) / 2 < chunks - 1 ? (chunks - 1) / 2 : 0;  // all costs inf / NaN: the middle
  const Box3 lbox = pre.shfl(best), rbox = suf.shfl(best + 1 > 31 ? 31 : best + 1);
  const int cut = __shfl_sync(0xffffffffu, b1, best);  // first leaf of the right side
  if (lane == 0) {
    const int nl = cut - r.first, nr = r.count - nl;
    auto leaf_count = [&](int c) { const int first = c * LEAF_PRIMS; return n_prims - first < LEAF_PRIMS ? n_prims - first : LEAF_PRIMS; };
    int n_new = (nl > 1 ? 1 : 0) + (nr > 1 ? 1 : 0);
    int id = n_new ? atomicAdd(n_nodes, n_new) : 0;
    int slot = n_new ? atomicAdd(n_out, n_new) : 0;
    int32_t c0, n0, c1, n1;
    if (nl == 1) { c0 = shift + r.first * LEAF_PRIMS; n0 = leaf_count(r.first); }
    else { c0 = id; n0 = 0; out[slot++] = Range{r.first, nl, id, r.depth + 1}; id++; }
    if (nr == 1) { c1 = shift + cut * LEAF_PRIMS; n1 = leaf_count(cut); }
    else { c1 = id; n1 = 0; out[slot] = Range{cut, nr, id, r.depth + 1}; }
#if RTK_BVH_WIDE
    // (the device tree keeps its binary shape: two of the four child slots)
    BvhNode4 nd;
    for (int k = 0; k < 4; k++) {
      nd.lox[k] = nd.loy[k] = nd.loz[k] = HUGE_VALF;
      nd.hix[k] = nd.hiy[k] = nd.hiz[k] = -HUGE_VALF;
      nd.ref[k] = BVH_EMPTY;
      nd._pad[k] = 0;
    }
    nd.lox[0] = bvh_below(lbox.lo[0]); nd.loy[0] = bvh_below(lbox.lo[1]); nd.loz[0] = bvh_below(lbox.lo[2]);
    nd.hix[0] = bvh_above(lbox.hi[0]); nd.hiy[0] = bvh_above(lbox.hi[1]); nd.hiz[0] = bvh_above(lbox.hi[2]);
    nd.lox[1] = bvh_below(rbox.lo[0]); nd.loy[1] = bvh_below(rbox.lo[1]); nd.loz[1] = bvh_below(rbox.lo[2]);
    nd.hix[1] = bvh_above(rbox.hi[0]); nd.hiy[1] = bvh_above(rbox.hi[1]); nd.hiz[1] = bvh_above(rbox.hi[2]);
    nd.ref[0] = n0 > 0 ? ~((c0 << 3) | n0) : c0;
    nd.ref[1] = n1 > 0 ? ~((c1 << 3) | n1) : c1;
#else
    BvhNode nd;
    for (int a = 0; a < 3; a++) {
      nd.lo0[a] = bvh_below(lbox.lo[a]); nd.hi0[a] = bvh_above(lbox.hi[a]);
      nd.lo1[a] = bvh_below(rbox.lo[a]); nd.hi1[a] = bvh_above(rbox.hi[a]);
    }
#if !RTK_BVH_F32
    nd._pad[0] = nd._pad[1] = nd._pad[2] = nd._pad[3] = 0;
#endif
    nd.c0 = c0; nd.n0 = n0; nd.c1 = c1; nd.n1 = n1;
#endif
    nodes[r.node] = nd;
    atomicMax(max_depth, r.depth + 1);
  }
}

__global__ void emit_kernel(In in, const uint32_t* __restrict__ sorted_idx, LeafRec* __restrict__ recs) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= in.n) return;
  const int32_t node = in.cand[sorted_idx[k]];
  const rtc_node nd = in.nodes[node];
  LeafRec r;
  r.node = node; r.xform = nd.xform; r.parent = in.parent[node];
  r.kind_flags = ((uint32_t)nd.kind << 24) | (nd.flags & 0xffffffu);
  r.aux = nd.aux; r._pad = 0;
  for (int c = 0; c < 9; c++) r.g[c] = 0.0;
  if (nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE)
    for (int c = 0; c < 9; c++) r.g[c] = in.tri_geom[9 * (size_t)nd.aux + c];
  recs[k] = r;
}

inline size_t node_count(int n_prims) {
  const int nc = (n_prims + LEAF_PRIMS - 1) / LEAF_PRIMS;
  return nc > 1 ? (size_t)(nc - 1) : 1;
}

// Builds nodes[node_count(n)] and recs[n] (device memory) on `stream`. Returns cudaSuccess and *ok = false when the
// host builder has to take over (non-finite box, tree deeper than the traversal stack).
inline cudaError_t build(const In& in, int32_t shift, TravNode* nodes, LeafRec* recs, cudaStream_t stream, bool* ok, int* depth_out) {
  *ok = false;
  const int n = in.n, nc = (n + LEAF_PRIMS - 1) / LEAF_PRIMS;
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  size_t sort_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, n, 0, 63, stream);
  const size_t o_box = 0, o_keys = o_box + up(sizeof(double) * 6 * (size_t)n), o_keys2 = o_keys + up(8 * (size_t)n),
               o_idx = o_keys2 + up(8 * (size_t)n), o_idx2 = o_idx + up(4 * (size_t)n), o_cbox = o_idx2 + up(4 * (size_t)n),
               o_qa = o_cbox + up(sizeof(double) * 6 * (size_t)nc), o_qb = o_qa + up(sizeof(Range) * (size_t)nc),
               o_misc = o_qb + up(sizeof(Range) * (size_t)nc), o_sort = o_misc + 256, total = o_sort + up(sort_bytes);
  char* tmp = nullptr;
  cudaError_t e = cudaMallocAsync((void**)&tmp, total, stream);
  if (e != cudaSuccess) return e;
  double* box = (double*)(tmp + o_box);
  unsigned long long *keys = (unsigned long long*)(tmp + o_keys), *keys2 = (unsigned long long*)(tmp + o_keys2);
  uint32_t *idx = (uint32_t*)(tmp + o_idx), *idx2 = (uint32_t*)(tmp + o_idx2);
  double* cbox = (double*)(tmp + o_cbox);
  Range *qa = (Range*)(tmp + o_qa), *qb = (Range*)(tmp + o_qb);
  long long* bounds = (long long*)(tmp + o_misc);      // 6 x 8 bytes
  int* flags = (int*)(tmp + o_misc + 64);              // [0] bad box, [1] depth, [2] nodes allocated, [3] ranges queued
  const long long init_bounds[6] = {0x7fffffffffffffffLL, 0x7fffffffffffffffLL, 0x7fffffffffffffffLL,
                                    (long long)0x8000000000000000ull, (long long)0x8000000000000000ull, (long long)0x8000000000000000ull};
  const int T = 256;
  auto blocks = [&](int m) { return (m + T - 1) / T; };
  struct Cleanup { char* p; cudaStream_t s; ~Cleanup() { cudaFreeAsync(p, s); } } cleanup{tmp, stream};
#define DEVB_TRY(x) do { e = (x); if (e != cudaSuccess) return e; } while (0)
  DEVB_TRY(cudaMemcpyAsync(bounds, init_bounds, sizeof init_bounds, cudaMemcpyHostToDevice, stream));
  DEVB_TRY(cudaMemsetAsync(flags, 0, 64, stream));
  prim_box_kernel<<<blocks(n), T, 0, stream>>>(in, box, flags);
  bounds_kernel<<<148 * 4, T, 0, stream>>>(box, n, bounds);
  morton_kernel<<<blocks(n), T, 0, stream>>>(box, n, bounds, keys, idx);
  DEVB_TRY(cub::DeviceRadixSort::SortPairs(tmp + o_sort, sort_bytes, keys, keys2, idx, idx2, n, 0, 63, stream));
  if (nc >= 2) {
    cluster_box_kernel<<<blocks(nc), T, 0, stream>>>(box, idx2, n, nc, cbox);
    const Range root{0, nc, 0, 1};
    const int one = 1;
    DEVB_TRY(cudaMemcpyAsync(qa, &root, sizeof root, cudaMemcpyHostToDevice, stream));
    DEVB_TRY(cudaMemcpyAsync(flags + 2, &one, sizeof one, cudaMemcpyHostToDevice, stream));  // node 0 is the root
    int n_in = 1;
    for (int level = 0; n_in > 0 && level < 64; level++) {
      DEVB_TRY(cudaMemsetAsync(flags + 3, 0, sizeof(int), stream));
      sah_level_kernel<<<blocks(n_in * 32), T, 0, stream>>>(qa, n_in, qb, flags + 3, cbox, n, shift, nodes, flags + 2, flags + 1);
      DEVB_TRY(cudaMemcpyAsync(&n_in, flags + 3, sizeof(int), cudaMemcpyDeviceToHost, stream));
      DEVB_TRY(cudaStreamSynchronize(stream));
      Range* t = qa; qa = qb; qb = t;
    }
  } else {
    // a single leaf: the root's first child holds it, the second is empty (n1 < 0); its box is never culled wrongly
    // by an all-covering extent
#if RTK_BVH_WIDE
    BvhNode4 root;
    for (int k = 0; k < 4; k++) {
      root.lox[k] = root.loy[k] = root.loz[k] = k == 0 ? -HUGE_VALF : HUGE_VALF;
      root.hix[k] = root.hiy[k] = root.hiz[k] = k == 0 ? HUGE_VALF : -HUGE_VALF;
      root.ref[k] = BVH_EMPTY;
      root._pad[k] = 0;
    }
    root.ref[0] = ~((shift << 3) | n);
#else
    BvhNode root;
    for (int a = 0; a < 3; a++) {
      root.lo0[a] = bvh_below(-1.7976931348623157e308); root.hi0[a] = bvh_above(1.7976931348623157e308);
      root.lo1[a] = bvh_above(1.7976931348623157e308); root.hi1[a] = bvh_below(-1.7976931348623157e308);
    }
    root.c0 = shift; root.n0 = n; root.c1 = 0; root.n1 = -1;
#endif
    DEVB_TRY(cudaMemcpyAsync(nodes, &root, sizeof root, cudaMemcpyHostToDevice, stream));
  }
  emit_kernel<<<blocks(n), T, 0, stream>>>(in, idx2, recs);
  DEVB_TRY(cudaGetLastError());
  int h[2] = {0, 0};
  DEVB_TRY(cudaMemcpyAsync(h, flags, sizeof h, cudaMemcpyDeviceToHost, stream));
  DEVB_TRY(cudaStreamSynchronize(stream));
#undef DEVB_TRY
  if (depth_out) *depth_out = h[1];
  *ok = h[0] == 0 && h[1] < BVH_STACK - 2;
  return cudaSuccess;
}

}  // namespace devbuild
}  // namespace rtk
