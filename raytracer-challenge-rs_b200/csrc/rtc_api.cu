// rtc_api.cu — the C-ABI of include/rtc.h over the sm_100a kernels. Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo --fmad=false -Xcompiler -fPIC,-ffp-contract=off
//        -shared -o librtc_cuda.so rtc_api.cu
// No CPU fallback: every compute entry point needs a CUDA device.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstdio>
#include <cctype>
#include <cerrno>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <sys/stat.h>
#include <unistd.h>

#include "rt_accel_device.cuh"
#include "rt_stream.cuh"

using namespace rtk;

namespace {

thread_local std::string g_err;
thread_local bool g_host_build_only = false;  // rtc_scene_create retrying after the device builder declined

int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define CUDA_TRY(expr)                                                                          \
  do {                                                                                          \
    cudaError_t e_ = (expr);                                                                    \
    if (e_ != cudaSuccess) {                                                                    \
      return fail(RTC_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_));            \
    }                                                                                           \
  } while (0)

constexpr int N_COUNTERS = 8;       // what a frame resets and rtc_scene_finish reads
constexpr int N_COUNTERS_ALLOC = 24;  // [8, 24): RTK_STREAM_STATS instrumentation (experiments)
#ifndef RTK_WAVE_MIN_RANKS
#define RTK_WAVE_MIN_RANKS 3
#endif

}  // namespace

struct rtc_scene {
  int device = 0;
  char* slab = nullptr;  // one device allocation holding every table
  size_t slab_bytes = 0;
  SceneView view{};
  bool has_csg = false;
  bool big_accel = false;  // BVH + leaf records well beyond L2: launch the 32-warps-per-SM kernel
  unsigned long long* d_counters = nullptr;
  double* d_canvas = nullptr;  // cached device canvas for rtc_render
  size_t canvas_doubles = 0;
  double* d_probe = nullptr;   // 32 doubles scratch for the debug entries
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  uint32_t launches = 0;
  int sm_count = 148;
  // wavefront path (rt_wave.cuh): queue arena, and what rtc_scene_finish needs to redo the frame with
  // the recursive kernel should a queue have overflowed
  char* wave_arena = nullptr;
  size_t wave_bytes = 0;
  // streaming wavefront: the arena's child-colour slots hold sentinels for this (pixels, depth) layout; 0 = they do not
  // (fresh or recycled arena, a level-synchronous frame, an abandoned frame): re-fill before the next streaming frame
  size_t wave_clean_px = 0;
  int wave_clean_depth = 0;
  bool lean = false;           // every-ray spheres and planes only: the kernels without BVH / generic leaf code
  bool tri_sph_only = false;   // spheres/planes every ray + a BVH of triangles and spheres: no generic leaf code
  bool has_branching = false;  // some material reflects AND refracts: shade_hit spawns two children, the tree grows as 2^depth
  bool last_wave = false;
  uint32_t stat_flags = 0;     // RTC_STAT_* of the frame in flight
  cudaStream_t busy_stream = nullptr;  // stream of a frame queued by rtc_render_device and not yet finished ...
  bool busy = false;                   // ... rtc_scene_destroy and a render on another stream wait for it
  CameraD last_cam{};
  RenderParams last_P{};
  double* last_out = nullptr;
};

constexpr size_t HOST_CANVAS_CTRL = 4096;  // arrival words of up to 64 ranks behind the pixels, one per 64-byte line
struct rtc_host_canvas {
  std::string name;
  double* ptr = nullptr;
  size_t bytes = 0;        // pixels + control
  int device = 0;
  bool creator = false, registered = false;
};

struct rtc_canvas {
  int device = 0;
  double* ptr = nullptr;
  size_t n = 0;
  bool opened = false;  // came from an IPC handle
};

namespace {

// Structural validation of a flattened world; also derives csg_parent[] and the CSG nesting depth.
int validate(const rtc_scene_desc* d, std::vector<int32_t>& csg_parent, bool& has_csg) {
  if (!d) return fail(RTC_ERR_INVALID, "desc is NULL");
  if (d->abi_version != RTC_ABI_VERSION) return fail(RTC_ERR_INVALID, "abi_version mismatch");
  if (d->n_nodes > 0 && !d->nodes) return fail(RTC_ERR_INVALID, "nodes is NULL");
  if (d->n_nodes > 0x7fffff00u) return fail(RTC_ERR_INVALID, "too many nodes");
  // every non-empty table has a pointer — before any table is read (here, in build_accel, or by the upload)
  if ((d->n_xforms && !d->xforms) || (d->n_boxes && !d->boxes) || (d->n_limits && !d->limits) ||
      (d->n_triangles && (!d->tri_p1e1e2 || !d->tri_normals)) || (d->n_materials && !d->materials) ||
      (d->n_patterns && !d->patterns) || (d->n_lights && !d->lights))
    return fail(RTC_ERR_INVALID, "a non-empty table has a NULL pointer");
  const int32_t n = (int32_t)d->n_nodes;
  csg_parent.assign((size_t)n, -1);
  has_csg = false;
  // stack of enclosing aggregates: (end index, is_csg, node index)
  struct Open { int32_t end; bool csg; int32_t node; };
  std::vector<Open> open;
  for (int32_t i = 0; i < n; i++) {
    while (!open.empty() && open.back().end == i) open.pop_back();
    const rtc_node& nd = d->nodes[i];
    // (the message prefix is only formatted on failure: a snprintf per node was 60 ms of a million-node scene)
    auto where = [i]() { return "node " + std::to_string(i) + ": "; };
    int32_t limit = open.empty() ? n : open.back().end;
    for (auto it = open.rbegin(); it != open.rend(); ++it)
      if (it->csg) { csg_parent[(size_t)i] = it->node; break; }
    if (nd.kind < RTC_SPHERE || nd.kind > RTC_CSG) return fail(RTC_ERR_INVALID, where() + "bad kind");
    bool aggregate = nd.kind == RTC_GROUP || nd.kind == RTC_CSG;
    if (!aggregate) {
      if (nd.skip != i + 1) return fail(RTC_ERR_INVALID, where() + "leaf skip must be index+1");
      if (nd.xform < 0 || (uint32_t)nd.xform >= d->n_xforms) return fail(RTC_ERR_INVALID, where() + "xform out of range");
      if (nd.material < 0 || (uint32_t)nd.material >= d->n_materials) return fail(RTC_ERR_INVALID, where() + "material out of range");
      if ((nd.kind == RTC_TRIANGLE || nd.kind == RTC_SMOOTH_TRIANGLE) && (nd.aux < 0 || (uint32_t)nd.aux >= d->n_triangles))
        return fail(RTC_ERR_INVALID, where() + "triangle index out of range");
      if ((nd.kind == RTC_CYLINDER || nd.kind == RTC_CONE) && (nd.aux < 0 || (uint32_t)nd.aux >= d->n_limits))
        return fail(RTC_ERR_INVALID, where() + "limits index out of range");
    } else {
      if (nd.skip <= i || nd.skip > limit) return fail(RTC_ERR_INVALID, where() + "skip out of range");
      if (nd.aux < 0 || (uint32_t)nd.aux >= d->n_boxes) return fail(RTC_ERR_INVALID, where() + "box index out of range");
      if (nd.kind == RTC_CSG) {
        has_csg = true;
        if (nd.xform < 0 || (uint32_t)nd.xform >= d->n_xforms) return fail(RTC_ERR_INVALID, where() + "xform out of range");
        if (nd.right <= i + 1 || nd.right >= nd.skip) return fail(RTC_ERR_INVALID, where() + "CSG right index out of range");
        if (d->nodes[i + 1].skip != nd.right) return fail(RTC_ERR_INVALID, where() + "CSG left subtree must end at right");
        if (d->nodes[nd.right].skip != nd.skip) return fail(RTC_ERR_INVALID, where() + "CSG right subtree must end at skip");
        if (((nd.flags >> 8) & 3u) > 2u) return fail(RTC_ERR_INVALID, where() + "bad CSG op");
        int depth = 1;
        for (auto& o : open) depth += o.csg ? 1 : 0;
        if (depth > CSG_MAX_DEPTH) return fail(RTC_ERR_UNSUPPORTED, where() + "CSG nesting deeper than 4");
      }
      if (nd.skip > i + 1) open.push_back({nd.skip, nd.kind == RTC_CSG, i});
    }
    if (nd.skip > limit) return fail(RTC_ERR_INVALID, where() + "subtree crosses its parent");
  }
  for (uint32_t m = 0; m < d->n_materials; m++) {
    int32_t p = d->materials[m].pattern;
    if (p >= 0 && (uint32_t)p >= d->n_patterns) return fail(RTC_ERR_INVALID, "material pattern out of range");
  }
  return RTC_OK;
}

size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }


int read_counters(rtc_scene* s, cudaStream_t stream, rtc_stats* stats) {
  unsigned long long h[N_COUNTERS];
  CUDA_TRY(cudaMemcpyAsync(h, s->d_counters, sizeof h, cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(cudaStreamSynchronize(stream));
#if defined(RTK_DEVICE_STATS)
  {
    static unsigned long long host[256][8];
    unsigned long long* dev = nullptr;
    if (cudaGetSymbolAddress((void**)&dev, g_dev_stats) == cudaSuccess && cudaMemcpy(host, dev, sizeof host, cudaMemcpyDeviceToHost) == cudaSuccess) {
      unsigned long long sum[8] = {0};
      for (int m = 0; m < 256; m++) for (int k = 0; k < 8; k++) sum[k] += host[m][k];
      cudaMemset(dev, 0, sizeof host);
      const unsigned long long rays = h[0] + h[1] + h[2] + h[3];
      // every-ray prims: one 128-byte PrimRec each; BVH: 128-byte nodes, 96-byte leaf records (prim_tests counts both kinds of prim)
      std::fprintf(stderr, "device stats: rays %llu searches %llu bvh nodes %llu prim tests %llu chain calls %llu chain boxes %llu | per ray: %.2f nodes, %.2f prims\n",
                   rays, sum[DST_walks], sum[DST_bvh_nodes], sum[DST_prim_tests], sum[DST_chain_calls], sum[DST_chain_boxes],
                   rays ? (double)sum[DST_bvh_nodes] / rays : 0.0, rays ? (double)sum[DST_prim_tests] / rays : 0.0);
    }
  }
#endif
#if defined(RTK_STREAM_STATS)
  if (s->stat_flags & RTC_STAT_WAVEFRONT) {
    unsigned long long d[16];
    CUDA_TRY(cudaMemcpy(d, s->d_counters + 8, sizeof d, cudaMemcpyDeviceToHost));
    std::fprintf(stderr, "stream stats: batches %llu items %llu idle polls %llu | clocks/warp-sum: acquire %llu stamp-wait %llu work %llu | stamp spins %llu partial batches %llu CAS failures %llu\n",
                 d[0], d[1], d[2], d[3], d[4], d[5], d[6], d[7], d[8]);
  }
#endif
  if (h[7]) {
    s->stat_flags |= RTC_STAT_WAVE_OVERFLOWED;
    CUDA_TRY(cudaMemsetAsync(s->d_counters + 7, 0, sizeof(unsigned long long), stream));
  }
  if (stats) {
    stats->rays_primary = h[0]; stats->rays_shadow = h[1]; stats->rays_reflect = h[2]; stats->rays_refract = h[3];
    stats->overflow_count = h[4]; stats->nan_count = h[5];
  }
  if (h[4]) return fail(RTC_ERR_OVERFLOW, "a bounded intersection list overflowed on the device (CSG/containers, cap 64)");
  if (h[5]) return fail(RTC_ERR_NAN, "NaN t reached the intersection sort (the reference panics: intersection.rs:113)");
  return RTC_OK;
}

int make_camera(const rtc_camera* c, CameraD& cam) {
  if (!c || c->hsize == 0 || c->vsize == 0) return fail(RTC_ERR_INVALID, "bad camera");
  cam.pixel_size = c->pixel_size; cam.half_width = c->half_width; cam.half_height = c->half_height;
  for (int i = 0; i < 12; i++) cam.inv[i] = c->inverse[i];
  // &self.transform_inverse * Point::origin()  (camera.rs:65; matrix.rs:232-245), same expression
  const double z = 0.0;
  volatile double ox = c->inverse[0] * z + c->inverse[1] * z + c->inverse[2] * z + c->inverse[3];
  volatile double oy = c->inverse[4] * z + c->inverse[5] * z + c->inverse[6] * z + c->inverse[7];
  volatile double oz = c->inverse[8] * z + c->inverse[9] * z + c->inverse[10] * z + c->inverse[11];
  cam.origin = v3(ox, oy, oz);
  cam.hsize = c->hsize; cam.vsize = c->vsize;
  return RTC_OK;
}

int make_params(const rtc_camera* c, const rtc_opts* o, RenderParams& P, bool& faithful) {
  rtc_opts d{};
  d.aa_samples = 1; d.max_depth = 5;
  if (o) d = *o;
  if (!(d.aa_samples == 1 || d.aa_samples == 2 || d.aa_samples == 4 || d.aa_samples == 8 || d.aa_samples == 16))
    return fail(RTC_ERR_INVALID, "aa_samples must be 1, 2, 4, 8 or 16 (camera.rs:227-233)");
  if ((int)d.max_depth + 1 > MAX_FRAMES) return fail(RTC_ERR_UNSUPPORTED, "max_depth > 7");
  P.row_begin = d.row_begin; P.row_end = d.row_end;
  if (P.row_begin == 0 && P.row_end == 0) P.row_end = c->vsize;
  if (P.row_end > c->vsize) P.row_end = c->vsize;
  if (P.row_begin > P.row_end) return fail(RTC_ERR_INVALID, "row_begin > row_end");
  P.nranks = d.nranks ? d.nranks : 1;
  P.rank = d.rank;
  if (P.rank >= P.nranks) return fail(RTC_ERR_INVALID, "rank >= nranks");
  P.tile_rows = d.tile_rows ? d.tile_rows : 8;
  uint32_t n_tiles = (c->vsize + P.tile_rows - 1) / P.tile_rows;
  uint32_t owned_tiles = n_tiles > P.rank ? (n_tiles - P.rank + P.nranks - 1) / P.nranks : 0;
  P.owned_rows = owned_tiles * P.tile_rows;
  P.aa = d.aa_samples;
  P.average = (d.aa_samples > 1 || (d.flags & RTC_OPT_AVERAGE)) ? 1u : 0u;
  P.max_depth = (int32_t)d.max_depth;
  P.flags = d.flags;
  faithful = (d.flags & RTC_OPT_FAITHFUL_TRAVERSAL) != 0;
  return RTC_OK;
}

// Centre-out dispatch order of a gx x gy grid of tiles, cached per (device, gx, gy) for the life of the process
// (64 KB at 1080p; building it costs a sort, which the per-step scene of the end-to-end path must not pay).
struct OrderCache {
  std::mutex mu;
  struct Entry { int dev; uint32_t gx, gy; uint32_t* ptr; };
  std::vector<Entry> entries;
  const uint32_t* get(int dev, uint32_t gx, uint32_t gy) {
    if (gx == 0 || gy == 0 || gx > 0xffffu || gy > 0xffffu) return nullptr;
    std::lock_guard<std::mutex> g(mu);
    for (const Entry& e : entries)
      if (e.dev == dev && e.gx == gx && e.gy == gy) return e.ptr;
    std::vector<std::pair<double, uint32_t>> key((size_t)gx * gy);
    for (uint32_t y = 0; y < gy; y++)
      for (uint32_t x = 0; x < gx; x++) {
        const double dx = (x + 0.5) / gx - 0.5, dy = (y + 0.5) / gy - 0.5;
        key[(size_t)y * gx + x] = {dx * dx + dy * dy, (y << 16) | x};
      }
    std::stable_sort(key.begin(), key.end(), [](const std::pair<double, uint32_t>& a, const std::pair<double, uint32_t>& b) { return a.first < b.first; });
    std::vector<uint32_t> h(key.size());
    for (size_t i = 0; i < key.size(); i++) h[i] = key[i].second;
    uint32_t* d = nullptr;
    if (cudaMalloc((void**)&d, h.size() * sizeof(uint32_t)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (cudaMemcpy(d, h.data(), h.size() * sizeof(uint32_t), cudaMemcpyHostToDevice) != cudaSuccess) { cudaGetLastError(); cudaFree(d); return nullptr; }
    entries.push_back({dev, gx, gy, d});
    return d;
  }
};
OrderCache g_order_cache;

// One launch over owned rows [oy0, oy0 + rows) of the launch domain (rows == 0: the whole domain).
int launch_rows(rtc_scene* s, const CameraD& cam, const RenderParams& P, bool faithful, double* out, cudaStream_t stream,
                uint32_t oy0, uint32_t rows) {
  RenderParams Pk = P;
  Pk.faithful = faithful ? 1u : 0u;
  Pk.oy0 = oy0;
  if (rows == 0) rows = P.owned_rows > oy0 ? P.owned_rows - oy0 : 0;
  if (oy0 + rows < Pk.owned_rows) Pk.owned_rows = oy0 + rows;
  dim3 block(128), grid((cam.hsize + 15) / 16, (rows + 7) / 8);
  {
    static const char* env = std::getenv("RTC_TILE_ORDER");  // 0: row-major (experiments)
    // not for accelerators far beyond L2: there row-major neighbours share DRAM-resident nodes, and rings around
    // the centre touch the whole mesh at once (1 M triangles at 4K on two GPUs: 31.6 -> 35.4 ms)
    Pk.order = ((env && env[0] == '0') || s->big_accel) ? nullptr : g_order_cache.get(s->device, grid.x, grid.y);
  }
  if (grid.x > 0 && grid.y > 0) {
    if (s->has_csg) {
      if (faithful) render_kernel<true, true><<<grid, block, 0, stream>>>(s->view, cam, Pk, out, s->d_counters);
      else if (s->big_accel) render_kernel<true, false, RTK_BIG_MINB><<<grid, block, 0, stream>>>(s->view, cam, Pk, out, s->d_counters);
      else render_kernel<true, false><<<grid, block, 0, stream>>>(s->view, cam, Pk, out, s->d_counters);
    } else if (s->lean && !faithful) {
      render_kernel<false, 2><<<grid, block, 0, stream>>>(s->view, cam, Pk, out, s->d_counters);
    } else if (s->tri_sph_only && !faithful) {
      if (s->big_accel) render_kernel<false, 4, RTK_BIG_MINB><<<grid, block, 0, stream>>>(s->view, cam, Pk, out, s->d_counters);
      else render_kernel<false, 4, RTK_BVH_MINB><<<grid, block, 0, stream>>>(s->view, cam, Pk, out, s->d_counters);
    } else {
      if (faithful) render_kernel<false, true><<<grid, block, 0, stream>>>(s->view, cam, Pk, out, s->d_counters);
      else if (s->big_accel) render_kernel<false, false, RTK_BIG_MINB><<<grid, block, 0, stream>>>(s->view, cam, Pk, out, s->d_counters);
      else render_kernel<false, false, RTK_BVH_MINB><<<grid, block, 0, stream>>>(s->view, cam, Pk, out, s->d_counters);
    }
    s->launches++;
  }
  CUDA_TRY(cudaGetLastError());
  return RTC_OK;
}

#ifndef RTK_STREAM_MIN_RANKS
#define RTK_STREAM_MIN_RANKS 2
#endif

// Scratch arenas are big (GBs at 4K): keep one freed arena per device, like the canvases.
struct ArenaPool {
  std::mutex mu;
  char* ptr[16] = {nullptr};
  size_t bytes[16] = {0};
  size_t clean_px[16] = {0};
  int clean_depth[16] = {0};
  char* take(int dev, size_t need, size_t& got, size_t& cpx, int& cdepth) {
    std::lock_guard<std::mutex> g(mu);
    if (dev >= 0 && dev < 16 && ptr[dev] && bytes[dev] >= need) {
      char* p = ptr[dev];
      got = bytes[dev];
      cpx = clean_px[dev];
      cdepth = clean_depth[dev];
      ptr[dev] = nullptr;
      bytes[dev] = 0;
      return p;
    }
    return nullptr;
  }
  void give(int dev, char* p, size_t n, size_t cpx, int cdepth) {
    {
      std::lock_guard<std::mutex> g(mu);
      if (dev >= 0 && dev < 16 && bytes[dev] < n) {
        std::swap(ptr[dev], p);
        std::swap(bytes[dev], n);
        clean_px[dev] = cpx;
        clean_depth[dev] = cdepth;
      }
    }
    if (p) cudaFree(p);
  }
};
ArenaPool g_arena_pool;

int ensure_arena(rtc_scene* s, size_t need) {
  if (s->wave_bytes >= need) return RTC_OK;
  if (s->wave_arena) g_arena_pool.give(s->device, s->wave_arena, s->wave_bytes, s->wave_clean_px, s->wave_clean_depth);
  s->wave_arena = nullptr;
  s->wave_bytes = 0;
  s->wave_clean_px = 0;
  s->wave_clean_depth = 0;
  size_t got = 0;
  s->wave_arena = g_arena_pool.take(s->device, need, got, s->wave_clean_px, s->wave_clean_depth);
  if (s->wave_arena) {
    s->wave_bytes = got;
  } else {
    CUDA_TRY(cudaMalloc((void**)&s->wave_arena, need));
    s->wave_bytes = need;
    // the streaming wavefront reads ready stamps out of queue slots nobody has written yet: start from zero
    // (epochs are never 0); a recycled arena only holds stamps of earlier, different epochs
    CUDA_TRY(cudaMemsetAsync(s->wave_arena, 0, need, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
  }
  return RTC_OK;
}

template <bool HAS_CSG, int MINB, int MODE = 0>
void launch_wave_level(rtc_scene* s, const CameraD& cam, const RenderParams& P, const WaveArgs& A, bool primary, double* out,
                       cudaStream_t stream) {
  const int grid = s->sm_count * MINB;
  if (primary) wave_forward_kernel<HAS_CSG, true, MINB, MODE><<<grid, 128, 0, stream>>>(s->view, cam, P, A, out, s->d_counters);
  else wave_forward_kernel<HAS_CSG, false, MINB, MODE><<<grid, 128, 0, stream>>>(s->view, cam, P, A, out, s->d_counters);
  s->launches++;
}

// The frame as a level-synchronous wavefront (rt_wave.cuh): 1 + max_depth forward passes and max_depth
// combine passes, queued back to back; no host synchronisation in between.
int launch_wave(rtc_scene* s, const CameraD& cam, const RenderParams& P, double* out, cudaStream_t stream) {
  const int D = P.max_depth;
  const size_t px = (size_t)P.owned_rows * cam.hsize;
  if (px == 0) return RTC_OK;
  if (px > 0x3fffffffu) return fail(RTC_ERR_UNSUPPORTED, "frame too large for the wavefront queues");
  // capacities: level 0 has one call per pixel; deeper levels get 2 per pixel (worst case 2^L: overflow →
  // counters[6] → rtc_scene_finish re-renders with the recursive kernel)
  size_t off = 256;  // [0,256): the counters below
  struct Lay { size_t items, recs, child; uint32_t cap_items, cap_recs; } lay[WAVE_MAX_LEVELS + 1] = {};
  for (int L = 0; L <= D; L++) {
    lay[L].cap_items = L == 0 ? 0u : (uint32_t)(2 * px);
    lay[L].cap_recs = L == D ? 0u : (L == 0 ? (uint32_t)px : (uint32_t)(2 * px));
    lay[L].items = off; off += align_up((size_t)lay[L].cap_items * sizeof(WaveItem));
    lay[L].recs = off; off += align_up((size_t)lay[L].cap_recs * sizeof(WaveRec));
    lay[L].child = off; off += align_up((size_t)lay[L].cap_recs * 6 * sizeof(double));
  }
  int rc = ensure_arena(s, off);
  if (rc) return rc;
  s->wave_clean_px = 0;  // these passes leave colours in the child slots
  char* base = s->wave_arena;
  unsigned long long* n_items = (unsigned long long*)base;  // [0,128): u64 per level (front | rear << 32)
  uint32_t* n_recs = (uint32_t*)(base + 128);               // [128,192)
  uint32_t* work = (uint32_t*)(base + 192);                 // [192,256)
  CUDA_TRY(cudaMemsetAsync(base, 0, 256, stream));
  WaveArgs A{};
  for (int L = 0; L <= D; L++) {
    A.lv[L].items = (WaveItem*)(base + lay[L].items);
    A.lv[L].recs = (WaveRec*)(base + lay[L].recs);
    A.lv[L].child = (double*)(base + lay[L].child);
    A.lv[L].cap_items = lay[L].cap_items;
    A.lv[L].cap_recs = lay[L].cap_recs;
  }
  A.n_items = n_items;
  A.n_recs = n_recs;
  A.max_depth = D;
  auto level = [&](int L) { return A.lv[L]; };
  for (int L = 0; L <= D; L++) {
    A.level = L;
    A.work = work + L;
    if (s->has_csg) {
      if (s->big_accel) launch_wave_level<true, RTK_BIG_MINB>(s, cam, P, A, L == 0, out, stream);
      else launch_wave_level<true, RTK_MIN_BLOCKS>(s, cam, P, A, L == 0, out, stream);
    } else if (s->lean) {
      launch_wave_level<false, RTK_MIN_BLOCKS, 2>(s, cam, P, A, L == 0, out, stream);
    } else if (s->tri_sph_only) {
      if (s->big_accel) launch_wave_level<false, RTK_BIG_MINB, 4>(s, cam, P, A, L == 0, out, stream);
      else launch_wave_level<false, RTK_BVH_MINB, 4>(s, cam, P, A, L == 0, out, stream);
    } else {
      if (s->big_accel) launch_wave_level<false, RTK_BIG_MINB>(s, cam, P, A, L == 0, out, stream);
      else launch_wave_level<false, RTK_BVH_MINB>(s, cam, P, A, L == 0, out, stream);
    }
  }
  for (int L = D - 1; L >= 0; L--) {
    const WaveLevel w = level(L);
    const int grid = s->sm_count * 4;
    if (L == 0) wave_combine_kernel<true><<<grid, 256, 0, stream>>>(w.recs, w.child, n_recs + L, w.cap_recs, nullptr, out, s->d_counters);
    else wave_combine_kernel<false><<<grid, 256, 0, stream>>>(w.recs, w.child, n_recs + L, w.cap_recs, level(L - 1).child, out, s->d_counters);
    s->launches++;
  }
  CUDA_TRY(cudaGetLastError());
  return RTC_OK;
}

std::atomic<uint32_t> g_wave_epoch{1};

// The same wavefront as ONE persistent launch without level barriers (rt_stream.cuh).
int launch_stream(rtc_scene* s, const CameraD& cam, const RenderParams& P, double* out, cudaStream_t stream) {
  const int D = P.max_depth;
  const size_t px = (size_t)P.owned_rows * cam.hsize;
  if (px == 0) return RTC_OK;
  if (px > 0x3fffffffu) return fail(RTC_ERR_UNSUPPORTED, "frame too large for the wavefront queues");
  size_t off = STREAM_CTRL_BYTES;  // the control block (rt_stream.cuh: one word per 128-byte line)
  StreamArgs A{};
  struct Lay { size_t items, recs, child; } lay[WAVE_MAX_LEVELS + 1] = {};
  for (int L = 0; L <= D; L++) {
    A.lv[L].cap_items = L == 0 ? 0u : (uint32_t)(2 * px);
    A.lv[L].cap_recs = L == D ? 0u : (L == 0 ? (uint32_t)px : (uint32_t)(2 * px));
    lay[L].items = off; off += align_up((size_t)A.lv[L].cap_items * sizeof(WaveItem));
    lay[L].recs = off; off += align_up((size_t)A.lv[L].cap_recs * sizeof(WaveRec));
    lay[L].child = off; off += align_up((size_t)A.lv[L].cap_recs * 6 * sizeof(double));
  }
  int rc = ensure_arena(s, off);
  if (rc) return rc;
  char* base = s->wave_arena;
  for (int L = 0; L <= D; L++) {
    A.lv[L].items = (WaveItem*)(base + lay[L].items);
    A.lv[L].recs = (WaveRec*)(base + lay[L].recs);
    A.lv[L].child = (double*)(base + lay[L].child);
  }
  if (s->wave_clean_px != px || s->wave_clean_depth != D) {
    // child colour slots start as sentinels ("nothing delivered"); a frame that completes leaves them that way
    for (int L = 0; L < D; L++) {
      const size_t n = (size_t)A.lv[L].cap_recs * 6;
      if (n) wave_fill_kernel<<<s->sm_count * 8, 256, 0, stream>>>((unsigned long long*)A.lv[L].child, n, WAVE_SENTINEL);
    }
    CUDA_TRY(cudaGetLastError());
    s->wave_clean_px = px;
    s->wave_clean_depth = D;
  }
  A.ctrl = base;
  A.max_depth = D;
  {
    static const char* env = std::getenv("RTC_STREAM_FLAGS");
    A.flags = env ? (uint32_t)std::atoi(env) : (STREAM_JOINT_DELIVER);
  }
  uint32_t e = g_wave_epoch.fetch_add(1) & 0x7fffffffu;
  if (e == 0) e = g_wave_epoch.fetch_add(1) & 0x7fffffffu;
  A.epoch = e;
  CUDA_TRY(cudaMemsetAsync(base, 0, STREAM_CTRL_BYTES, stream));
#if defined(RTK_STREAM_STATS)
  CUDA_TRY(cudaMemsetAsync(s->d_counters + 8, 0, 16 * sizeof(unsigned long long), stream));
#endif
  const int minb = s->big_accel ? RTK_BIG_MINB : ((s->has_csg || s->lean) ? RTK_MIN_BLOCKS : RTK_BVH_MINB);
  const int grid = s->sm_count * minb;
  if (s->has_csg) {
    if (s->big_accel) wave_stream_kernel<true, RTK_BIG_MINB><<<grid, 128, 0, stream>>>(s->view, cam, P, A, out, s->d_counters);
    else wave_stream_kernel<true, RTK_MIN_BLOCKS><<<grid, 128, 0, stream>>>(s->view, cam, P, A, out, s->d_counters);
  } else if (s->lean) {
    wave_stream_kernel<false, RTK_MIN_BLOCKS, 2><<<s->sm_count * RTK_MIN_BLOCKS, 128, 0, stream>>>(s->view, cam, P, A, out, s->d_counters);
  } else if (s->tri_sph_only) {
    if (s->big_accel) wave_stream_kernel<false, RTK_BIG_MINB, 4><<<grid, 128, 0, stream>>>(s->view, cam, P, A, out, s->d_counters);
    else wave_stream_kernel<false, RTK_BVH_MINB, 4><<<grid, 128, 0, stream>>>(s->view, cam, P, A, out, s->d_counters);
  } else {
    if (s->big_accel) wave_stream_kernel<false, RTK_BIG_MINB><<<grid, 128, 0, stream>>>(s->view, cam, P, A, out, s->d_counters);
    else wave_stream_kernel<false, RTK_BVH_MINB><<<grid, 128, 0, stream>>>(s->view, cam, P, A, out, s->d_counters);
  }
  s->launches++;
  CUDA_TRY(cudaGetLastError());
  return RTC_OK;
}

// RTC_WAVEFRONT=0/1 in the environment overrides the choice (experiments); otherwise RTC_OPT_WAVEFRONT /
// RTC_OPT_RECURSIVE in rtc_opts.flags, otherwise the measured default.
// Default (measured, profiles/r01h_shard_probe.log, r01k_ncu_teapot_summary.md): on a whole frame the recursive kernel
// is faster (no queue traffic: 2.06 vs 2.62 ms on the glass scene); once the frame is split over >= RTK_WAVE_MIN_RANKS
// GPUs the serial recursion of a pixel whose hits spawn TWO children (a material that reflects and refracts: up to
// 63 calls at depth 5) is the critical path — 0.81 ms for an eighth of that frame, 0.48 ms as a wavefront. Mirror-only or
// glass-only scenes recurse in chains of at most max_depth + 1 calls; there a single heavy BVH query bounds the
// frame and the wavefront only adds traffic (teapot scene, an eighth: 1.64 vs 2.29 ms), so they stay recursive.
bool use_wavefront(const rtc_scene* s, const RenderParams& P, bool faithful, uint32_t flags) {
  if (faithful || P.aa != 1 || P.average || P.max_depth < 1) return false;
  static const char* env = std::getenv("RTC_WAVEFRONT");
  if (env && env[0] == '0') return false;
  if (env && env[0] == '1') return true;
  if (flags & RTC_OPT_RECURSIVE) return false;
  if (flags & RTC_OPT_WAVEFRONT) return true;
  // ... and on accelerators far beyond L2 the wavefront wins even on a whole frame (1 M triangles at 4K: 55.1 vs
  // 62.3 ms): its passes are all-reflection / all-refraction warps of compacted rays, which traverse together
  // (mirror-only scenes — the OBJ stand-in — recurse in chains and never take the wavefront: a shard of them is one
  // silhouette warp's chain either way; forced through it, one GPU as rank 0 of 4 / 8 needs 1.66 / 1.55 ms with the
  // passes and 1.34 / 1.28 ms streaming against 1.09 / 1.06 ms recursive, profiles/r02bs_shard_probe_others.log)
  return s->has_branching && (P.nranks >= RTK_WAVE_MIN_RANKS || s->big_accel);
}

int launch_frame(rtc_scene* s, const CameraD& cam, const RenderParams& P, bool faithful, double* out, cudaStream_t stream) {
  // the counters, the wavefront arena and the ray-count events are per scene: a frame still in flight on ANOTHER
  // stream must be over before they are reset (same stream: stream order does it)
  if (s->busy && s->busy_stream != stream) CUDA_TRY(cudaStreamSynchronize(s->busy_stream));
  s->busy = true;
  s->busy_stream = stream;
  // counters[7] is sticky: wavefront frames whose queues overflowed since the last rtc_scene_finish
  CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, (N_COUNTERS - 1) * sizeof(unsigned long long), stream));
  CUDA_TRY(cudaEventRecord(s->ev0, stream));
  s->last_wave = use_wavefront(s, P, faithful, P.flags);
  s->stat_flags = s->last_wave ? RTC_STAT_WAVEFRONT : 0u;
  int rc;
  if (s->last_wave) {
    s->last_cam = cam; s->last_P = P; s->last_out = out;
    // the level-synchronous passes (rt_wave.cuh), or one barrier-free persistent launch (rt_stream.cuh): on request,
    // and by default for half or less of a frame over an accelerator far beyond L2 — there a level's tail is
    // long (a few heavy traversals) and the next level's rays are ready to fill it (one GPU rendering rank 0's share
    // of the million-triangle 4K frame, profiles/r02bo_shard_probe_mesh.log: a quarter 13.5 -> 12.0 ms, an eighth
    // 8.9 -> 7.2 ms; half a frame 22.3 -> 21.9 (two GPUs: 22.7 -> 22.1), the whole frame 39.7 -> 40.7; on the glass scene the passes win at
    // every share). A frame it gives up on is rendered again recursively like an overflowed one.
    static const char* env_stream = std::getenv("RTC_WAVE_STREAM");
    const bool streaming = env_stream ? env_stream[0] == '1'
                                      : ((P.flags & RTC_OPT_WAVE_STREAM) != 0 || (s->big_accel && P.nranks >= RTK_STREAM_MIN_RANKS) ||
                                         // (a BVH of generic shapes, cover.yml, an eighth of the frame: 0.65 ms against 0.72 / 0.79)
                                         (!s->lean && !s->has_csg && !s->tri_sph_only && P.nranks >= 8));
    rc = streaming ? launch_stream(s, cam, P, out, stream) : launch_wave(s, cam, P, out, stream);
  } else {
    rc = launch_rows(s, cam, P, faithful, out, stream, 0, 0);
  }
  if (rc) return rc;
  CUDA_TRY(cudaEventRecord(s->ev1, stream));
  return RTC_OK;
}

// A wavefront frame whose queues were too small (more than 2 calls per pixel at some level) is rendered
// again by the recursive kernel, which needs no queues. Must run before anything reads the canvas.
int resolve_wave(rtc_scene* s, cudaStream_t stream) {
  if (!s->last_wave) return RTC_OK;
  s->last_wave = false;
  unsigned long long dropped = 0;
  CUDA_TRY(cudaMemcpyAsync(&dropped, s->d_counters + 6, sizeof dropped, cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(cudaStreamSynchronize(stream));
  if (!dropped) return RTC_OK;
  if (std::getenv("RTC_TIMING"))
    std::fprintf(stderr, "wavefront abandoned: %llu full-queue events, %llu colours that never arrived; rendering the frame recursively\n",
                 dropped & 0xffffffffull, dropped >> 32);
  s->stat_flags |= RTC_STAT_WAVE_FALLBACK;
  s->wave_clean_px = 0;  // the abandoned frame left colours in child slots
  CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, (N_COUNTERS - 1) * sizeof(unsigned long long), stream));
  CUDA_TRY(cudaEventRecord(s->ev0, stream));
  int rc = launch_rows(s, s->last_cam, s->last_P, false, s->last_out, stream, 0, 0);
  if (rc) return rc;
  CUDA_TRY(cudaEventRecord(s->ev1, stream));
  return RTC_OK;
}

// Device canvases are expensive to allocate (49.8 MB at 1080p): keep one freed canvas per device.
struct CanvasPool {
  std::mutex mu;
  double* ptr[16] = {nullptr};
  size_t doubles[16] = {0};
  double* take(int dev, size_t need, size_t& got) {
    std::lock_guard<std::mutex> g(mu);
    if (dev >= 0 && dev < 16 && ptr[dev] && doubles[dev] >= need) {
      double* p = ptr[dev];
      got = doubles[dev];
      ptr[dev] = nullptr;
      doubles[dev] = 0;
      return p;
    }
    return nullptr;
  }
  void give(int dev, double* p, size_t n) {
    {
      std::lock_guard<std::mutex> g(mu);
      if (dev >= 0 && dev < 16 && doubles[dev] < n) {
        std::swap(ptr[dev], p);
        std::swap(doubles[dev], n);
      }
    }
    if (p) cudaFree(p);
  }
};
CanvasPool g_canvas_pool;

// Streams and events of destroyed scenes, per device: creating a scene per frame (the end-to-end path of a caller
// that hands over a fresh World every time) then costs no cudaStreamCreate / cudaEventCreate (≈ 60 µs together).
struct StreamPool {
  struct Item { cudaStream_t stream; cudaEvent_t ev0, ev1; };
  std::mutex mu;
  std::vector<Item> free_[16];
  bool take(int dev, Item& it) {
    std::lock_guard<std::mutex> g(mu);
    if (dev < 0 || dev >= 16 || free_[dev].empty()) return false;
    it = free_[dev].back();
    free_[dev].pop_back();
    return true;
  }
  bool give(int dev, const Item& it) {
    std::lock_guard<std::mutex> g(mu);
    if (dev < 0 || dev >= 16 || free_[dev].size() >= 8) return false;
    free_[dev].push_back(it);
    return true;
  }
};
StreamPool g_stream_pool;

// rtc_trace_rays_static: device staging for a ray batch without any allocation. One file-scope array per GPU that
// loads the library (320 KB, held until exit): [origins 3n | directions 3n | rgb 3n | t n] at RTC_TRACE_STATIC_MAX.
// In host code `g_trace_scratch` is only the symbol's host-side stand-in; the device address is looked up once per
// device with that device current (cudaGetSymbolAddress) and cached here.
static __device__ __align__(256) double g_trace_scratch[(size_t)RTC_TRACE_STATIC_MAX * 10];
struct TraceScratchTable {
  std::mutex mu;
  double* ptr[64] = {nullptr};
  int get(int device, double** out) {
    if (device < 0 || device >= 64) return fail(RTC_ERR_INVALID, "device index beyond the static scratch table (64)");
    std::lock_guard<std::mutex> g(mu);
    if (!ptr[device]) {
      int cur = -1;
      CUDA_TRY(cudaGetDevice(&cur));
      if (cur != device) return fail(RTC_ERR_CUDA, "static scratch set-up needs its device current");
      double* p = nullptr;
      CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_trace_scratch));
      if (!p) return fail(RTC_ERR_CUDA, "cudaGetSymbolAddress(g_trace_scratch) returned NULL");
      ptr[device] = p;
    }
    *out = ptr[device];
    return RTC_OK;
  }
};
TraceScratchTable g_trace_scratch_table;

}  // namespace

extern "C" {

const char* rtc_last_error(void) { return g_err.c_str(); }
int rtc_abi_version(void) { return RTC_ABI_VERSION; }
int rtc_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int rtc_scene_create(const rtc_scene_desc* d, int device, rtc_scene** out) {
  if (!out) return fail(RTC_ERR_INVALID, "out is NULL");
  *out = nullptr;
  std::vector<int32_t> csg_parent;
  bool has_csg = false;
  const auto t_begin = std::chrono::steady_clock::now();
  int rc = validate(d, csg_parent, has_csg);
  if (rc) return rc;
  const auto t_valid = std::chrono::steady_clock::now();
  if (rtc_device_count() <= device || device < 0) return fail(RTC_ERR_CUDA, "no such CUDA device (this library has no CPU fallback)");
  CUDA_TRY(cudaSetDevice(device));

  AccelBuild accel;
  {
    // RTC_DEVICE_BVH_MIN = n: from n BVH prims on, the host only walks the nodes and the tree is built on the GPU
    // (rt_accel_device.cuh). Off by default: built in milliseconds instead of 0.4 s for a million triangles, but its
    // Morton-contiguous nodes overlap more than the host's SAH tree — twice the node visits per ray on both mesh scenes
    // (profiles/r02r_leaf_stats.log), so a scene that is rendered more than a few times is better off with the host tree.
    const char* env = std::getenv("RTC_DEVICE_BVH_MIN");
    const size_t device_min = env ? (size_t)std::atoll(env) : (size_t)0;
    build_accel(d, accel, g_host_build_only ? 0 : device_min);
  }
  const size_t n_dev = accel.deferred ? accel.cand.size() : 0;
  const size_t n_bvh_nodes = accel.deferred ? devbuild::node_count((int)n_dev) : accel.trav_count();
  const size_t n_leaf_recs = accel.deferred ? n_dev : accel.leaf_recs.size();
  const auto t_accel = std::chrono::steady_clock::now();
  // one slab: [nodes | xforms | boxes | limits | tri_geom | tri_normals | materials | patterns | lights | csg_parent
  //            | parent | prims | bvh | recs | leaf_recs]
  struct Part { const void* src; size_t bytes; size_t off; };
  Part parts[15] = {
      {d->nodes, sizeof(rtc_node) * d->n_nodes, 0},
      {d->xforms, sizeof(double) * 12 * d->n_xforms, 0},
      {d->boxes, sizeof(double) * 6 * d->n_boxes, 0},
      {d->limits, sizeof(double) * 2 * d->n_limits, 0},
      {d->tri_p1e1e2, sizeof(double) * 9 * d->n_triangles, 0},
      {d->tri_normals, sizeof(double) * 9 * d->n_triangles, 0},
      {d->materials, sizeof(rtc_material) * d->n_materials, 0},
      {d->patterns, sizeof(rtc_pattern) * d->n_patterns, 0},
      {d->lights, sizeof(rtc_light) * d->n_lights, 0},
      {csg_parent.data(), sizeof(int32_t) * csg_parent.size(), 0},
      {accel.parent.data(), sizeof(int32_t) * accel.parent.size(), 0},
      {accel.prims.data(), sizeof(int32_t) * accel.prims.size(), 0},
      {accel.deferred ? nullptr : accel.trav_data(), sizeof(TravNode) * n_bvh_nodes, 0},
      {accel.recs.data(), sizeof(PrimRec) * accel.recs.size(), 0},
      {accel.deferred ? nullptr : accel.leaf_recs.data(), sizeof(LeafRec) * n_leaf_recs, 0}};
  size_t total = 0;
  for (int k = 0; k < 15; k++) {
    Part& p = parts[k];
    if (p.bytes && !p.src && !(accel.deferred && (k == 12 || k == 14))) return fail(RTC_ERR_INVALID, "a non-empty table has a NULL pointer");
    p.off = total;
    total += align_up(p.bytes);
  }
  if (total == 0) total = 256;
  rtc_scene* s = new rtc_scene();
  s->device = device;
  s->has_csg = has_csg;
  for (uint32_t m = 0; m < d->n_materials; m++)
    if (d->materials[m].reflective != 0.0 && d->materials[m].transparency != 0.0) s->has_branching = true;
  // 126 MB of L2 on B200: past ~64 MB of nodes + leaf records, traversal is DRAM-latency bound
  s->big_accel = sizeof(TravNode) * n_bvh_nodes + sizeof(LeafRec) * n_leaf_recs > ((size_t)64 << 20);
  if (const char* env = std::getenv("RTC_BIG_ACCEL")) s->big_accel = env[0] == '1';  // experiments: force the 32-warp kernels on/off
  s->slab_bytes = total;
  auto bail = [&](cudaError_t e, const char* what) {
    std::string m = std::string(what) + ": " + cudaGetErrorString(e);
    rtc_scene_destroy(s);
    return fail(RTC_ERR_CUDA, m);
  };
  cudaError_t e;
  // Stream-ordered allocation from the device's default pool (kept warm: release threshold = max),
  // so creating and destroying a scene per frame costs no cudaMalloc/cudaFree round trip.
  {
    static std::mutex mu;
    static bool pool_ready[64] = {false};
    std::lock_guard<std::mutex> g(mu);
    if (device < 64 && !pool_ready[device]) {
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        unsigned long long thr = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
      }
      cudaGetLastError();
      pool_ready[device] = true;
    }
  }
  {
    StreamPool::Item it;
    if (g_stream_pool.take(device, it)) { s->stream = it.stream; s->ev0 = it.ev0; s->ev1 = it.ev1; }
  }
  if (!s->stream && (e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
  { int n = 0; if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && n > 0) s->sm_count = n; }
  const size_t tail = align_up(N_COUNTERS_ALLOC * sizeof(unsigned long long)) + align_up(32 * sizeof(double));
  if ((e = cudaMallocAsync((void**)&s->slab, total + tail, s->stream)) != cudaSuccess) return bail(e, "cudaMallocAsync(scene)");
  s->d_counters = (unsigned long long*)(s->slab + total);
  s->d_probe = (double*)(s->slab + total + align_up(N_COUNTERS_ALLOC * sizeof(unsigned long long)));
  if (!s->ev0 && (e = cudaEventCreate(&s->ev0)) != cudaSuccess) return bail(e, "cudaEventCreate");
  if (!s->ev1 && (e = cudaEventCreate(&s->ev1)) != cudaSuccess) return bail(e, "cudaEventCreate");
  const auto t_built = std::chrono::steady_clock::now();
  if (total <= ((size_t)1 << 20)) {
    // small scene: stage on the host, one H2D copy (15 separate copies would cost more than the bytes)
    std::vector<char> stage(total, 0);
    for (auto& p : parts)
      if (p.bytes && p.src) std::memcpy(stage.data() + p.off, p.src, p.bytes);
    if ((e = cudaMemcpyAsync(s->slab, stage.data(), total, cudaMemcpyHostToDevice, s->stream)) != cudaSuccess) return bail(e, "cudaMemcpy(scene)");
    if ((e = cudaStreamSynchronize(s->stream)) != cudaSuccess) return bail(e, "cudaStreamSynchronize");
  } else {
    // large scene (meshes): each table straight from where it lies, no host-side gather of hundreds of MB
    for (auto& p : parts)
      if (p.bytes && p.src && (e = cudaMemcpyAsync(s->slab + p.off, p.src, p.bytes, cudaMemcpyHostToDevice, s->stream)) != cudaSuccess)
        return bail(e, "cudaMemcpy(scene table)");
    if ((e = cudaStreamSynchronize(s->stream)) != cudaSuccess) return bail(e, "cudaStreamSynchronize");
  }
  const auto t_uploaded = std::chrono::steady_clock::now();
  if (accel.deferred) {
    // the tree and the leaf records, on the GPU, from the tables that have just arrived
    char* tmp = nullptr;
    const size_t cand_bytes = align_up(sizeof(int32_t) * n_dev), fwd_bytes = align_up(sizeof(double) * accel.fwd.size());
    if ((e = cudaMallocAsync((void**)&tmp, cand_bytes + fwd_bytes, s->stream)) != cudaSuccess) return bail(e, "cudaMallocAsync(device build)");
    devbuild::In in;
    in.nodes = (const rtc_node*)(s->slab + parts[0].off);
    in.limits = (const double*)(s->slab + parts[3].off);
    in.tri_geom = (const double*)(s->slab + parts[4].off);
    in.parent = (const int32_t*)(s->slab + parts[10].off);
    in.cand = (const int32_t*)tmp;
    in.fwd = (const double*)(tmp + cand_bytes);
    in.n = (int32_t)n_dev;
    bool ok = false;
    int depth = 0;
    e = cudaMemcpyAsync(tmp, accel.cand.data(), sizeof(int32_t) * n_dev, cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(tmp + cand_bytes, accel.fwd.data(), sizeof(double) * accel.fwd.size(), cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess)
      e = devbuild::build(in, accel.n_always, (TravNode*)(s->slab + parts[12].off), (LeafRec*)(s->slab + parts[14].off), s->stream, &ok, &depth);
    cudaFreeAsync(tmp, s->stream);
    if (e != cudaSuccess) return bail(e, "device BVH build");
    if (std::getenv("RTC_TIMING"))
      std::fprintf(stderr, "device BVH build: %zu prims, %zu nodes, depth %d, %.2f ms%s\n", n_dev, n_bvh_nodes, depth,
                   std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_uploaded).count(),
                   ok ? "" : " — declined (non-finite box or tree too deep): host builder");
    if (!ok) {
      rtc_scene_destroy(s);
      g_host_build_only = true;
      const int rc2 = rtc_scene_create(d, device, out);
      g_host_build_only = false;
      return rc2;
    }
  }
  if (std::getenv("RTC_TIMING")) {
    auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
      return std::chrono::duration<double, std::milli>(b - a).count();
    };
    std::fprintf(stderr, "rtc_scene_create: validate %.2f ms, build_accel %.2f ms, alloc %.2f ms, upload %.2f ms (%.1f MB)\n",
                 ms(t_begin, t_valid), ms(t_valid, t_accel), ms(t_accel, t_built), ms(t_built, std::chrono::steady_clock::now()),
                 total / 1048576.0);
  }
  SceneView& v = s->view;
  v.nodes = (const rtc_node*)(s->slab + parts[0].off);
  v.xforms = (const double*)(s->slab + parts[1].off);
  v.boxes = (const double*)(s->slab + parts[2].off);
  v.limits = (const double*)(s->slab + parts[3].off);
  v.tri_geom = (const double*)(s->slab + parts[4].off);
  v.tri_normals = (const double*)(s->slab + parts[5].off);
  v.materials = (const rtc_material*)(s->slab + parts[6].off);
  v.patterns = (const rtc_pattern*)(s->slab + parts[7].off);
  v.lights = (const rtc_light*)(s->slab + parts[8].off);
  v.csg_parent = (const int32_t*)(s->slab + parts[9].off);
  v.parent = (const int32_t*)(s->slab + parts[10].off);
  v.prims = (const int32_t*)(s->slab + parts[11].off);
  v.bvh = (const TravNode*)(s->slab + parts[12].off);
  v.recs = (const PrimRec*)(s->slab + parts[13].off);
  v.leaf_recs = (const LeafRec*)(s->slab + parts[14].off);
  v.n_always = accel.n_always;
  v.n_sph = accel.n_sph;
  v.n_pl = accel.n_pl;
  for (int k = 0; k < SPH_CULL_MAX; k++) v.sph_cull[k] = accel.sph_cull[k];
  for (int k = 0; k < CUBE_DIR_MAX; k++) v.cube_dirs[k] = accel.cube_dirs[k];
  v.n_cube_dirs = accel.n_cube_dirs;
  v.bvh_ext[0] = v.bvh_ext[1] = v.bvh_ext[2] = 0.0;
  if (n_bvh_nodes > 0) {
    if (accel.deferred) {  // built on the device: fetch the root (the build has completed on s->stream)
      TravNode root;
      cudaError_t e = cudaMemcpyAsync(&root, s->slab + parts[12].off, sizeof root, cudaMemcpyDeviceToHost, s->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
      if (e != cudaSuccess) return bail(e, "device BVH root");
      bvh_extent(root, v.bvh_ext);
    } else {
      bvh_extent(accel.bvh[0], v.bvh_ext);
    }
  }
  v.n_bvh = (int32_t)n_bvh_nodes;
  v.n_nodes = (int32_t)d->n_nodes;
  v.n_lights = (int32_t)d->n_lights;
  v.has_csg = has_csg ? 1 : 0;
  s->lean = !has_csg && n_bvh_nodes == 0 && accel.n_always == accel.n_sph + accel.n_pl;
  s->tri_sph_only = !has_csg && n_bvh_nodes != 0 && accel.n_always == accel.n_sph + accel.n_pl && !(accel.deferred && accel.cand_generic);
  for (const LeafRec& lr : accel.leaf_recs) {
    const uint32_t k = lr.kind_flags >> 24;
    if (k != RTC_SPHERE && k != RTC_TRIANGLE && k != RTC_SMOOTH_TRIANGLE) { s->tri_sph_only = false; break; }
  }
  v.sane_indices = 1;
  for (uint32_t m = 0; m < d->n_materials; m++)
    if (!(d->materials[m].refractive_index > 0.0) || !std::isfinite(d->materials[m].refractive_index)) v.sane_indices = 0;
  *out = s;
  return RTC_OK;
}

void rtc_scene_destroy(rtc_scene* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  // a frame queued on a caller's stream may still be reading the slab, the arena and the canvas that go back to
  // their pools below
  if (s->busy) cudaStreamSynchronize(s->busy_stream);
  if (s->slab) {
    if (s->stream) cudaFreeAsync(s->slab, s->stream); else cudaFree(s->slab);
  }
  if (s->d_canvas) g_canvas_pool.give(s->device, s->d_canvas, s->canvas_doubles);
  if (s->wave_arena) g_arena_pool.give(s->device, s->wave_arena, s->wave_bytes, s->wave_clean_px, s->wave_clean_depth);
  // (the slab's cudaFreeAsync above is ordered on the stream; whoever takes the stream next allocates behind it)
  if (!(s->stream && s->ev0 && s->ev1 && g_stream_pool.give(s->device, StreamPool::Item{s->stream, s->ev0, s->ev1}))) {
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->stream) cudaStreamDestroy(s->stream);
  }
  delete s;
}

int rtc_render_device(rtc_scene* s, const rtc_camera* camera, const rtc_opts* opts, double* rgb_dev, void* stream_, rtc_stats* stats) {
  if (!s || !rgb_dev) return fail(RTC_ERR_INVALID, "NULL scene or canvas");
  auto t0 = std::chrono::steady_clock::now();
  CameraD cam;
  RenderParams P{};
  bool faithful = false;
  int rc = make_camera(camera, cam);
  if (rc) return rc;
  rc = make_params(camera, opts, P, faithful);
  if (rc) return rc;
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t stream = (cudaStream_t)stream_;
  s->launches = 0;
  rc = launch_frame(s, cam, P, faithful, rgb_dev, stream);
  if (rc) return rc;
  if (stats) {
    rc = rtc_scene_finish(s, stream_, stats);
    stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  return rc;
}

int rtc_scene_finish(rtc_scene* s, void* stream_, rtc_stats* stats) {
  if (!s) return fail(RTC_ERR_INVALID, "NULL scene");
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t stream = (cudaStream_t)stream_;
  if (stats) std::memset(stats, 0, sizeof *stats);
  int rc = resolve_wave(s, stream);
  if (rc) return rc;
  rc = read_counters(s, stream, stats);  // synchronises `stream`
  if (s->busy && s->busy_stream == stream) s->busy = false;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(RTC_ERR_CUDA, std::string("kernel: ") + cudaGetErrorString(e));
  if (stats) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, s->ev0, s->ev1) == cudaSuccess) stats->kernel_ms = ms; else cudaGetLastError();
    stats->launches = s->launches;
    stats->flags = s->stat_flags;
  }
  return rc;
}

// Device canvas of at least `need` doubles for this scene (recycled through the per-device pool).
static int ensure_canvas(rtc_scene* s, size_t need) {
  if (s->canvas_doubles >= need) return RTC_OK;
  if (s->d_canvas) g_canvas_pool.give(s->device, s->d_canvas, s->canvas_doubles);
  s->d_canvas = nullptr;
  s->canvas_doubles = 0;
  size_t got = 0;
  s->d_canvas = g_canvas_pool.take(s->device, need, got);
  if (s->d_canvas) {
    s->canvas_doubles = got;
  } else {
    CUDA_TRY(cudaMalloc((void**)&s->d_canvas, need * sizeof(double)));
    s->canvas_doubles = need;
  }
  return RTC_OK;
}

int rtc_render(rtc_scene* s, const rtc_camera* camera, const rtc_opts* opts, double* rgb_host, rtc_stats* stats) {
  if (!s || !rgb_host) return fail(RTC_ERR_INVALID, "NULL scene or canvas");
  auto t0 = std::chrono::steady_clock::now();
  CameraD cam;
  RenderParams P{};
  bool faithful = false;
  int rc = make_camera(camera, cam);
  if (rc) return rc;
  rc = make_params(camera, opts, P, faithful);
  if (rc) return rc;
  CUDA_TRY(cudaSetDevice(s->device));
  size_t need = (size_t)cam.hsize * cam.vsize * 3;
  {
    // Zero-copy canvas: if the caller's canvas is pinned, mapped host memory (cudaHostAlloc /
    // cudaHostRegister — what a pinned torch tensor is), the kernel stores its pixels straight into
    // it over PCIe (24 B per pixel, posted writes spread over the whole kernel) and there is no
    // device canvas and no device→host copy at all.
    cudaPointerAttributes attr;
    // (The wavefront renderer writes a pixel from whichever pass completes it; measured over PCIe that is
    // 9.8 ms against 3.9 ms into a device canvas, so it takes the staged path: device canvas, one DMA.)
    if (!use_wavefront(s, P, faithful, P.flags) && cudaPointerGetAttributes(&attr, rgb_host) == cudaSuccess &&
        attr.type == cudaMemoryTypeHost && attr.devicePointer) {
      s->launches = 0;
      rc = launch_frame(s, cam, P, faithful, (double*)attr.devicePointer, s->stream);
      if (rc) return rc;
      rtc_stats local;
      rc = rtc_scene_finish(s, s->stream, &local);
      local.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      if (stats) *stats = local;
      return rc;
    }
    cudaGetLastError();  // pageable memory: not an error, take the staged path
  }
  rc = ensure_canvas(s, need);
  if (rc) return rc;
  s->launches = 0;
  const size_t row_doubles = (size_t)cam.hsize * 3;
  if (P.nranks <= 1) {
    rc = launch_frame(s, cam, P, faithful, s->d_canvas, s->stream);
    if (rc) return rc;
    rc = resolve_wave(s, s->stream);  // before the canvas is read
    if (rc) return rc;
    size_t off = (size_t)P.row_begin * row_doubles, cnt = (size_t)(P.row_end - P.row_begin) * row_doubles;
    if (cnt) CUDA_TRY(cudaMemcpyAsync(rgb_host + off, s->d_canvas + off, cnt * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  } else {
    rc = launch_frame(s, cam, P, faithful, s->d_canvas, s->stream);
    if (rc) return rc;
    rc = resolve_wave(s, s->stream);  // before the canvas is read
    if (rc) return rc;
    // device → host: only the rows this call rendered (owned tiles clipped to the row range)
    if (P.row_begin == 0 && P.row_end == cam.vsize) {
      // whole frame: this rank's whole tiles leave as ONE strided DMA (plus the ragged last tile if it is ours)
      const size_t row_bytes = row_doubles * sizeof(double), tile_bytes = row_bytes * P.tile_rows;
      const uint32_t full_tiles = cam.vsize / P.tile_rows;
      const uint32_t mine = full_tiles > P.rank ? (full_tiles - P.rank + P.nranks - 1) / P.nranks : 0;
      const size_t first = (size_t)P.rank * tile_bytes;
      if (mine)
        CUDA_TRY(cudaMemcpy2DAsync((char*)rgb_host + first, tile_bytes * P.nranks, (const char*)s->d_canvas + first,
                                   tile_bytes * P.nranks, tile_bytes, mine, cudaMemcpyDeviceToHost, s->stream));
      if (cam.vsize % P.tile_rows && full_tiles % P.nranks == P.rank) {
        const size_t off = (size_t)full_tiles * tile_bytes;
        CUDA_TRY(cudaMemcpyAsync((char*)rgb_host + off, (const char*)s->d_canvas + off, row_bytes * (cam.vsize % P.tile_rows),
                                 cudaMemcpyDeviceToHost, s->stream));
      }
    } else {
      for (uint32_t t = P.rank; (size_t)t * P.tile_rows < cam.vsize; t += P.nranks) {
        uint32_t y0 = t * P.tile_rows, y1 = y0 + P.tile_rows;
        if (y0 < P.row_begin) y0 = P.row_begin;
        if (y1 > P.row_end) y1 = P.row_end;
        if (y0 >= y1) continue;
        size_t off = (size_t)y0 * row_doubles, cnt = (size_t)(y1 - y0) * row_doubles;
        CUDA_TRY(cudaMemcpyAsync(rgb_host + off, s->d_canvas + off, cnt * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
      }
    }
  }
  static const bool timing = std::getenv("RTC_TIMING") != nullptr;
  const auto t_queued = std::chrono::steady_clock::now();
  rtc_stats local;
  rc = rtc_scene_finish(s, s->stream, &local);
  local.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  if (timing)
    std::fprintf(stderr, "rtc_render(staged): queue + resolve %.3f ms, kernels %.3f ms, whole call %.3f ms (rank %u of %u)\n",
                 std::chrono::duration<double, std::milli>(t_queued - t0).count(), local.kernel_ms, local.total_ms, P.rank, P.nranks);
  if (stats) *stats = local;
  return rc;
}

int rtc_quantize_device(int device, const double* rgb_dev, size_t n, uint8_t* rgb8_dev, void* stream) {
  if (!rgb_dev || !rgb8_dev) return fail(RTC_ERR_INVALID, "NULL buffer");
  CUDA_TRY(cudaSetDevice(device));
  if (n == 0) return RTC_OK;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  quantize_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(rgb_dev, n, rgb8_dev);
  CUDA_TRY(cudaGetLastError());
  return RTC_OK;
}

// File-output path: render, quantise on the device, and move 3 B per pixel instead of 24.
int rtc_render_rgb8(rtc_scene* s, const rtc_camera* camera, const rtc_opts* opts, uint8_t* rgb8_host, rtc_stats* stats) {
  if (!s || !rgb8_host) return fail(RTC_ERR_INVALID, "NULL scene or image");
  auto t0 = std::chrono::steady_clock::now();
  CameraD cam;
  RenderParams P{};
  bool faithful = false;
  int rc = make_camera(camera, cam);
  if (rc) return rc;
  rc = make_params(camera, opts, P, faithful);
  if (rc) return rc;
  if (P.nranks > 1) return fail(RTC_ERR_INVALID, "rtc_render_rgb8 renders whole row ranges (nranks must be 1)");
  CUDA_TRY(cudaSetDevice(s->device));
  const size_t need = (size_t)cam.hsize * cam.vsize * 3;
  rc = ensure_canvas(s, need);
  if (rc) return rc;
  s->launches = 0;
  rc = launch_frame(s, cam, P, faithful, s->d_canvas, s->stream);
  if (rc) return rc;
  rc = resolve_wave(s, s->stream);  // before the canvas is read
  if (rc) return rc;
  const size_t off = (size_t)P.row_begin * cam.hsize * 3, cnt = (size_t)(P.row_end - P.row_begin) * cam.hsize * 3;
  if (cnt) {
    uint8_t* d8 = nullptr;
    CUDA_TRY(cudaMallocAsync((void**)&d8, cnt, s->stream));
    rc = rtc_quantize_device(s->device, s->d_canvas + off, cnt, d8, s->stream);
    if (rc == RTC_OK) {
      s->launches++;
      cudaError_t e = cudaMemcpyAsync(rgb8_host + off, d8, cnt, cudaMemcpyDeviceToHost, s->stream);
      if (e != cudaSuccess) rc = fail(RTC_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaFreeAsync(d8, s->stream);
    if (rc) return rc;
  }
  rtc_stats local;
  rc = rtc_scene_finish(s, s->stream, &local);
  local.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  if (stats) *stats = local;
  return rc;
}

int rtc_trace_rays(rtc_scene* s, size_t n, const double* origins, const double* directions, uint32_t remaining,
                   double* rgb_out, double* t_out) {
  if (!s || !origins || !directions) return fail(RTC_ERR_INVALID, "NULL argument");
  if ((int)remaining + 1 > MAX_FRAMES) return fail(RTC_ERR_UNSUPPORTED, "remaining > 7");
  if (n == 0) return RTC_OK;
  CUDA_TRY(cudaSetDevice(s->device));
  double *d_o = nullptr, *d_d = nullptr, *d_rgb = nullptr, *d_t = nullptr;
  auto cleanup = [&]() { cudaFree(d_o); cudaFree(d_d); cudaFree(d_rgb); cudaFree(d_t); };
  cudaError_t e;
#define TRY_OR_CLEAN(expr) if ((e = (expr)) != cudaSuccess) { cleanup(); return fail(RTC_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e)); }
  TRY_OR_CLEAN(cudaMalloc((void**)&d_o, 3 * n * sizeof(double)));
  TRY_OR_CLEAN(cudaMalloc((void**)&d_d, 3 * n * sizeof(double)));
  TRY_OR_CLEAN(cudaMalloc((void**)&d_rgb, 3 * n * sizeof(double)));
  TRY_OR_CLEAN(cudaMalloc((void**)&d_t, n * sizeof(double)));
  TRY_OR_CLEAN(cudaMemcpyAsync(d_o, origins, 3 * n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  TRY_OR_CLEAN(cudaMemcpyAsync(d_d, directions, 3 * n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  TRY_OR_CLEAN(cudaMemsetAsync(s->d_counters, 0, N_COUNTERS * sizeof(unsigned long long), s->stream));
  unsigned blocks = (unsigned)((n + 127) / 128);
  if (s->has_csg) trace_kernel<true><<<blocks, 128, 0, s->stream>>>(s->view, n, d_o, d_d, (int)remaining, d_rgb, d_t, s->d_counters);
  else trace_kernel<false><<<blocks, 128, 0, s->stream>>>(s->view, n, d_o, d_d, (int)remaining, d_rgb, d_t, s->d_counters);
  TRY_OR_CLEAN(cudaGetLastError());
  if (rgb_out) TRY_OR_CLEAN(cudaMemcpyAsync(rgb_out, d_rgb, 3 * n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  if (t_out) TRY_OR_CLEAN(cudaMemcpyAsync(t_out, d_t, n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  int rc = read_counters(s, s->stream, nullptr);
  cleanup();
#undef TRY_OR_CLEAN
  return rc;
}

// No counterpart in the reference: rtc_trace_rays for callers that must not allocate (CUDA graph capture around the
// call site, tight serving loops). Same kernel, same stream, same counters; the staging lives in g_trace_scratch.
// No cudaMalloc / cudaFree / allocating library call on this path. Every call shares the array: calls that overlap
// on one GPU race (this one is synchronous, so only concurrent host threads can make them overlap).
int rtc_trace_rays_static(rtc_scene* s, size_t n, const double* origins, const double* directions, uint32_t remaining,
                          double* rgb_out, double* t_out) {
  if (!s || !origins || !directions) return fail(RTC_ERR_INVALID, "NULL argument");
  if ((int)remaining + 1 > MAX_FRAMES) return fail(RTC_ERR_UNSUPPORTED, "remaining > 7");
  if (n > (size_t)RTC_TRACE_STATIC_MAX)
    return fail(RTC_ERR_INVALID, "rtc_trace_rays_static: n = " + std::to_string(n) + " exceeds RTC_TRACE_STATIC_MAX = " +
                                     std::to_string((int)RTC_TRACE_STATIC_MAX) + " rays (use rtc_trace_rays, or split the batch)");
  if (n == 0) return RTC_OK;
  CUDA_TRY(cudaSetDevice(s->device));
  double* base = nullptr;
  int rc = g_trace_scratch_table.get(s->device, &base);
  if (rc) return rc;
  const size_t cap = (size_t)RTC_TRACE_STATIC_MAX;
  double *d_o = base, *d_d = base + 3 * cap, *d_rgb = base + 6 * cap, *d_t = base + 9 * cap;
  CUDA_TRY(cudaMemcpyAsync(d_o, origins, 3 * n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  CUDA_TRY(cudaMemcpyAsync(d_d, directions, 3 * n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  // the array keeps the previous call's results: the kernel overwrites rgb/t of every ray < n, and the rest is
  // never copied out; cleared all the same so that a faulting launch cannot hand back stale colours
  CUDA_TRY(cudaMemsetAsync(d_rgb, 0, 4 * cap * sizeof(double), s->stream));
  CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, N_COUNTERS * sizeof(unsigned long long), s->stream));
  unsigned blocks = (unsigned)((n + 127) / 128);
  if (s->has_csg) trace_kernel<true><<<blocks, 128, 0, s->stream>>>(s->view, n, d_o, d_d, (int)remaining, d_rgb, d_t, s->d_counters);
  else trace_kernel<false><<<blocks, 128, 0, s->stream>>>(s->view, n, d_o, d_d, (int)remaining, d_rgb, d_t, s->d_counters);
  CUDA_TRY(cudaGetLastError());
  if (rgb_out) CUDA_TRY(cudaMemcpyAsync(rgb_out, d_rgb, 3 * n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  if (t_out) CUDA_TRY(cudaMemcpyAsync(t_out, d_t, n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  return read_counters(s, s->stream, nullptr);
}

int rtc_probe_hit(rtc_scene* s, const double* o, const double* dir, double* out20) {
  if (!s || !o || !dir || !out20) return fail(RTC_ERR_INVALID, "NULL argument");
  CUDA_TRY(cudaSetDevice(s->device));
  RayD r;
  r.o = v3(o[0], o[1], o[2]);
  r.d = v3(dir[0], dir[1], dir[2]);
  CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, N_COUNTERS * sizeof(unsigned long long), s->stream));
  if (s->has_csg) probe_kernel<true><<<1, 32, 0, s->stream>>>(s->view, r, s->d_probe, s->d_counters);
  else probe_kernel<false><<<1, 32, 0, s->stream>>>(s->view, r, s->d_probe, s->d_counters);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpyAsync(out20, s->d_probe, 20 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  return read_counters(s, s->stream, nullptr);
}

int rtc_is_shadowed(rtc_scene* s, const double* p, uint32_t light_index) {
  if (!s || !p) return fail(RTC_ERR_INVALID, "NULL argument");
  if ((int32_t)light_index >= s->view.n_lights) return fail(RTC_ERR_INVALID, "light index out of range");
  CUDA_TRY(cudaSetDevice(s->device));
  CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, N_COUNTERS * sizeof(unsigned long long), s->stream));
  int* d_out = reinterpret_cast<int*>(s->d_probe);
  if (s->has_csg) shadow_kernel<true><<<1, 32, 0, s->stream>>>(s->view, v3(p[0], p[1], p[2]), (int)light_index, d_out, s->d_counters);
  else shadow_kernel<false><<<1, 32, 0, s->stream>>>(s->view, v3(p[0], p[1], p[2]), (int)light_index, d_out, s->d_counters);
  CUDA_TRY(cudaGetLastError());
  int h = 0;
  CUDA_TRY(cudaMemcpyAsync(&h, d_out, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
  int rc = read_counters(s, s->stream, nullptr);
  return rc ? rc : h;
}

int rtc_canvas_create(int device, size_t n_doubles, rtc_canvas** out) {
  if (!out || n_doubles == 0) return fail(RTC_ERR_INVALID, "bad canvas size");
  CUDA_TRY(cudaSetDevice(device));
  rtc_canvas* c = new rtc_canvas();
  c->device = device;
  c->n = n_doubles;
  cudaError_t e = cudaMalloc((void**)&c->ptr, n_doubles * sizeof(double));
  if (e != cudaSuccess) { delete c; return fail(RTC_ERR_CUDA, std::string("cudaMalloc(canvas): ") + cudaGetErrorString(e)); }
  cudaMemset(c->ptr, 0, n_doubles * sizeof(double));
  *out = c;
  return RTC_OK;
}
int rtc_canvas_export(rtc_canvas* c, unsigned char handle[64]) {
  if (!c || c->opened) return fail(RTC_ERR_INVALID, "cannot export this canvas");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  CUDA_TRY(cudaSetDevice(c->device));
  CUDA_TRY(cudaIpcGetMemHandle(&h, c->ptr));
  std::memcpy(handle, &h, 64);
  return RTC_OK;
}
int rtc_canvas_open(int device, const unsigned char handle[64], size_t n_doubles, rtc_canvas** out) {
  if (!out) return fail(RTC_ERR_INVALID, "out is NULL");
  CUDA_TRY(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, 64);
  void* p = nullptr;
  CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  rtc_canvas* c = new rtc_canvas();
  c->device = device; c->ptr = (double*)p; c->n = n_doubles; c->opened = true;
  *out = c;
  return RTC_OK;
}
double* rtc_canvas_ptr(rtc_canvas* c) { return c ? c->ptr : nullptr; }
int rtc_canvas_download(rtc_canvas* c, double* host, size_t off, size_t n) {
  if (!c || !host || off + n > c->n) return fail(RTC_ERR_INVALID, "bad canvas range");
  CUDA_TRY(cudaSetDevice(c->device));
  CUDA_TRY(cudaMemcpy(host, c->ptr + off, n * sizeof(double), cudaMemcpyDeviceToHost));
  return RTC_OK;
}
int rtc_canvas_push_tiles(rtc_canvas* c, const double* local_dev, uint32_t hsize, uint32_t vsize, uint32_t tile_rows,
                          uint32_t rank, uint32_t nranks, void* stream_) {
  if (!c || !local_dev || tile_rows == 0 || nranks == 0 || rank >= nranks) return fail(RTC_ERR_INVALID, "bad tile push");
  if ((size_t)hsize * vsize * 3 > c->n) return fail(RTC_ERR_INVALID, "canvas too small for the frame");
  CUDA_TRY(cudaSetDevice(c->device));
  cudaStream_t stream = (cudaStream_t)stream_;
  const size_t row_bytes = (size_t)hsize * 3 * sizeof(double), tile_bytes = row_bytes * tile_rows;
  const uint32_t full_tiles = vsize / tile_rows;                       // tiles [0, full_tiles) are whole
  const uint32_t mine = full_tiles > rank ? (full_tiles - rank + nranks - 1) / nranks : 0;
  const size_t first = (size_t)rank * tile_bytes;
  if (mine)
    CUDA_TRY(cudaMemcpy2DAsync((char*)c->ptr + first, tile_bytes * nranks, (const char*)local_dev + first, tile_bytes * nranks,
                               tile_bytes, mine, cudaMemcpyDeviceToDevice, stream));
  if (vsize % tile_rows && full_tiles % nranks == rank) {              // the ragged last tile is this rank's
    const size_t off = (size_t)full_tiles * tile_bytes;
    CUDA_TRY(cudaMemcpyAsync((char*)c->ptr + off, (const char*)local_dev + off, row_bytes * (vsize % tile_rows),
                             cudaMemcpyDeviceToDevice, stream));
  }
  return RTC_OK;
}
void rtc_canvas_destroy(rtc_canvas* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->opened) cudaIpcCloseMemHandle(c->ptr); else cudaFree(c->ptr);
  delete c;
}

int rtc_host_canvas_create(const char* name, size_t n_doubles, int device, int create, rtc_host_canvas** out) {
  if (!out || !name || name[0] != '/' || n_doubles == 0) return fail(RTC_ERR_INVALID, "bad host canvas (name must start with '/')");
  *out = nullptr;
  CUDA_TRY(cudaSetDevice(device));
  const size_t bytes = ((n_doubles * sizeof(double) + 4095) & ~(size_t)4095) + HOST_CANVAS_CTRL;
  int fd = shm_open(name, create ? (O_CREAT | O_RDWR) : O_RDWR, 0600);
  if (fd < 0) return fail(RTC_ERR_INVALID, std::string("shm_open(") + name + "): " + std::strerror(errno));
  if (create && ftruncate(fd, (off_t)bytes) != 0) {
    const std::string m = std::string("ftruncate: ") + std::strerror(errno);
    close(fd);
    shm_unlink(name);
    return fail(RTC_ERR_INVALID, m);
  }
  void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) {
    if (create) shm_unlink(name);
    return fail(RTC_ERR_INVALID, std::string("mmap: ") + std::strerror(errno));
  }
  rtc_host_canvas* c = new rtc_host_canvas();
  c->name = name; c->ptr = (double*)p; c->bytes = bytes; c->device = device; c->creator = create != 0;
  *out = c;
  return RTC_OK;
}
// NUMA node of a CUDA device from sysfs (-1: unknown / no NUMA)
static int gpu_numa_node(int device) {
  char bdf[32] = {0};
  if (cudaDeviceGetPCIBusId(bdf, sizeof bdf, device) != cudaSuccess) { cudaGetLastError(); return -1; }
  for (char* q = bdf; *q; q++) *q = (char)std::tolower((unsigned char)*q);
  const std::string path = std::string("/sys/bus/pci/devices/") + bdf + "/numa_node";
  FILE* f = std::fopen(path.c_str(), "r");
  if (!f) return -1;
  int node = -1;
  if (std::fscanf(f, "%d", &node) != 1) node = -1;
  std::fclose(f);
  return node;
}
int rtc_host_canvas_place(rtc_host_canvas* c, uint32_t hsize, uint32_t vsize, uint32_t tile_rows, uint32_t rank, uint32_t nranks) {
  if (!c || c->registered || tile_rows == 0 || nranks == 0 || rank >= nranks) return fail(RTC_ERR_INVALID, "bad host canvas placement (before the first rtc_host_canvas_ptr only)");
  if ((size_t)hsize * vsize * 3 * sizeof(double) > c->bytes - HOST_CANVAS_CTRL) return fail(RTC_ERR_INVALID, "canvas too small for the frame");
  const int node = gpu_numa_node(c->device);
  const size_t row_bytes = (size_t)hsize * 3 * sizeof(double), page = 4096;
  char* base = reinterpret_cast<char*>(c->ptr);
  for (uint32_t t = rank; (size_t)t * tile_rows < vsize; t += nranks) {
    const size_t y0 = (size_t)t * tile_rows, y1 = std::min<size_t>(y0 + tile_rows, vsize);
    size_t b0 = y0 * row_bytes, b1 = y1 * row_bytes;
    if (node >= 0 && node < 64) {
      // whole pages inside the tile (a page shared with a neighbouring tile goes wherever it is touched first)
      const size_t p0 = (b0 + page - 1) & ~(page - 1), p1 = b1 & ~(page - 1);
      if (p1 > p0) {
        unsigned long mask = 1ul << node;
        (void)syscall(SYS_mbind, base + p0, p1 - p0, 2 /* MPOL_BIND */, &mask, 65ul, 0u);  // best effort
      }
    }
    std::memset(base + b0, 0, b1 - b0);  // first touch
  }
  if (rank == 0) std::memset(base + c->bytes - HOST_CANVAS_CTRL, 0, HOST_CANVAS_CTRL);
  return RTC_OK;
}
double* rtc_host_canvas_ptr(rtc_host_canvas* c) {
  if (!c) return nullptr;
  if (!c->registered) {
    if (cudaSetDevice(c->device) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    cudaError_t e = cudaHostRegister(c->ptr, c->bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e != cudaSuccess) {
      fail(RTC_ERR_CUDA, std::string("cudaHostRegister(host canvas): ") + cudaGetErrorString(e));
      return nullptr;
    }
    c->registered = true;
  }
  return c->ptr;
}
static volatile uint32_t* host_canvas_word(rtc_host_canvas* c, uint32_t rank) {
  return reinterpret_cast<volatile uint32_t*>(reinterpret_cast<char*>(c->ptr) + c->bytes - HOST_CANVAS_CTRL + 64 * (size_t)rank);
}
int rtc_host_canvas_arrive(rtc_host_canvas* c, uint32_t rank, uint32_t frame) {
  if (!c || rank >= 64) return fail(RTC_ERR_INVALID, "bad host canvas arrival");
  __atomic_store_n(host_canvas_word(c, rank), frame, __ATOMIC_RELEASE);
  return RTC_OK;
}
int rtc_host_canvas_wait(rtc_host_canvas* c, uint32_t nranks, uint32_t frame, double timeout_s) {
  if (!c || nranks == 0 || nranks > 64) return fail(RTC_ERR_INVALID, "bad host canvas wait");
  const auto t0 = std::chrono::steady_clock::now();
  for (uint32_t r = 0; r < nranks; r++) {
    uint32_t spins = 0;
    // (frames count up; a rank that is already past `frame` has arrived at it)
    while ((int32_t)(__atomic_load_n(host_canvas_word(c, r), __ATOMIC_ACQUIRE) - frame) < 0) {
      if ((++spins & 0x3ffu) == 0u && std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s)
        return fail(RTC_ERR_INVALID, "rtc_host_canvas_wait: rank " + std::to_string(r) + " did not arrive at frame " + std::to_string(frame));
    }
  }
  return RTC_OK;
}
void rtc_host_canvas_destroy(rtc_host_canvas* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->registered) cudaHostUnregister(c->ptr);
  if (c->ptr) munmap(c->ptr, c->bytes);
  if (c->creator) shm_unlink(c->name.c_str());
  delete c;
}

int rtc_debug_normalize(int device, size_t n, const double* xyz_host, double* out_host) {
  if (!xyz_host || !out_host) return fail(RTC_ERR_INVALID, "NULL buffer");
  CUDA_TRY(cudaSetDevice(device));
  if (n == 0) return RTC_OK;
  double* d = nullptr;
  CUDA_TRY(cudaMalloc((void**)&d, 2 * 3 * n * sizeof(double)));
  cudaError_t e = cudaMemcpy(d, xyz_host, 3 * n * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    normalize_kernel<<<(int)((n + 255) / 256 < 4096 ? (n + 255) / 256 : 4096), 256>>>(d, n, d + 3 * n);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(out_host, d + 3 * n, 3 * n * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return fail(RTC_ERR_CUDA, cudaGetErrorString(e));
  return RTC_OK;
}

int rtc_measure_fp64(int device, int fused, double* gflops_out) {
  if (!gflops_out) return fail(RTC_ERR_INVALID, "NULL out");
  CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  double* d = nullptr;
  CUDA_TRY(cudaMalloc((void**)&d, 64));
  cudaEvent_t a, b;
  CUDA_TRY(cudaEventCreate(&a));
  CUDA_TRY(cudaEventCreate(&b));
  const int iters = 8192, threads = 256, blocks = prop.multiProcessorCount * 8;
  double best = 0.0;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(a);
    if (fused) fp64_peak_kernel<true><<<blocks, threads>>>(d, iters, 1.0000001, 1e-9);
    else fp64_peak_kernel<false><<<blocks, threads>>>(d, iters, 1.0000001, 1e-9);
    cudaEventRecord(b);
    cudaError_t e = cudaEventSynchronize(b);
    if (e != cudaSuccess) { cudaFree(d); return fail(RTC_ERR_CUDA, cudaGetErrorString(e)); }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    double flops = (double)blocks * threads * (double)iters * 8.0 * 2.0;
    double g = flops / (ms * 1e-3) / 1e9;
    if (rep > 0 && g > best) best = g;
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  cudaFree(d);
  *gflops_out = best;
  return RTC_OK;
}

}  // extern "C"
