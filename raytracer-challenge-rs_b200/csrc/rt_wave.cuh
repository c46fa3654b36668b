// rt_wave.cuh — the Whitted recursion as a level-synchronous WAVEFRONT (sm_100a).
//
// The one-thread-per-pixel kernel (rt_kernels.cuh) evaluates a pixel's whole reflected_color /
// refracted_color tree (world.rs:107-134; up to 63 color_at calls at depth 5) on ONE thread. A glass
// pixel is then a ~1 ms serial chain, which is what bounds the frame once the image is sharded over
// several GPUs (an eighth of the 1080p glass frame still takes 1.1 ms instead of 0.37). Here every
// color_at call is its own work item:
//
//   level 0   one item per pixel (ray_for_pixel, camera.rs:57-69)
//   forward   level L: closest hit → shade_hit's light loop (world.rs:41-58) → containers query →
//             schlick / refraction set-up, all through the SAME Tracer state machine and the same
//             single search site as the recursive kernel, but with remaining = max_depth - L and no
//             descent: a call that needs children stores its suspended shade_hit as a 64-byte
//             WaveRec and appends the reflection / refraction rays (world.rs:111, :131) to level
//             L+1's queue (warp-aggregated: two ballots, one atomicAdd per warp and queue); a call
//             without children writes its colour straight to its parent's child slot (level 0: the
//             canvas — local, peer-mapped or pinned host memory, as before).
//   combine   levels max_depth-1 … 0: shade_hit's last line (world.rs:62-67) for every WaveRec, in the
//             reference's association (surface + reflected·k) + refracted·k', result to the
//             parent's slot / the canvas.
//
// Every arithmetic expression and its order per call is the recursive kernel's, so the frame is
// bit-identical (tests/test_gpu_parity.py::test_wavefront_*). All launches are persistent grids
// (SMs x resident CTAs, 32 items per warp from a device-side counter) queued back to back on one stream:
// no host round trip between levels. Queue capacity is optimistic (2 items per pixel and level, the
// worst case is 2^L); an overflow raises a flag and rtc_scene_finish re-renders the frame with the
// recursive kernel.
#pragma once
#include "rt_kernels.cuh"

namespace rtk {

struct WaveItem {   // 64 B: a pending color_at(ray, max_depth - level)
  RayD ray;
  uint32_t parent;  // WaveRec index in the previous level
  uint32_t slot;    // 0: its reflection child, 1: its refraction child
  uint32_t _pad[2];
};
struct WaveRec {    // 64 B: a suspended shade_hit (Frame, rt_shade.cuh) waiting for its children
  V3 surface;
  double reflective, transparency, reflectance;
  uint32_t parent, slot;  // where this call's own colour goes (level 0: parent = pixel index in the canvas)
  uint32_t bits, _pad;    // FR_NEED_REFRACT | FR_USE_SCHLICK | WR_HAS_REFLECT
};
enum : uint32_t { WR_HAS_REFLECT = 4u };
static_assert(sizeof(WaveItem) == 64 && sizeof(WaveRec) == 64, "wavefront record layout");

constexpr int WAVE_MAX_LEVELS = MAX_FRAMES;  // levels 0 … max_depth

struct WaveLevel {
  WaveItem* items;    // this level's queue (level 0: none, items are pixels)
  WaveRec* recs;      // this level's suspended calls
  double* child;      // [rec][2][3]: colours returned by the children of this level's recs
  uint32_t cap_items, cap_recs;
};
struct WaveArgs {
  WaveLevel lv[WAVE_MAX_LEVELS + 1];
  unsigned long long* n_items;  // per level: low word = reflection items (stored from the front of the level's
                                // queue), high word = refraction items (stored from its END downwards); one atomic
  uint32_t* n_recs;
  uint32_t* work;         // per pass: items handed out so far
  int32_t level;          // the level whose queue this pass drains
  int32_t max_depth;
  uint32_t _pad[2];
};

// counters[6]: items dropped because a queue was full (→ the host re-renders with the recursive kernel)
__device__ __forceinline__ void wave_store_color(double* dst, V3 c) {
  dst[0] = c.x; dst[1] = c.y; dst[2] = c.z;
}

// One forward pass over the queue of level A.level. PRIMARY: items are the pixels of this rank's tiles, a
// warp = one 8x4 patch. Work is handed out 32 items at a time from a device counter (patches / queue
// slices differ in cost by more than 10x; a static split leaves the chip waiting for the unluckiest warp).
//
// A call that needs children reserves its WaveRec and queue slots the moment Tracer::advance produces
// them, straight from registers: the lanes that reach the emission together (__activemask) elect a
// leader, which does ONE atomicAdd per counter for the group (per-lane atomics on one address serialise
// in a single L2 slice: 6.0 ms instead of 3.9), and every lane writes its 64-byte record and its
// children's 64-byte items with 16-byte stores. No shared memory: L1 keeps its full size.
template <bool HAS_CSG, bool PRIMARY, int MINB, int MODE = 0>
__global__ void __launch_bounds__(128, MINB) wave_forward_kernel(SceneView S, CameraD cam, RenderParams P, WaveArgs A,
                                                                 double* __restrict__ canvas,
                                                                 unsigned long long* __restrict__ counters) {
  // a queue overflowed in an earlier pass: the frame is going to be re-rendered, and the queues have holes
  if (!PRIMARY && *reinterpret_cast<const volatile unsigned long long*>(counters + 6) != 0ull) return;
  const uint32_t lane = threadIdx.x & 31u;
  Scratch<HAS_CSG> scratch;
  RayCounters rc = {0u, 0u, 0u, 0u};
  uint32_t err = 0, dropped = 0;
  const int l = PRIMARY ? 0 : A.level;
  // patches: 8 px wide, 4 rows tall over the owned-row domain
  const uint32_t pw = (cam.hsize + 7u) >> 3, ph = (P.owned_rows + 3u) >> 2;
  // A level's queue holds the reflection rays from its front and the refraction rays from its end: the warps of
  // the next pass are then all-reflection (wall hits: closest + shadow) or all-refraction (inside glass:
  // closest + containers) and their lanes finish together far more often.
  uint32_t n_front = 0, n_rear = 0;
  if (!PRIMARY) {
    const unsigned long long q = A.n_items[l];
    n_front = min((uint32_t)q, A.lv[l].cap_items);
    n_rear = min((uint32_t)(q >> 32), A.lv[l].cap_items - n_front);
  }
  const uint32_t n = PRIMARY ? pw * ph * 32u : n_front + n_rear;
  uint32_t parent = 0, slot = 0;
  // item idx of this pass → this lane's ray / parent / slot; false: nothing to trace (pixel outside the frame)
  auto load_item = [&](uint32_t idx, RayD& ray) -> bool {
    if (PRIMARY) {
      const uint32_t patch = idx >> 5, li = idx & 31u;
      const uint32_t px = (patch % pw) * 8u + (li & 7u), oy = (patch / pw) * 4u + (li >> 3);
      const uint32_t k = oy / P.tile_rows, r = oy - k * P.tile_rows;
      const uint32_t py = (P.rank + k * P.nranks) * P.tile_rows + r;
      if (!(px < cam.hsize && oy < P.owned_rows && py >= P.row_begin && py < P.row_end)) return false;
      rc.primary++;
      ray = ray_for_pixel(cam, px, py, 0.5, 0.5);
      parent = py * cam.hsize + px;  // Canvas::pixel_idx, canvas.rs:44-48
    } else {
      const uint32_t at = idx < n_front ? idx : A.lv[l].cap_items - 1u - (idx - n_front);
      const double2* q = reinterpret_cast<const double2*>(A.lv[l].items + at);
      const double2 a = __ldcs(q), b = __ldcs(q + 1), c = __ldcs(q + 2);
      const int4 h = __ldcs(reinterpret_cast<const int4*>(q + 3));
      ray.o = v3(a.x, a.y, b.x); ray.d = v3(b.y, c.x, c.y);
      parent = (uint32_t)h.x; slot = (uint32_t)h.y;
    }
    return true;
  };
  // Tracer::advance<WAVE> hands a call that needs children here (divergent code: only the lanes that
  // emit in this very advance() are converged at this point)
  auto put = [&](const Frame& f, const RayD& refl, bool has_reflect) {
    const bool has_refr = (f.bits & FR_NEED_REFRACT) != 0;
    const unsigned m = __activemask();
    const unsigned m_refl = __ballot_sync(m, has_reflect), m_refr = __ballot_sync(m, has_refr);
    const int leader = __ffs(m) - 1;
    uint32_t rec0 = 0, item0 = 0;
    const uint32_t g_refl = __popc(m_refl), g_refr = __popc(m_refr);
    unsigned long long q0 = 0;
    if ((int)lane == leader) {
      rec0 = atomicAdd(A.n_recs + l, (uint32_t)__popc(m));
      q0 = atomicAdd(A.n_items + l + 1, ((unsigned long long)g_refr << 32) | g_refl);  // both ends in one atomic
    }
    rec0 = __shfl_sync(m, rec0, leader);
    q0 = __shfl_sync(m, q0, leader);
    const uint32_t front0 = (uint32_t)q0, back0 = (uint32_t)(q0 >> 32);
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t ri = rec0 + __popc(m & lt);
    const uint32_t cap = A.lv[l + 1].cap_items;
    // the two ends have not met iff everything reserved so far, this group included, fits the queue
    if (ri >= A.lv[l].cap_recs || (unsigned long long)front0 + g_refl + back0 + g_refr > cap) { dropped++; return; }
    const uint32_t i_refl = front0 + __popc(m_refl & lt), i_refr = cap - 1u - (back0 + __popc(m_refr & lt));
    (void)item0;
    double2* R = reinterpret_cast<double2*>(A.lv[l].recs + ri);
    R[0] = make_double2(f.surface.x, f.surface.y);
    R[1] = make_double2(f.surface.z, f.reflective);
    R[2] = make_double2(f.transparency, f.reflectance);
    reinterpret_cast<uint4*>(R)[3] = make_uint4(parent, slot, f.bits | (has_reflect ? WR_HAS_REFLECT : 0u), 0u);
    if (has_reflect) {
      double2* w = reinterpret_cast<double2*>(A.lv[l + 1].items + i_refl);
      w[0] = make_double2(refl.o.x, refl.o.y);
      w[1] = make_double2(refl.o.z, refl.d.x);
      w[2] = make_double2(refl.d.y, refl.d.z);
      reinterpret_cast<uint4*>(w)[3] = make_uint4(ri, 0u, 0u, 0u);
    }
    if (has_refr) {
      double2* w = reinterpret_cast<double2*>(A.lv[l + 1].items + i_refr);
      w[0] = make_double2(f.refr_ray.o.x, f.refr_ray.o.y);
      w[1] = make_double2(f.refr_ray.o.z, f.refr_ray.d.x);
      w[2] = make_double2(f.refr_ray.d.y, f.refr_ray.d.z);
      reinterpret_cast<uint4*>(w)[3] = make_uint4(ri, 1u, 0u, 0u);
    }
  };
  auto deliver = [&](V3 c) {
    double* dst = l == 0 ? canvas + (size_t)parent * 3 : A.lv[l - 1].child + ((size_t)parent * 2 + slot) * 3;
    wave_store_color(dst, c);
  };
  Tracer tr;
  RayD ray;
  ray.o = ray.d = v3(0.0, 0.0, 0.0);
#if defined(RTK_WAVE_REFILL)
  // Lane-level streams: a lane that has finished its call takes the next item at the top of the very next
  // turn (the lanes that are idle together fetch together: one atomicAdd per group), so the warp keeps
  // meeting at the one search site with most lanes busy, in whatever phase of whatever call each is.
  bool busy = false, drained = false;
  for (;;) {
    if (!busy && !drained) {
      const unsigned m = __activemask();
      const int leader = __ffs(m) - 1;
      uint32_t base = 0;
      if ((int)lane == leader) base = atomicAdd(A.work, (uint32_t)__popc(m));
      base = __shfl_sync(m, base, leader);
      const uint32_t idx = base + __popc(m & ((1u << lane) - 1u));
      if (idx >= n) {
        drained = true;
      } else if (load_item(idx, ray)) {
        tr.start(ray, A.max_depth - l, nullptr, &err);
        busy = true;
      }
    }
    __syncwarp();
    if (!__any_sync(0xffffffffu, busy)) {
      if (__all_sync(0xffffffffu, drained)) break;
      continue;
    }
    if (busy) {
      search<HAS_CSG, MODE>(S, tr.qray, tr.q, scratch.csg(), &scratch.bvh, &err);
      if (tr.template advance<true>(S, nullptr, scratch.neg, rc, &err, put)) {
        if (!tr.w_emit) deliver(tr.result);
        busy = false;
      }
    }
    __syncwarp();
  }
#else
  // Warp batches of 32 consecutive items. The next batch's index is requested before the current batch
  // is processed: the atomic's round trip to L2 overlaps the work instead of preceding it.
  uint32_t next = 0;
  if (lane == 0) next = atomicAdd(A.work, 32u);
  for (;;) {
    const uint32_t base = __shfl_sync(0xffffffffu, next, 0);
    if (base >= n) break;
    if (lane == 0) next = atomicAdd(A.work, 32u);
    if (base + lane < n && load_item(base + lane, ray)) {
      tr.start(ray, A.max_depth - l, nullptr, &err);
      for (;;) {
        search<HAS_CSG, MODE>(S, tr.qray, tr.q, scratch.csg(), &scratch.bvh, &err);
        if (tr.template advance<true>(S, nullptr, scratch.neg, rc, &err, put)) break;
      }
      if (!tr.w_emit) deliver(tr.result);
    }
    __syncwarp();
  }
#endif
  flush_counters(rc, err, counters);
  if (dropped) { atomicAdd(counters + 6, (unsigned long long)dropped); atomicAdd(counters + 7, 1ull); }
}

// shade_hit's last line (world.rs:62-67) for the suspended calls of one level.
template <bool LEVEL0>
__global__ void __launch_bounds__(256) wave_combine_kernel(const WaveRec* __restrict__ recs, const double* __restrict__ child,
                                                           const uint32_t* __restrict__ n_recs, uint32_t cap,
                                                           double* __restrict__ parent_child, double* __restrict__ canvas,
                                                           const unsigned long long* __restrict__ counters) {
  if (*reinterpret_cast<const volatile unsigned long long*>(counters + 6) != 0ull) return;  // see wave_forward_kernel
  uint32_t n = *n_recs;
  if (n > cap) n = cap;
  for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
    const WaveRec R = recs[r];
    const double* c = child + (size_t)r * 6;
    V3 reflected = v3(0.0, 0.0, 0.0);
    if (R.bits & WR_HAS_REFLECT) reflected = v3(c[0], c[1], c[2]) * R.reflective;             // world.rs:113
    V3 refracted = v3(0.0, 0.0, 0.0);
    if (R.bits & FR_NEED_REFRACT) refracted = v3(c[3], c[4], c[5]) * R.transparency;          // world.rs:133
    const V3 res = combine(R.surface, reflected, refracted, R.reflectance, (R.bits & FR_USE_SCHLICK) != 0);
    double* dst = LEVEL0 ? canvas + (size_t)R.parent * 3 : parent_child + ((size_t)R.parent * 2 + R.slot) * 3;
    wave_store_color(dst, res);
  }
}

}  // namespace rtk
