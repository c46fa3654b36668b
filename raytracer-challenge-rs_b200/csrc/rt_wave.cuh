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
// (SMs x resident CTAs, grid-stride over a device-side count) queued back to back on one stream: no
// host round trip between levels. Queue capacity is optimistic (2 items per pixel and level, the
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
  uint32_t* n_items;      // device counters, one per level
  uint32_t* n_recs;
  uint32_t* work;         // per pass: items handed out so far
  int32_t level;          // the level whose queue this pass drains
  int32_t max_depth;
  uint32_t inline_min;    // keep the first child in-lane when at least this many lanes of the warp emit (33: never)
  uint32_t _pad;
};

// counters[6]: items dropped because a queue was full (→ the host re-renders with the recursive kernel)
__device__ __forceinline__ void wave_store_color(double* dst, V3 c) {
  dst[0] = c.x; dst[1] = c.y; dst[2] = c.z;
}

// Per-thread staging of an emitted call in SHARED memory (field-major, so lane-consecutive = bank-
// consecutive): the suspended shade_hit and both child rays leave the register file the moment
// Tracer::advance produces them, and come back after the warp has agreed on queue slots.
constexpr int WAVE_STAGE_DOUBLES = 18;  // surface 3, reflective, transparency, reflectance, reflection ray 6, refraction ray 6
struct WaveStage {
  double d[WAVE_STAGE_DOUBLES][128];
  uint32_t bits[128];
};

// One forward pass over the queue of level A.level. PRIMARY: items are the pixels of this rank's tiles, a
// warp = one 8x4 patch. Work is handed out 32 items at a time from a device counter (patches / queue
// slices differ in cost by more than 10x; a static split leaves the chip waiting for the unluckiest warp).
//
// A call that needs children may KEEP one of them: the lane stores its WaveRec and carries on with the
// reflection ray (the refraction ray when there is no reflection) as a call of the next level, in the
// same pass — no queue round trip for mirror chains; only the second child of a glass hit is queued.
// A thread's serial chain is then at most max_depth + 1 calls, against 2^(max_depth+1) - 1 recursively.
// All lanes of a warp move down one level per step, so the level is warp-uniform. Keeping pays while
// most of the warp carries on (A.inline_min); a sparse warp queues everything instead.
template <bool HAS_CSG, bool PRIMARY, int MINB>
__global__ void __launch_bounds__(128, MINB) wave_forward_kernel(SceneView S, CameraD cam, RenderParams P, WaveArgs A,
                                                                 double* __restrict__ canvas,
                                                                 unsigned long long* __restrict__ counters) {
  __shared__ WaveStage stage;
  // a queue overflowed in an earlier pass: the frame is going to be re-rendered, and the queues have holes
  if (!PRIMARY && *reinterpret_cast<const volatile unsigned long long*>(counters + 6) != 0ull) return;
  const uint32_t lane = threadIdx.x & 31u, tid = threadIdx.x;
  Scratch<HAS_CSG> scratch;
  RayCounters rc = {0u, 0u, 0u, 0u};
  uint32_t err = 0, dropped = 0;
  // patches: 8 px wide, 4 rows tall over the owned-row domain
  const uint32_t pw = (cam.hsize + 7u) >> 3, ph = (P.owned_rows + 3u) >> 2;
  const uint32_t n = PRIMARY ? pw * ph * 32u : min(A.n_items[A.level], A.lv[A.level].cap_items);
  // the next batch's index is requested before the current batch is processed: the atomic's round trip
  // to L2 overlaps the work instead of preceding it
  uint32_t next = 0;
  if (lane == 0) next = atomicAdd(A.work, 32u);
  for (;;) {
    const uint32_t base = __shfl_sync(0xffffffffu, next, 0);
    if (base >= n) break;
    if (lane == 0) next = atomicAdd(A.work, 32u);
    const uint32_t idx = base + lane;
    bool active = idx < n;
    RayD ray;
    ray.o = ray.d = v3(0.0, 0.0, 0.0);
    uint32_t parent = 0, slot = 0;
    if (PRIMARY) {
      const uint32_t patch = base >> 5;
      const uint32_t px = (patch % pw) * 8u + (lane & 7u), oy = (patch / pw) * 4u + (lane >> 3);
      const uint32_t k = oy / P.tile_rows, r = oy - k * P.tile_rows;
      const uint32_t py = (P.rank + k * P.nranks) * P.tile_rows + r;
      active = px < cam.hsize && oy < P.owned_rows && py >= P.row_begin && py < P.row_end;
      if (active) {
        rc.primary++;
        ray = ray_for_pixel(cam, px, py, 0.5, 0.5);
        parent = py * cam.hsize + px;  // Canvas::pixel_idx, canvas.rs:44-48
      }
    } else if (active) {
      const WaveItem* it = A.lv[A.level].items + idx;
      const double2* q = reinterpret_cast<const double2*>(it);
      const double2 a = __ldcs(q), b = __ldcs(q + 1), c = __ldcs(q + 2);
      const int4 h = __ldcs(reinterpret_cast<const int4*>(q + 3));
      ray.o = v3(a.x, a.y, b.x); ray.d = v3(b.y, c.x, c.y);
      parent = (uint32_t)h.x; slot = (uint32_t)h.y;
    }
    for (int l = PRIMARY ? 0 : A.level;; l++) {  // the chain of kept children; l is warp-uniform
      bool emit = false;
      V3 result = v3(0.0, 0.0, 0.0);
      if (active) {
        Tracer tr;
        tr.start(ray, A.max_depth - l, nullptr, &err);
        auto put = [&](const Frame& f, const RayD& refl, bool has_reflect) {
          stage.d[0][tid] = f.surface.x; stage.d[1][tid] = f.surface.y; stage.d[2][tid] = f.surface.z;
          stage.d[3][tid] = f.reflective; stage.d[4][tid] = f.transparency; stage.d[5][tid] = f.reflectance;
          stage.d[6][tid] = refl.o.x; stage.d[7][tid] = refl.o.y; stage.d[8][tid] = refl.o.z;
          stage.d[9][tid] = refl.d.x; stage.d[10][tid] = refl.d.y; stage.d[11][tid] = refl.d.z;
          stage.d[12][tid] = f.refr_ray.o.x; stage.d[13][tid] = f.refr_ray.o.y; stage.d[14][tid] = f.refr_ray.o.z;
          stage.d[15][tid] = f.refr_ray.d.x; stage.d[16][tid] = f.refr_ray.d.y; stage.d[17][tid] = f.refr_ray.d.z;
          stage.bits[tid] = f.bits | (has_reflect ? WR_HAS_REFLECT : 0u);
        };
        for (;;) {
          search<HAS_CSG, false>(S, tr.qray, tr.q, scratch.csg(), &scratch.bvh, &err);
          if (tr.template advance<true>(S, nullptr, scratch.neg, rc, &err, put)) break;
        }
        emit = tr.w_emit;
        result = tr.result;
      }
      __syncwarp();
      if (active && !emit) {
        double* dst = l == 0 ? canvas + (size_t)parent * 3 : A.lv[l - 1].child + ((size_t)parent * 2 + slot) * 3;
        wave_store_color(dst, result);
      }
      // warp-aggregated allocation: one record per emitting lane and a queue slot per queued child. The
      // first child is KEPT (the lane carries on with it) only if at least A.inline_min lanes emit: a
      // mostly idle warp is better off handing everything to the queue, where the next pass packs it.
      const uint32_t m_emit = __ballot_sync(0xffffffffu, emit);
      if (m_emit == 0u) break;
      const bool keep = (uint32_t)__popc(m_emit) >= A.inline_min;
      const uint32_t bits = emit ? stage.bits[tid] : 0u;
      const bool has_refl = (bits & WR_HAS_REFLECT) != 0, has_refr = (bits & FR_NEED_REFRACT) != 0;
      // queued: (reflection unless kept) and (refraction unless it is the kept first child)
      const bool q_refl = has_refl && !keep, q_refr = has_refr && (has_refl || !keep);
      const uint32_t m_qa = __ballot_sync(0xffffffffu, q_refl), m_qb = __ballot_sync(0xffffffffu, q_refr);
      const uint32_t n_qa = __popc(m_qa);
      uint32_t rec0 = 0, item0 = 0;
      if (lane == 0) {
        rec0 = atomicAdd(A.n_recs + l, (uint32_t)__popc(m_emit));
        if (m_qa | m_qb) item0 = atomicAdd(A.n_items + l + 1, n_qa + (uint32_t)__popc(m_qb));
      }
      rec0 = __shfl_sync(0xffffffffu, rec0, 0);
      item0 = __shfl_sync(0xffffffffu, item0, 0);
      const uint32_t lt = (1u << lane) - 1u;
      active = false;
      if (emit) {
        const uint32_t ri = rec0 + __popc(m_emit & lt);
        const uint32_t qa = item0 + __popc(m_qa & lt), qb = item0 + n_qa + __popc(m_qb & lt);
        const uint32_t cap = A.lv[l + 1].cap_items;
        const bool fits = ri < A.lv[l].cap_recs && (!q_refl || qa < cap) && (!q_refr || qb < cap);
        if (!fits) {
          dropped++;
        } else {
          double2* R = reinterpret_cast<double2*>(A.lv[l].recs + ri);
          R[0] = make_double2(stage.d[0][tid], stage.d[1][tid]);
          R[1] = make_double2(stage.d[2][tid], stage.d[3][tid]);
          R[2] = make_double2(stage.d[4][tid], stage.d[5][tid]);
          reinterpret_cast<uint4*>(R)[3] = make_uint4(parent, slot, bits, 0u);
          if (q_refl) {
            double2* w = reinterpret_cast<double2*>(A.lv[l + 1].items + qa);
            w[0] = make_double2(stage.d[6][tid], stage.d[7][tid]);
            w[1] = make_double2(stage.d[8][tid], stage.d[9][tid]);
            w[2] = make_double2(stage.d[10][tid], stage.d[11][tid]);
            reinterpret_cast<uint4*>(w)[3] = make_uint4(ri, 0u, 0u, 0u);
          }
          if (q_refr) {
            double2* w = reinterpret_cast<double2*>(A.lv[l + 1].items + qb);
            w[0] = make_double2(stage.d[12][tid], stage.d[13][tid]);
            w[1] = make_double2(stage.d[14][tid], stage.d[15][tid]);
            w[2] = make_double2(stage.d[16][tid], stage.d[17][tid]);
            reinterpret_cast<uint4*>(w)[3] = make_uint4(ri, 1u, 0u, 0u);
          }
          if (keep) {  // this lane carries on with its first child
            const int o = has_refl ? 6 : 12;
            ray.o = v3(stage.d[o][tid], stage.d[o + 1][tid], stage.d[o + 2][tid]);
            ray.d = v3(stage.d[o + 3][tid], stage.d[o + 4][tid], stage.d[o + 5][tid]);
            parent = ri;
            slot = has_refl ? 0u : 1u;
            active = true;
          }
        }
      }
      __syncwarp();
      if (!keep) break;
    }
  }
  flush_counters(rc, err, counters);
  if (dropped) atomicAdd(counters + 6, (unsigned long long)dropped);
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
