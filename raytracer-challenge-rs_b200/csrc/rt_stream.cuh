// rt_stream.cuh — the wavefront of rt_wave.cuh WITHOUT level barriers: ONE persistent launch per frame (sm_100a).
//
// The level-synchronous wavefront runs 1 + max_depth forward passes and max_depth combine passes; every pass ends in
// the tail of its slowest batch (≈ 23 µs each on the glass scene) and the next level cannot start before it. Here the
// same work items (one per color_at call, world.rs:70-81) flow through the same per-level two-ended queues, but:
//
//   * one persistent grid (SMs x resident CTAs); a warp takes a TICKET of 32 consecutive slots from the shallowest
//     level that has any published (pixels first: breadth first keeps the queues long and the batches full), deeper
//     levels as soon as they fill — a level never waits for the one above it to drain. A ticket is two atomicAdds that
//     cannot fail destructively: up to 32 TOKENS of the queue end (items published and not handed out; a grab that
//     finds too few puts the excess back), then the end's head by what was granted. (A CAS that reserves "only what
//     exists" fails for every warp but one per L2 round trip under the demand of 2 368 warps: measured, 40 M failures
//     for 50 k batches, 230 ms per frame; an unconditional atomicAdd on the head lets idle warps race far ahead of the
//     producers and sit on tickets that are served last.) The holder runs whichever of its slots are stamped;
//   * queue slots are READY-STAMPED with the frame's epoch and the queue end they belong to. Producers write records
//     and rays while they trace and publish ONCE PER ROUND: after the warp has finished its items, the lanes that
//     emitted store-release the 16-byte item headers (two MEMBAR.ALL.GPU per round instead of one per lane: a
//     device-scope release costs ≈ 0.65 µs per warp instruction, 20 µs when 32 diverged lanes each do their own —
//     scripts/microbench/membar_rate.cu) and add the level's tokens, which is what consumers look at;
//   * the combine is CONTINUATION-PASSING and FENCE-FREE: a suspended shade_hit (WaveRec) counts the children that
//     have arrived (relaxed atomic); the thread that arrives last evaluates shade_hit's last line (world.rs:62-67,
//     same association) with its own colour in registers and the sibling's from the parent's child slot, and carries
//     the result one level up, up to the canvas. Child slots hold a SENTINEL (a NaN payload no arithmetic produces)
//     until a colour is delivered; the last arriver polls the sibling's three doubles until none is the sentinel and
//     puts the sentinel back — 8-byte stores are atomic, so no fence has to order "colour before count";
//   * acquire fences are never used: on sm_100a fence.acq_rel / fence.sc / ld.acquire at device scope are
//     MEMBAR + CCTL.IVALL, i.e. they throw away the SM's whole L1 for every warp on it. Consumers poll with
//     ld.relaxed.gpu and read published data with L2 loads (__ldcg) that are control-dependent on the value polled;
//   * termination: `finished` (items fully processed) equals the number of items ever created (pixels + the queues'
//     reserved counts). A warp that finds no work reads `finished` FIRST and the (monotonic) reserved counts after
//     it: equality then means nothing is in flight or queued.
//
// Per call the arithmetic and its order are the recursive kernel's, so the frame is bit-identical and the per-kind ray
// counts are equal (tests/test_gpu_parity.py::test_wavefront_is_bit_identical runs both wavefront renderers).
// Queue capacities and the overflow fallback are those of rt_wave.cuh: a full queue raises counters[6], every warp
// leaves, and the host re-renders the frame with the recursive kernel (and re-fills the child slots with sentinels).
#pragma once
#include "rt_wave.cuh"

namespace rtk {

constexpr unsigned long long WAVE_SENTINEL = 0x7ff8dead7ff8deadull;  // "no colour delivered yet" (a NaN with a payload)
constexpr size_t STREAM_CTRL_BYTES = 8192;                           // control block: one 128-byte line per word

struct StreamArgs {
  WaveLevel lv[WAVE_MAX_LEVELS + 1];
  char* ctrl;      // see the accessors below
  uint32_t epoch;  // this frame's stamp (1 … 2^31 - 1; the arena is zeroed when allocated)
  int32_t max_depth;
  uint32_t flags;  // STREAM_* (policy switches; RTC_STREAM_FLAGS in the environment overrides the default: experiments)
  uint32_t _pad;
};
enum : uint32_t {
  STREAM_BIG_TICKETS = 1u,   // 128 slots per ticket while a queue holds plenty
  STREAM_EAGER_PUBLISH = 2u, // when work is scarce, emission groups of >= 4 lanes publish at once instead of at the round's end
  STREAM_JOINT_DELIVER = 4u, // the lanes of a round deliver their colours together (converged) after the round's tracing
  STREAM_DEPTH_FIRST = 8u,   // tickets from the DEEPEST level that has a batch (pixels last) instead of the shallowest
};
// per level: reserved so far (front | rear << 32; front = reflection items from the queue's start, rear = refraction
// items from its end), TOKENS of the two ends (items published and not yet handed out), heads of the two ends,
// records; one word per 128-byte line so that the producers' atomics, the consumers' tickets and the polling loads do
// not queue up on one L2 line
__device__ __forceinline__ unsigned long long* sc_items(const StreamArgs& A, int l) { return reinterpret_cast<unsigned long long*>(A.ctrl + 128 * l); }
__device__ __forceinline__ int* sc_tokens(const StreamArgs& A, int l, uint32_t rear) { return reinterpret_cast<int*>(A.ctrl + (rear ? 6144 : 1024) + 128 * l); }
__device__ __forceinline__ uint32_t* sc_pop(const StreamArgs& A, int l, uint32_t rear) { return reinterpret_cast<uint32_t*>(A.ctrl + (rear ? 3072 : 2048) + 128 * l); }
__device__ __forceinline__ uint32_t* sc_recs(const StreamArgs& A, int l) { return reinterpret_cast<uint32_t*>(A.ctrl + 4096 + 128 * l); }
__device__ __forceinline__ unsigned long long* sc_finished(const StreamArgs& A) { return reinterpret_cast<unsigned long long*>(A.ctrl + 5120); }

__device__ __forceinline__ void st_release_u4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.release.gpu.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void st_relaxed_u4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void red_add_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const void* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Child colour slots of [first, first + n) doubles ← sentinel (host: before the first streaming frame in an arena, and
// after a frame that was abandoned).
__global__ void wave_fill_kernel(unsigned long long* __restrict__ p, size_t n, unsigned long long v) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = v;
}

template <bool HAS_CSG, int MINB, int MODE = 0>
__global__ void __launch_bounds__(128, MINB) wave_stream_kernel(SceneView S, CameraD cam, RenderParams P, StreamArgs A,
                                                                double* __restrict__ canvas,
                                                                unsigned long long* __restrict__ counters) {
  const uint32_t lane = threadIdx.x & 31u;
  Scratch<HAS_CSG> scratch;
  RayCounters rc = {0u, 0u, 0u, 0u};
  uint32_t err = 0;
  const int D = A.max_depth;
  const uint32_t pw = (cam.hsize + 7u) >> 3, ph = (P.owned_rows + 3u) >> 2;
  const uint32_t n0 = pw * ph * 32u;  // level 0: one slot per lane of every 8x4 patch of the owned-row domain
  const unsigned long long* overflow = counters + 6;
  constexpr uint32_t NONE = 0xffffffffu;

  int l = 0;                      // level of the ticket in hand (warp-uniform)
  uint32_t parent = 0, slot = 0;  // of this lane's item
  uint32_t pub_a = NONE, pub_b = NONE, pub_ri = 0;  // items this lane has written but not published yet

  auto give_up = [&](unsigned long long why) {  // let the host render the frame recursively (why: 1 a queue is full, 2^32 a colour never arrived)
    atomicAdd(counters + 6, why);
    atomicAdd(counters + 7, 1ull);
  };
  // A finished call's colour goes to its parent: the thread that arrives last at a parent evaluates it and carries the
  // result further up (world.rs:62-67 per level), down to the canvas.
  auto deliver = [&](V3 c, int lvl, uint32_t par, uint32_t sl) {
    for (;;) {
      if (lvl == 0) {
        wave_store_color(canvas + (size_t)par * 3, c);
        return;
      }
      const WaveLevel& up = A.lv[lvl - 1];
      // the record was published (arrival count 0) before this call's item was
      const double2* R = reinterpret_cast<const double2*>(up.recs + par);
      const uint4 h = __ldcg(reinterpret_cast<const uint4*>(R + 3));
      const bool two = (h.z & WR_HAS_REFLECT) && (h.z & FR_NEED_REFRACT);
      V3 other = v3(0.0, 0.0, 0.0);
      if (two) {
        unsigned long long* mine = reinterpret_cast<unsigned long long*>(up.child + ((size_t)par * 2 + sl) * 3);
        unsigned long long* theirs = reinterpret_cast<unsigned long long*>(up.child + ((size_t)par * 2 + (sl ^ 1u)) * 3);
        unsigned long long bx = (unsigned long long)__double_as_longlong(c.x), by = (unsigned long long)__double_as_longlong(c.y),
                           bz = (unsigned long long)__double_as_longlong(c.z);
        // (a colour that IS the sentinel bit pattern can only come from scene data carrying that NaN: keep it a NaN)
        if (bx == WAVE_SENTINEL) bx ^= 1ull;
        if (by == WAVE_SENTINEL) by ^= 1ull;
        if (bz == WAVE_SENTINEL) bz ^= 1ull;
        mine[0] = bx; mine[1] = by; mine[2] = bz;
        uint32_t* arrived = reinterpret_cast<uint32_t*>(up.recs + par) + 15;  // WaveRec::_pad
        if (atomicAdd(arrived, 1u) == 0u) return;  // the sibling is still out: it will find this colour
        // last in. The sibling's count can be visible before its colour is (nothing orders them): poll the doubles.
        unsigned long long ox = ld_relaxed_u64(theirs), oy = ld_relaxed_u64(theirs + 1), oz = ld_relaxed_u64(theirs + 2);
        uint32_t spins = 0;
        while (ox == WAVE_SENTINEL || oy == WAVE_SENTINEL || oz == WAVE_SENTINEL) {
          if (++spins > (1u << 20)) { give_up(1ull << 32); return; }
          __nanosleep(40u);
          ox = ld_relaxed_u64(theirs); oy = ld_relaxed_u64(theirs + 1); oz = ld_relaxed_u64(theirs + 2);
        }
        other = v3(__longlong_as_double((long long)ox), __longlong_as_double((long long)oy), __longlong_as_double((long long)oz));
        mine[0] = WAVE_SENTINEL; mine[1] = WAVE_SENTINEL; mine[2] = WAVE_SENTINEL;      // both slots back to "empty"
        theirs[0] = WAVE_SENTINEL; theirs[1] = WAVE_SENTINEL; theirs[2] = WAVE_SENTINEL;
      }
      const double2 r0 = __ldcg(R), r1 = __ldcg(R + 1), r2 = __ldcg(R + 2);
      const V3 refl_c = sl == 0u ? c : other, refr_c = sl == 0u ? other : c;
      V3 reflected = v3(0.0, 0.0, 0.0), refracted = v3(0.0, 0.0, 0.0);
      if (h.z & WR_HAS_REFLECT) reflected = refl_c * r1.y;    // world.rs:113
      if (h.z & FR_NEED_REFRACT) refracted = refr_c * r2.x;   // world.rs:133
      c = combine(v3(r0.x, r0.y, r1.x), reflected, refracted, r2.y, (h.z & FR_USE_SCHLICK) != 0);
      lvl -= 1;
      par = h.x;
      sl = h.y;
    }
  };
  // Publish what the calling lanes have written (converged among `m`): ONE store-release per lane — its first item's
  // header — is one MEMBAR.ALL.GPU for the group and orders the lane's record and rays before it; a second item's header
  // follows as a plain strong store (program order after the release). Then the level's tokens.
  bool scarce = false;  // work is short (this warp has been polling): children are published the moment they exist
  auto publish = [&](unsigned m) {
    const bool pa = pub_a != NONE, pb = pub_b != NONE;
    const uint32_t ma = __ballot_sync(m, pa), mb = __ballot_sync(m, pb);
    if ((ma | mb) == 0u) return;
    if (pa | pb) {
      WaveItem* first = A.lv[l + 1].items + (pa ? pub_a : pub_b);
      st_release_u4(reinterpret_cast<double2*>(first) + 3, pub_ri, pa ? 0u : 1u, (A.epoch << 1) | (pa ? 0u : 1u), 0u);
      if (pa && pb) st_relaxed_u4(reinterpret_cast<double2*>(A.lv[l + 1].items + pub_b) + 3, pub_ri, 1u, (A.epoch << 1) | 1u, 0u);
    }
    if ((int)lane == __ffs(m) - 1) {
      if (ma) atomicAdd(sc_tokens(A, l + 1, 0u), (int)__popc(ma));
      if (mb) atomicAdd(sc_tokens(A, l + 1, 1u), (int)__popc(mb));
    }
    pub_a = pub_b = NONE;
  };
  // Tracer::advance<WAVE> hands a call that needs children here (divergent: only the lanes that emit in this very
  // advance() are converged). One atomicAdd per counter for the group; record and rays are written straight from
  // registers, the item headers (stamps) at the end of the round.
  auto put = [&](const Frame& f, const RayD& refl, bool has_reflect) {
    const bool has_refr = (f.bits & FR_NEED_REFRACT) != 0;
    const unsigned m = __activemask();
    const unsigned m_refl = __ballot_sync(m, has_reflect), m_refr = __ballot_sync(m, has_refr);
    const int leader = __ffs(m) - 1;
    const uint32_t g_refl = __popc(m_refl), g_refr = __popc(m_refr);
    uint32_t rec0 = 0;
    unsigned long long q0 = 0;
    if ((int)lane == leader) {
      rec0 = atomicAdd(sc_recs(A, l), (uint32_t)__popc(m));
      q0 = atomicAdd(sc_items(A, l + 1), ((unsigned long long)g_refr << 32) | g_refl);
    }
    rec0 = __shfl_sync(m, rec0, leader);
    q0 = __shfl_sync(m, q0, leader);
    const uint32_t front0 = (uint32_t)q0, back0 = (uint32_t)(q0 >> 32);
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t ri = rec0 + __popc(m & lt);
    const uint32_t cap = A.lv[l + 1].cap_items;
    if (rec0 + (uint32_t)__popc(m) > A.lv[l].cap_recs || (unsigned long long)front0 + g_refl + back0 + g_refr > cap) {
      // the group does not fit: the frame is going to be re-rendered; tell every warp now (they are polling)
      if ((int)lane == leader) give_up(1ull);
      return;
    }
    double2* R = reinterpret_cast<double2*>(A.lv[l].recs + ri);
    R[0] = make_double2(f.surface.x, f.surface.y);
    R[1] = make_double2(f.surface.z, f.reflective);
    R[2] = make_double2(f.transparency, f.reflectance);
    reinterpret_cast<uint4*>(R)[3] = make_uint4(parent, slot, f.bits | (has_reflect ? WR_HAS_REFLECT : 0u), 0u);  // arrivals: 0
    pub_ri = ri;
    if (has_reflect) {
      pub_a = front0 + __popc(m_refl & lt);
      double2* w = reinterpret_cast<double2*>(A.lv[l + 1].items + pub_a);
      w[0] = make_double2(refl.o.x, refl.o.y);
      w[1] = make_double2(refl.o.z, refl.d.x);
      w[2] = make_double2(refl.d.y, refl.d.z);
    }
    if (has_refr) {
      pub_b = cap - 1u - (back0 + __popc(m_refr & lt));
      double2* w = reinterpret_cast<double2*>(A.lv[l + 1].items + pub_b);
      w[0] = make_double2(f.refr_ray.o.x, f.refr_ray.o.y);
      w[1] = make_double2(f.refr_ray.o.z, f.refr_ray.d.x);
      w[2] = make_double2(f.refr_ray.d.y, f.refr_ray.d.z);
    }
    // with idle warps about, a group worth a fence does not wait for the end of the round (the rest of the round is
    // tens of µs on the frame's critical path, once per level)
    if ((A.flags & STREAM_EAGER_PUBLISH) && scarce && __popc(m) >= 4) publish(m);
  };
  // lanes 1..D look at one level each, in parallel
  auto frame_done = [&]() -> bool {  // converged; has every item ever created been processed?
    const unsigned long long fin = ld_relaxed_u64(sc_finished(A));
    if (fin < (unsigned long long)n0) return false;  // (also the control dependency that orders the loads below after `fin`)
    unsigned created = 0;
    if (lane >= 1u && (int)lane <= D) {
      const unsigned long long q = ld_relaxed_u64(sc_items(A, (int)lane));
      const uint32_t cap = A.lv[lane].cap_items;
      const uint32_t nf = min((uint32_t)q, cap), nr = min((uint32_t)(q >> 32), cap - nf);
      created = nf + nr;
    }
    unsigned long long total = n0;  // (per-level counts fit 32 bits; their sum over <= 7 levels needs 64)
    for (int k = 1; k <= D; k++) total += __shfl_sync(0xffffffffu, created, k);
    return fin == total;
  };

  Tracer tr;
  RayD ray;
  ray.o = ray.d = v3(0.0, 0.0, 0.0);
  uint32_t idle = 0;
#if defined(RTK_STREAM_STATS)
  // counters[8..]: tickets, items, idle polls, clocks acquiring, clocks waiting for stamps, clocks processing (incl.
  // publication), stamp polls, rounds that ran part of a ticket, tickets taken beyond what was published
  unsigned long long st_batches = 0, st_items = 0, st_idle = 0, st_acq = 0, st_spin = 0, st_work = 0, st_spins = 0, st_partial = 0, st_over = 0;
#define RTK_ST(x) x
#else
#define RTK_ST(x)
#endif
  uint32_t pending = 0, t_base = 0, t_rear = 0, t_left = 0, polls = 0;
  for (;;) {
    RTK_ST(const long long tc0 = clock64();)
    if (pending == 0u && t_left > 0u) {  // the next 32 slots of a long ticket
      t_base += 32u;
      t_left -= 32u;
      pending = 0xffffffffu;
      polls = 0;
    }
    if (pending == 0u) {
      // ---- take a ticket: pixels first, then the shallowest level with a full batch published, then (after an idle
      // poll) the shallowest level with anything published ------------------------------------------------------------
      // A ticket is two atomics, neither of which can fail destructively: take up to 32 TOKENS of a queue end (tokens =
      // items published and not handed out; a grab that finds too few puts back what it took too many of), then
      // advance that end's head by what was granted. Heads therefore never pass the published count: every slot of a
      // ticket is at worst a few stores from being stamped.
      unsigned long long ov = 0;
      int tf = 0, tr_ = 0;
      uint32_t p0 = 0;
      if (lane == 0u) {
        p0 = ld_relaxed_u32(sc_pop(A, 0, 0u));
        ov = ld_relaxed_u64(overflow);
      } else if ((int)lane <= D) {
        tf = (int)ld_relaxed_u32(reinterpret_cast<const uint32_t*>(sc_tokens(A, (int)lane, 0u)));
        tr_ = (int)ld_relaxed_u32(reinterpret_cast<const uint32_t*>(sc_tokens(A, (int)lane, 1u)));
      }
      if (__shfl_sync(0xffffffffu, ov != 0ull ? 1 : 0, 0)) break;
      const bool level_lane = lane >= 1u && (int)lane <= D;
      const uint32_t m_pix = __ballot_sync(0xffffffffu, lane == 0u && p0 < n0);
      const uint32_t m_full = __ballot_sync(0xffffffffu, level_lane && (tf >= 32 || tr_ >= 32));
      const uint32_t m_plenty = __ballot_sync(0xffffffffu, level_lane && (tf >= 16384 || tr_ >= 16384));
      const uint32_t m_any = __ballot_sync(0xffffffffu, level_lane && (tf > 0 || tr_ > 0));
      uint32_t got = 0;  // slots granted
      uint32_t b_base = 0, b_rear = 0;
      int picker;
      if (A.flags & STREAM_DEPTH_FIRST) {
        // deepest full batch, else a pixel ticket, else (after an idle poll) the deepest level with anything
        const uint32_t m_pick = m_full ? m_full : (m_pix ? m_pix : (idle > 0u ? m_any : 0u));
        picker = m_pick ? 31 - __clz(m_pick) : -1;
      } else {
        const uint32_t m_pick = m_pix ? m_pix : (m_full ? m_full : (idle > 0u ? m_any : 0u));
        picker = m_pick ? __ffs(m_pick) - 1 : -1;
      }
      if ((int)lane == picker) {
        if (lane == 0u) {
          // pixels: 128 at a time while there are plenty (the acquisition is three L2 round trips: 2.4 µs)
          const uint32_t want = ((A.flags & STREAM_BIG_TICKETS) && p0 + 65536u <= n0) ? 128u : 32u;
          b_base = atomicAdd(sc_pop(A, 0, 0u), want);
          got = b_base < n0 ? min(want, n0 - b_base) : 0u;
        } else {
          b_rear = (tf >= 32 || (tr_ < 32 && tf >= tr_)) ? 0u : 1u;
          int* tok = sc_tokens(A, (int)lane, b_rear);
          const int want = ((A.flags & STREAM_BIG_TICKETS) && (b_rear ? tr_ : tf) >= 16384) ? 128 : 32;
          const int old = atomicAdd(tok, -want);
          const int grant = old >= want ? want : (old >= 32 ? (old & ~31) : ((old > 0 && m_full == 0u) ? old : 0));
          if (grant < want) atomicAdd(tok, want - grant);
          RTK_ST(if (grant == 0) st_over++;)
          if (grant > 0) {
            b_base = atomicAdd(sc_pop(A, (int)lane, b_rear), (uint32_t)grant);
            got = (uint32_t)grant;
          }
        }
      }
      if (picker >= 0) {
        got = __shfl_sync(0xffffffffu, got, picker);
        b_base = __shfl_sync(0xffffffffu, b_base, picker);
        b_rear = __shfl_sync(0xffffffffu, b_rear, picker);
      }
      if (got == 0u) {
        if (picker < 0 && frame_done()) break;
        RTK_ST(st_idle++; st_acq += (unsigned long long)(clock64() - tc0);)
        if (picker < 0) {
          idle++;
          __nanosleep(idle < 8u ? 100u : 400u);
        }
        continue;
      }
      // scarce: nothing but this ticket was on offer, or the warp had to wait for it
      scarce = idle > 0u || (m_pix == 0u && m_plenty == 0u);
      idle = 0;
      l = picker;
      t_base = b_base;
      t_rear = b_rear;
      pending = got >= 32u ? 0xffffffffu : ((1u << got) - 1u);
      t_left = got > 32u ? got - 32u : 0u;
      polls = 0;
      RTK_ST(st_batches++;)
    }
    RTK_ST(const long long tc1 = clock64(); st_acq += (unsigned long long)(tc1 - tc0);)
    // ---- which of the ticket's lanes can run now ---------------------------------------------------------------------
    bool have = false;
    const bool mine = (pending >> lane) & 1u;
    const double2* q = nullptr;
    if (l == 0) {
      if (mine) {
        const uint32_t idx = t_base + lane;
        const uint32_t patch = idx >> 5, li = idx & 31u;
        const uint32_t px = (patch % pw) * 8u + (li & 7u), oy = (patch / pw) * 4u + (li >> 3);
        const uint32_t k = oy / P.tile_rows, r = oy - k * P.tile_rows;
        const uint32_t py = (P.rank + k * P.nranks) * P.tile_rows + r;
        if (idx < n0 && px < cam.hsize && oy < P.owned_rows && py >= P.row_begin && py < P.row_end) {
          rc.primary++;
          ray = ray_for_pixel(cam, px, py, 0.5, 0.5);
          parent = py * cam.hsize + px;  // Canvas::pixel_idx, canvas.rs:44-48
          slot = 0;
          have = true;
        }
      }
      const uint32_t live = __ballot_sync(0xffffffffu, have);
      if (lane == 0u && live != pending) red_add_u64(sc_finished(A), (unsigned long long)__popc(pending & ~live));  // slots without a pixel
      pending = live;
    } else {
      const uint32_t k = t_base + lane;
      const uint32_t at = t_rear ? A.lv[l].cap_items - 1u - k : k;
      q = reinterpret_cast<const double2*>(A.lv[l].items + at);
      // the stamp names the frame AND the queue end: the other end's items live in the same array, from the far side.
      // (Stamp first, rays after it has been seen: four loads issued together could return a fresh stamp with stale rays.)
      if (mine) have = ld_relaxed_u32(reinterpret_cast<const uint32_t*>(q + 3) + 2) == ((A.epoch << 1) | t_rear);
    }
    const uint32_t ready = __ballot_sync(0xffffffffu, have);
    if (l != 0 && ready != pending) {
      // every slot of the ticket has been published by COUNT, but publication is not in slot order: a slot can be a
      // few stores from its stamp. Poll a few times for the whole ticket, then run what there is.
      polls++;
      RTK_ST(st_spins++;)
      if (ready == 0u || polls < 4u) {
        if ((polls & 15u) == 0u && __shfl_sync(0xffffffffu, lane == 0u && ld_relaxed_u64(overflow) != 0ull ? 1 : 0, 0)) break;
        RTK_ST(st_spin += (unsigned long long)(clock64() - tc1);)
        __nanosleep(polls < 8u ? 40u : 200u);
        continue;
      }
      RTK_ST(st_partial++;)
    }
    polls = 0;
    if (l != 0 && have) {
      const double2 a = __ldcg(q), b = __ldcg(q + 1), c = __ldcg(q + 2);
      const uint4 h = __ldcg(reinterpret_cast<const uint4*>(q + 3));
      ray.o = v3(a.x, a.y, b.x); ray.d = v3(b.y, c.x, c.y);
      parent = h.x; slot = h.y;
    }
    RTK_ST(const long long tc2 = clock64(); st_spin += (unsigned long long)(tc2 - tc1); st_items += __popc(ready);)
    if (have) {
      tr.start(ray, D - l, nullptr, &err);
      for (;;) {
        search<HAS_CSG, MODE>(S, tr.qray, tr.q, scratch.csg(), &scratch.bvh, &err);
        if (tr.template advance<true>(S, nullptr, scratch.neg, rc, &err, put)) break;
      }
      if (!(A.flags & STREAM_JOINT_DELIVER) && !tr.w_emit) deliver(tr.result, l, parent, slot);
    }
    __syncwarp();
    // the arrival atomics, sibling polls and record loads of a delivery are dependent L2 round trips: lane by lane as
    // each finishes its trace they add up; together they overlap
    if ((A.flags & STREAM_JOINT_DELIVER) && have && !tr.w_emit) deliver(tr.result, l, parent, slot);
    __syncwarp();
    // ---- publish what this round emitted: records and rays (plain stores above) before the stamps (store-release,
    // one MEMBAR per kind for the warp), then the level's published count -------------------------------------------------
    publish(0xffffffffu);
    RTK_ST(st_work += (unsigned long long)(clock64() - tc2);)
    pending &= ~ready;
    // the lanes' reservations (atomics on the reserved counts, whose results they have already used) are at L2 by now
    if (lane == 0u) red_add_u64(sc_finished(A), (unsigned long long)__popc(ready));
  }
#if defined(RTK_STREAM_STATS)
  st_spins = __reduce_add_sync(0xffffffffu, (unsigned)st_spins);
  if (lane == 0) {
    atomicAdd(counters + 8, st_batches); atomicAdd(counters + 9, st_items); atomicAdd(counters + 10, st_idle);
    atomicAdd(counters + 11, st_acq); atomicAdd(counters + 12, st_spin); atomicAdd(counters + 13, st_work);
    atomicAdd(counters + 14, st_spins / 32u); atomicAdd(counters + 15, st_partial); atomicAdd(counters + 16, st_over);
  }
#endif
  flush_counters(rc, err, counters);
}

}  // namespace rtk
