// rt_wide.cuh -- the default kernels of the RT path.
//
// One persistent kernel per vx_start.  Design (DESIGN.md "Kernels"):
//   * persistent warps: grid = #SM x resident CTAs; warps pull chunks of 32 jobs (pixels, or rays
//     of an explicit list) from one global atomic cursor, so no warp ever waits for a tail CTA;
//   * every lane is a small state machine (primary -> shadow -> bounce -> bounce shadow) around ONE
//     generic traversal core: closest-hit and any-hit rays of different path stages share the same
//     instructions, so a warp stays busy although its lanes are at different stages;
//   * a lane always wants exactly one of three kinds of work: a NODE step (test the four children
//     of a wide node), a TRIANGLE step (one ray/triangle test of its pending leaf), or an ADVANCE
//     step (its ray has ended: shade, spawn the next ray of the path, or take a new job by
//     ballot / prefix-popc from the warp's chunk -- the compaction / refill).  Each iteration the
//     warp votes (two ballots) and runs the step that most lanes want (SCHED_VOTE); the classic
//     while-while loop with a refill threshold is kept as SCHED_WHILE for A/B measurements;
//   * 4-wide nodes (rt_repack.h Node4): one 128-byte line per node, fetched as seven 16-byte
//     read-only loads; the ray octant picks the near/far plane by address, so a child costs six
//     FMAs and two 3-input min/max;
//   * traversal stack: the first kSmStack entries per thread live in shared memory
//     ([entry][thread], conflict-free), deeper entries spill to a local array.
// The ray/triangle test and all shading arithmetic are the shared, no-FMA functions of
// rt_device.cuh; the box test uses FMA on purpose (it is conservative and never reaches an output).
#pragma once

#include "rt_device.cuh"

namespace srt {

#ifndef SRT_SM_STACK
#define SRT_SM_STACK 16
#endif
#ifndef SRT_MIN_BLOCKS
#define SRT_MIN_BLOCKS 3
#endif
constexpr int kWideThreads = 256;
constexpr int kSmStack = SRT_SM_STACK;  // shared-memory stack entries per thread
constexpr int kLocalStack = 80;  // spill entries per thread (local memory)
constexpr int kKeepActive = 20;  // SCHED_WHILE: leave the traversal loop when fewer lanes are busy
constexpr int kAdvanceAt = 16;   // SCHED_VOTE: run the advance step once this many lanes want it
constexpr int kTriAt = 8;        // SCHED_VOTE: run the triangle step once this many lanes want it
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr uint32_t kChunk = 32;  // jobs per atomic fetch

enum { SCHED_VOTE = 0, SCHED_WHILE = 1, SCHED_DEFER = 2 };

struct WideScene {
  const float4* nodes;  // Node4 array, 8 float4 each
  const float4* tris;   // srt_tri_rec_t array, 3 float4 each
  int tune;             // SCHED_VOTE: advance threshold | triangle threshold << 8; SCHED_WHILE: keep-active
};

enum { ST_PRIMARY = 0, ST_SHADOW1 = 1, ST_BOUNCE = 2, ST_SHADOW2 = 3 };

struct RayState {
  V3 o, d, idir, ood;
  float tmin;
  uint32_t ox, oy, oz;  // float4 index of the near plane per axis within a node: {0,1} {2,3} {4,5}
  bool any;             // any-hit ray: stop at the first accepted triangle
  bool found;           // any-hit result
};

__device__ __forceinline__ void ray_setup(RayState& r, V3 o, V3 d, float tmin, bool any) {
  r.o = o;
  r.d = d;
  r.idir = mk(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
  r.ood = mk(o.x * r.idir.x, o.y * r.idir.y, o.z * r.idir.z);
  r.tmin = tmin;
  r.ox = r.idir.x < 0.0f ? 1u : 0u;
  r.oy = r.idir.y < 0.0f ? 3u : 2u;
  r.oz = r.idir.z < 0.0f ? 5u : 4u;
  r.any = any;
  r.found = false;
}

// What a lane is doing inside the scene: a child ref as stored in Node4 -- an inner node index
// (bit 31 clear), one triangle (bit 31 set) -- or kNone when its ray has ended.
struct Cursor {
  uint32_t ref;
  __device__ __forceinline__ bool wants_node() const { return (int32_t)ref >= 0; }
  __device__ __forceinline__ bool wants_tri() const { return (int32_t)ref < 0 && ref != kNone; }
  __device__ __forceinline__ bool busy() const { return ref != kNone; }
  __device__ __forceinline__ void clear() { ref = kNone; }
  __device__ __forceinline__ void go(uint32_t r) { ref = r; }
};

// Shared-memory accesses are written as PTX so that they stay `st.shared` / `ld.shared` with a 32-bit
// address (as a struct member the pointer degraded to generic 64-bit ST.E / LD.E).
__device__ __forceinline__ void sts64(uint32_t addr, uint32_t x, uint32_t y) {
  asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
  return v;
}

// 32-byte read-only load (ld.global.nc.v8.f32 -> LDG.E.256.CONSTANT on sm_100; p must be 32-byte aligned).
struct F8 {
  float4 a, b;
};
__device__ __forceinline__ F8 ldg256(const float4* p) {
  F8 r;
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.a.z), "=f"(r.a.w), "=f"(r.b.x), "=f"(r.b.y), "=f"(r.b.z), "=f"(r.b.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float4 sel4(bool c, const float4 a, const float4 b) {
  return make_float4(c ? a.x : b.x, c ? a.y : b.y, c ? a.z : b.z, c ? a.w : b.w);
}

// One traversal stack per thread: shared entries first ([entry][thread]), local spill above.  The spill
// array is a separate kernel-local array (passed by pointer) so that `top` and `base` stay in registers.
// The stack pointer IS the shared-space byte address of the next free entry (entries of one thread lie
// kStackStride apart): a push is a store and one add, and the depth never has to be turned into an address.
// Addresses at or above base + kSmStack * kStackStride do not exist in shared memory; they stand for the
// entries of the spill array.
constexpr uint32_t kStackStride = kWideThreads * 8u;  // bytes between consecutive entries of one thread
constexpr uint32_t kSmStackBytes = (uint32_t)kSmStack * kStackStride;
struct Stack {
  uint32_t base;  // shared-space byte address of this thread's entry 0
  uint32_t top;   // address of the next free entry
  uint2* spill;   // kLocalStack entries in local memory
  __device__ __forceinline__ void reset() { top = base; }
  __device__ __forceinline__ bool empty() const { return top == base; }
  __device__ __forceinline__ void push_slow(uint32_t ref, float tn) {
    const uint32_t off = top - base;
    if (off < kSmStackBytes) sts64(top, ref, __float_as_uint(tn));
    else if (off < kSmStackBytes + (uint32_t)kLocalStack * kStackStride) spill[(off - kSmStackBytes) / kStackStride] = make_uint2(ref, __float_as_uint(tn));
    else __trap();  // the host checked 3 * depth <= capacity before launching
    top += kStackStride;
  }
  __device__ __forceinline__ uint2 pop() {
    top -= kStackStride;
    const uint32_t off = top - base;
    return off < kSmStackBytes ? lds64(top) : spill[(off - kSmStackBytes) / kStackStride];
  }
};

// Pop until an entry survives the current closest distance; kNone when the stack is empty.
__device__ __forceinline__ uint32_t pop_next(Stack& st, float best_t) {
  const float limit = best_t * kPad;
  while (!st.empty()) {
    const uint2 e = st.pop();
    if (__uint_as_float(e.y) <= limit) return e.x;
  }
  return kNone;
}

// ---- NODE step: test the four children, continue with the nearest, push the others ---------------
// Box test per child: entry = max(near planes, tmin), exit = min(far planes, best_t); the child
// is visited when entry <= exit * kPad.  Boxes were inflated by the scene delta at repack time.
template <bool STATS, bool LD256>
__device__ __forceinline__ void node_step(const WideScene& sc, Cursor& cu, const RayState& r, float best_t, Stack& st,
                                          Stats& stats) {
  // float4 index of the node's first plane, kept in 32 bits (the host refuses images of 2^32 float4 = 64 GiB):
  // every load address is one 32-bit add and one widening multiply-add onto the uniform base pointer.
  const uint32_t n0 = cu.ref * 8u;
  // Two forms of the plane fetch, chosen per kernel (LD256):
  //  * six 16-byte loads, near / far plane picked by ADDRESS from the ray octant: no extra arithmetic, but every load
  //    instruction costs L1 data-pipe wavefronts for each distinct node line among the warp's lanes;
  //  * three 32-byte loads (LDG.E.256, sm_100), one per axis, near / far picked by 24 selects on the ALU pipe.
  // Measured: coherent rays (camera rays, C2) are bound by the ALU pipe and lose 6 % to the
  // selects; the shadow + bounce kernels run the L1 data pipe at 80 % of peak and gain 3-4 % from the wide loads;
  // the HBM-resident random-ray list (C5) loses 12 %.
  float4 nx, fx, ny, fy, nz, fz;
  if (LD256) {
    const F8 px = ldg256(sc.nodes + n0), py = ldg256(sc.nodes + (n0 + 2u)), pz = ldg256(sc.nodes + (n0 + 4u));
    const bool sx = r.ox != 0u, sy = r.oy != 2u, sz = r.oz != 4u;  // near plane = hi plane on that axis
    nx = sel4(sx, px.b, px.a); fx = sel4(sx, px.a, px.b);
    ny = sel4(sy, py.b, py.a); fy = sel4(sy, py.a, py.b);
    nz = sel4(sz, pz.b, pz.a); fz = sel4(sz, pz.a, pz.b);
  } else {
    nx = __ldg(sc.nodes + (n0 + r.ox)); fx = __ldg(sc.nodes + ((n0 + r.ox) ^ 1u));
    ny = __ldg(sc.nodes + (n0 + r.oy)); fy = __ldg(sc.nodes + ((n0 + r.oy) ^ 1u));
    nz = __ldg(sc.nodes + (n0 + r.oz)); fz = __ldg(sc.nodes + ((n0 + r.oz) ^ 1u));
  }
  const float4 rf = __ldg(sc.nodes + (n0 + 6u));
  if (STATS) { stats.nodes += 1; if (step_leader()) stats.wnode += 1; }
#define SRT_CHILD(c, tn, hit)                                                                              \
  const float tn = fmaxf(fmaxf(__fmaf_rn(nx.c, r.idir.x, -r.ood.x), __fmaf_rn(ny.c, r.idir.y, -r.ood.y)), \
                         fmaxf(__fmaf_rn(nz.c, r.idir.z, -r.ood.z), r.tmin));                             \
  const bool hit = tn <= fminf(fminf(__fmaf_rn(fx.c, r.idir.x, -r.ood.x), __fmaf_rn(fy.c, r.idir.y, -r.ood.y)), \
                               fminf(__fmaf_rn(fz.c, r.idir.z, -r.ood.z), best_t)) * kPad;
  SRT_CHILD(x, t0, h0)
  SRT_CHILD(y, t1, h1)
  SRT_CHILD(z, t2, h2)
  SRT_CHILD(w, t3, h3)
#undef SRT_CHILD
  if (!(h0 | h1 | h2 | h3)) {
    cu.go(pop_next(st, best_t));
    return;
  }
  // nearest hit child becomes the next position; the other hit children are pushed with their
  // entry distance (checked again against the then-current closest distance when popped).
  // Written as compare + select chains (no data-dependent branch: the four-way pick used to compile
  // into a branch tree that the lanes of a warp took in different directions).
  const float inf = __int_as_float(0x7f800000);
  const float k0 = h0 ? t0 : inf, k1 = h1 ? t1 : inf, k2 = h2 ? t2 : inf, k3 = h3 ? t3 : inf;
  const float kmin = fminf(fminf(k0, k1), fminf(k2, k3));
  const uint32_t c0 = __float_as_uint(rf.x), c1 = __float_as_uint(rf.y), c2 = __float_as_uint(rf.z),
                 c3 = __float_as_uint(rf.w);
  const bool e0 = k0 == kmin, e1 = !e0 && k1 == kmin, e2 = !e0 && !e1 && k2 == kmin;  // lowest index wins a tie
  uint32_t next = c3;
  next = e2 ? c2 : next;
  next = e1 ? c1 : next;
  next = e0 ? c0 : next;
  const bool p0 = h0 && !e0, p1 = h1 && !e1, p2 = h2 && !e2, p3 = h3 && (e0 || e1 || e2);
  if (st.top - st.base <= kSmStackBytes - 3u * kStackStride) {  // common case: predicated stores, no branches
    if (p3) { sts64(st.top, c3, __float_as_uint(t3)); st.top += kStackStride; }
    if (p2) { sts64(st.top, c2, __float_as_uint(t2)); st.top += kStackStride; }
    if (p1) { sts64(st.top, c1, __float_as_uint(t1)); st.top += kStackStride; }
    if (p0) { sts64(st.top, c0, __float_as_uint(t0)); st.top += kStackStride; }
  } else {
    if (p3) st.push_slow(c3, t3);
    if (p2) st.push_slow(c2, t2);
    if (p1) st.push_slow(c1, t1);
    if (p0) st.push_slow(c0, t0);
  }
  cu.go(next);
}

// ---- TRIANGLE step: the ray/triangle test of the triangle the lane is at ----------------------------
// tri_test: tests triangle `ref`; true when an any-hit ray found its occluder (the ray is over).
template <bool STATS>
__device__ __forceinline__ bool tri_test(const WideScene& sc, uint32_t ref, RayState& r, Hit& best, Stats& stats) {
  const uint32_t rec = ref & 0x7FFFFFFFu;
  if (STATS) { stats.tris += 1; if (step_leader()) stats.wtri += 1; }
  if (r.any) {
    const float4* tp = sc.tris + 3 * (size_t)rec;
    const float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
    float t, u, v;
    if (intersect_tri(r0, r1, r2, r.o, r.d, t, u, v) && t > r.tmin && t < best.t) {
      r.found = true;
      return true;
    }
  } else {
    consider(sc.tris, rec, r.o, r.d, r.tmin, best);
  }
  return false;
}

template <bool STATS>
__device__ __forceinline__ void tri_step(const WideScene& sc, Cursor& cu, RayState& r, Hit& best, Stack& st,
                                         Stats& stats) {
  if (tri_test<STATS>(sc, cu.ref, r, best, stats)) {
    st.reset();
    cu.clear();
    return;
  }
  cu.go(pop_next(st, best.t));
}

// ---- job pool: a warp-uniform window [next, end) of job ids, refilled kChunk at a time ------------
struct Pool {
  uint32_t next, end;  // warp-uniform
  bool dry;            // the global cursor ran past the last job
};

// Hands one job id to every lane with want == true (~0u when none is left).  Warp-convergent.
__device__ __forceinline__ uint32_t take_jobs(Pool& p, bool want, unsigned long long* cursor, uint32_t total_jobs) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t mask = __ballot_sync(0xffffffffu, want);
  uint32_t job = ~0u;
  if (mask == 0u) return job;
  const uint32_t rank = __popc(mask & ((1u << lane) - 1u));
  uint32_t n = __popc(mask);
  const uint32_t have = p.end - p.next;
  if (want && rank < have) job = p.next + rank;
  if (n <= have) {
    p.next += n;
    return job;
  }
  // window exhausted: fetch the next chunk (n - have <= 32 = kChunk, so one fetch is enough)
  n -= have;
  uint32_t base = 0;
  if (!p.dry) {
    if (lane == 0) base = (uint32_t)atomicAdd(cursor, (unsigned long long)kChunk);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= total_jobs) p.dry = true;
  }
  if (p.dry) {
    p.next = p.end = 0;
    return job;
  }
  p.next = base;
  p.end = min(base + kChunk, total_jobs);
  if (want && rank >= have) {
    const uint32_t j = base + (rank - have);
    if (j < p.end) job = j;
  }
  p.next = min(p.next + n, p.end);
  return job;
}

// ---- the warp scheduler ----------------------------------------------------------------------------
// advance(): warp-convergent; for every lane whose ray has ended it finishes / continues the lane's
// job, then hands new jobs to lanes without one (sets have_job, cu, r, best).
template <int SCHED, bool STATS, bool LD256, class Advance>
__device__ __forceinline__ void run_warp(const WideScene& sc, Cursor& cu, RayState& r, Hit& best, Stack& st,
                                         Stats& stats, const Pool& pool, const bool& have_job, Advance advance) {
  if (SCHED == SCHED_WHILE) {
#ifdef SRT_SPEC
    // while-while with up to SRT_SPEC (1..3) postponed triangles per lane (speculative traversal): a lane that reaches
    // a triangle while others of its warp still walk nodes parks it and walks on; it leaves the node loop when its
    // parking slots are full and it meets another triangle, or when it runs out of nodes.  Triangles come in bursts
    // (a leaf-level Node4 holds up to four single-triangle children), so one slot does not keep a lane in the loop.
    // The triangle phase tests the parked triangles slot by slot, each a warp-wide step.  Test order cannot change a
    // result (closest hit = minimum over all accepted triangles).
    uint32_t pend0 = kNone, pend1 = kNone, pend2 = kNone;
    int np = 0;
    for (;;) {
      advance();
      if (pool.dry && !__any_sync(0xffffffffu, have_job)) return;
      for (;;) {
        while (cu.wants_node()) {
          node_step<STATS, LD256>(sc, cu, r, best.t, st, stats);
          while (np < SRT_SPEC && cu.wants_tri()) {
            if (np == 0) pend0 = cu.ref; else if (np == 1) pend1 = cu.ref; else pend2 = cu.ref;
            ++np;
            cu.go(pop_next(st, best.t));
          }
        }
#pragma unroll
        for (int k = 0; k < SRT_SPEC; ++k) {
          if (k < np) {
            const uint32_t t = k == 0 ? pend0 : (k == 1 ? pend1 : pend2);
            if (tri_test<STATS>(sc, t, r, best, stats)) {  // an any-hit ray found its occluder
              st.reset();
              cu.clear();
              np = 0;
            }
          }
        }
        np = 0;
        while (cu.wants_tri()) tri_step<STATS>(sc, cu, r, best, st, stats);
        if (__popc(__ballot_sync(0xffffffffu, cu.busy())) < sc.tune) break;
      }
    }
#else
    for (;;) {
      advance();
      if (pool.dry && !__any_sync(0xffffffffu, have_job)) return;
      for (;;) {
        while (cu.wants_node()) node_step<STATS, LD256>(sc, cu, r, best.t, st, stats);
        while (cu.wants_tri()) tri_step<STATS>(sc, cu, r, best, st, stats);
        if (__popc(__ballot_sync(0xffffffffu, cu.busy())) < sc.tune) break;
      }
    }
#endif
  } else if (SCHED == SCHED_DEFER) {
    // Vote scheduler with one deferred triangle per lane.  A lane that reaches a triangle while the warp is busy
    // with node steps parks it in `pend` and goes on with the next entry of its stack instead of idling; the
    // triangle step runs once `tri_at` lanes hold a triangle (or no lane wants a node).  Closest hit is the
    // minimum over every accepted triangle, so the order of the tests cannot change a result -- a late test only
    // means the closest distance shrinks later and a few more nodes are visited.
    uint32_t pend = kNone;
    const int adv_at = sc.tune & 0xff, tri_at = (sc.tune >> 8) & 0xff;
    for (;;) {
      if (cu.wants_tri() && pend == kNone && !st.empty()) {
        pend = cu.ref;
        cu.go(pop_next(st, best.t));
      }
      if (!cu.busy() && pend != kNone) { cu.go(pend); pend = kNone; }
      const bool wn = cu.wants_node(), wt = pend != kNone || cu.wants_tri();
      const bool wa = !cu.busy() && (have_job || !pool.dry);
      const uint32_t mn = __ballot_sync(0xffffffffu, wn), mt = __ballot_sync(0xffffffffu, wt);
      const uint32_t ma = __ballot_sync(0xffffffffu, wa);
      if ((mn | mt | ma) == 0u) return;
      if (__popc(ma) >= adv_at || (mn | mt) == 0u) {
        advance();
      } else if (__popc(mt) >= tri_at || mn == 0u) {
        if (wt) {
          const bool own = pend == kNone;
          if (tri_test<STATS>(sc, own ? cu.ref : pend, r, best, stats)) {
            st.reset();
            cu.clear();
          } else if (own) {
            cu.go(pop_next(st, best.t));
          }
          pend = kNone;
        }
      } else {
        if (wn) node_step<STATS, LD256>(sc, cu, r, best.t, st, stats);
      }
    }
  } else {
    for (;;) {
      const bool wn = cu.wants_node(), wt = cu.wants_tri();
      const bool wa = !wn && !wt && (have_job || !pool.dry);
      const uint32_t mn = __ballot_sync(0xffffffffu, wn), mt = __ballot_sync(0xffffffffu, wt);
      const uint32_t ma = __ballot_sync(0xffffffffu, wa);
      if ((mn | mt | ma) == 0u) return;
      // minority work (triangle tests, path advance) is batched until enough lanes want it
      if (__popc(ma) >= (sc.tune & 0xff) || (mn | mt) == 0u) {
        advance();
      } else if (__popc(mt) >= ((sc.tune >> 8) & 0xff) || mn == 0u) {
        if (wt) tri_step<STATS>(sc, cu, r, best, st, stats);
      } else {
        if (wn) node_step<STATS, LD256>(sc, cu, r, best.t, st, stats);
      }
    }
  }
}

// ---- L2 warm-up: one prefetch per 128-byte line of the scene (used when the scene fits in L2, so
// that first touches inside the traversal kernels cost an L2 hit instead of an HBM round trip) ----
__global__ void __launch_bounds__(256) rt_prefetch_l2_kernel(const char* a, size_t lines_a, const char* b, size_t lines_b) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < lines_a + lines_b; i += stride) {
    const char* p = i < lines_a ? a + i * 128 : b + (i - lines_a) * 128;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
  }
}

// ---- render kernel: every lane carries one pixel through all samples and path stages --------------
struct PathState {
  uint32_t x, ly, pix;  // pixel, local (strip) row, frame pixel index (RNG key)
  uint32_t s, stage;    // sample index, ST_*
  V3 sum;               // radiance summed over finished samples
  V3 n;                 // oriented normal at the primary hit (kept for the bounce direction)
  float ndl;            // n.l at the hit whose shadow ray is in flight
  float k1;             // direct term at the primary hit
  uint32_t alb1, alb2;  // packed albedo at the primary / bounce hit
};

// SPLIT false: a job is a pixel, its samples run one after the other in the lane and the pixel is written here.
// SPLIT true (multi-sample frames): a job is one (pixel, sample); its radiance goes to samples[s * pixels + pixel]
// and rt_resolve_kernel adds the samples in order -- the same additions in the same order, spread over spp x as
// many, spp x shorter jobs whose 32-job windows hold one sample index of one 8x4 tile: a shorter tail for the
// persistent kernel and lanes that stay in step (C3 at 4 spp: 6.9 ms instead of 8.1).
template <int SCHED, bool STATS, bool SPLIT>
__device__ __forceinline__ void wide_render_body(const kernel_arg_t& a, const WideScene& sc, uint32_t local_rows,
                                                 uint32_t total_jobs, unsigned long long* counters,
                                                 float4* __restrict__ samples) {
  constexpr bool split = SPLIT;
  extern __shared__ uint2 s_stack[];
  uint2 spill[kLocalStack];
  Stack st;
  st.base = (uint32_t)__cvta_generic_to_shared(s_stack + threadIdx.x);
  st.spill = spill;
  st.reset();
  const uint32_t W = a.dst_width, tiles_x = (W + 7u) / 8u;
  const uint32_t seed_hash = mix32(a.seed);
  const Camera cam = make_camera(a);
  Pool pool = {0u, 0u, false};
  PathState ps = {};
  RayState r;
  Hit best;
  Stats stats = {0u, 0u, 0u, 0u, 0u};
  Cursor cu;
  uint32_t n_shadow = 0u, n_bounce = 0u;
  bool have_job = false;
  cu.clear();
  best.t = 0.0f; best.u = 0.0f; best.v = 0.0f; best.prim = SRT_MISS; best.rec = 0u;
  ray_setup(r, cam.eye, cam.fwd, 0.0f, false);

  auto start_ray = [&](V3 o, V3 d, float tmax, bool any) {
    if (STATS) stats.rays += 1;
    ray_setup(r, o, d, 0.0f, any);
    best.t = tmax; best.prim = SRT_MISS; best.rec = 0u;
    st.reset();
    cu.go(0u);  // root
  };
  auto start_primary = [&]() {
    float jx = 0.5f, jy = 0.5f;
    if (a.spp > 1) { jx = rnd01(seed_hash, ps.pix, ps.s, 0); jy = rnd01(seed_hash, ps.pix, ps.s, 1); }
    ps.stage = ST_PRIMARY;
    start_ray(cam.eye, camera_dir(cam, ps.x, global_row(a, ps.ly), jx, jy), kTMax, false);
  };

  auto advance = [&]() {
    // ---- lanes whose ray has ended advance their path ----------------------------------------
    if (have_job && !cu.busy()) {
      V3 rad = ld3(a.sky);
      bool sample_done = false;
      if (ps.stage == ST_PRIMARY || ps.stage == ST_BOUNCE) {
        const bool bounce = ps.stage == ST_BOUNCE;
        if (best.prim == SRT_MISS) {
          if (bounce) {  // c2 = sky
            const V3 alb = unpack_albedo(ps.alb1), c2 = ld3(a.sky);
            rad = alb * ps.k1;
            rad.x = rad.x + (alb.x * 0.5f) * c2.x;
            rad.y = rad.y + (alb.y * 0.5f) * c2.y;
            rad.z = rad.z + (alb.z * 0.5f) * c2.z;
          }
          sample_done = true;
        } else {
          const float4 r1 = __ldg(sc.tris + 3 * (size_t)best.rec + 1), r2 = __ldg(sc.tris + 3 * (size_t)best.rec + 2);
          const Surface sf = surface_at(a, r1, r2, r.o, r.d, best.t);
          if (bounce) ps.alb2 = __float_as_uint(r1.w);
          else { ps.alb1 = __float_as_uint(r1.w); ps.n = sf.n; }
          ps.ndl = sf.ndl;
          ps.stage = bounce ? ST_SHADOW2 : ST_SHADOW1;
          if (sf.ndl > 0.0f && a.mode >= SRT_MODE_SHADOW) {
            ++n_shadow;
            start_ray(sf.po, sf.ld, sf.dist, true);
          } else {
            // no shadow ray: continue below as "shadow ray ended, light visible"
            r.o = sf.po;
            r.found = false;
            if (!(sf.ndl > 0.0f)) ps.ndl = 0.0f;  // direct term = ambient exactly
          }
        }
      }
      if (!sample_done && !cu.busy() && (ps.stage == ST_SHADOW1 || ps.stage == ST_SHADOW2)) {
        // direct term: ambient + n.l * visibility (visibility is 0 or 1, so the product is exact)
        const float k = (ps.ndl > 0.0f) ? (r.found ? a.ambient : a.ambient + ps.ndl) : a.ambient;
        if (ps.stage == ST_SHADOW1) {
          ps.k1 = k;
          if (a.mode >= SRT_MODE_BOUNCE) {
            ++n_bounce;
            ps.stage = ST_BOUNCE;
            start_ray(r.o, bounce_dir(seed_hash, ps.pix, ps.s, ps.n), kTMax, false);
          } else {
            rad = unpack_albedo(ps.alb1) * k;
            sample_done = true;
          }
        } else {
          const V3 alb = unpack_albedo(ps.alb1), c2 = unpack_albedo(ps.alb2) * k;
          rad = alb * ps.k1;
          rad.x = rad.x + (alb.x * 0.5f) * c2.x;
          rad.y = rad.y + (alb.y * 0.5f) * c2.y;
          rad.z = rad.z + (alb.z * 0.5f) * c2.z;
          sample_done = true;
        }
      }
      if (sample_done && split) {
        samples[(size_t)ps.s * ((size_t)local_rows * W) + (size_t)ps.ly * W + ps.x] = make_float4(rad.x, rad.y, rad.z, 0.0f);
        have_job = false;
      } else if (sample_done) {
        ps.sum = ps.sum + rad;
        if (++ps.s < a.spp) {
          start_primary();
        } else {
          const V3 c = ps.sum / (float)a.spp;
          *reinterpret_cast<uint32_t*>(a.cbuf_addr + (uint64_t)ps.x * a.cbuf_stride + (uint64_t)ps.ly * a.cbuf_pitch) =
              pack_argb(c);
          have_job = false;
        }
      }
    }
    // ---- lanes without a pixel take one from the warp's window (ballot + prefix popc) ---------
    const uint32_t job = take_jobs(pool, !have_job, counters + CTR_CURSOR, total_jobs);
    if (job != ~0u) {
      // 32 consecutive jobs = one 8x4 pixel tile (and, when split, one sample of it)
      const uint32_t group = job >> 5, l = job & 31u;
      const uint32_t tile = split ? group / a.spp : group;
      ps.x = (tile % tiles_x) * 8u + (l & 7u);
      ps.ly = (tile / tiles_x) * 4u + (l >> 3);
      if (ps.x < W && ps.ly < local_rows) {
        ps.pix = global_row(a, ps.ly) * W + ps.x;
        ps.s = split ? group - tile * a.spp : 0u;
        ps.sum = mk(0.0f, 0.0f, 0.0f);
        have_job = true;
        start_primary();
      }
    }
  };

  run_warp<SCHED, STATS, false>(sc, cu, r, best, st, stats, pool, have_job, advance);

  n_shadow = __reduce_add_sync(0xffffffffu, n_shadow);
  n_bounce = __reduce_add_sync(0xffffffffu, n_bounce);
  if ((threadIdx.x & 31u) == 0u) {
    if (n_shadow) atomicAdd(counters + CTR_RAYS_SHADOW, (unsigned long long)n_shadow);
    if (n_bounce) atomicAdd(counters + CTR_RAYS_BOUNCE, (unsigned long long)n_bounce);
  }
  if (STATS) flush_stats(stats, counters);
}

template <int SCHED, bool STATS>
__global__ void __launch_bounds__(kWideThreads, SRT_MIN_BLOCKS)
rt_wide_render_kernel(const __grid_constant__ kernel_arg_t a, const WideScene sc, uint32_t local_rows,
                      uint32_t total_jobs, unsigned long long* counters, float4* samples) {
  wide_render_body<SCHED, STATS, false>(a, sc, local_rows, total_jobs, counters, samples);
}

template <int SCHED, bool STATS>
__global__ void __launch_bounds__(kWideThreads, SRT_MIN_BLOCKS)
rt_wide_render_split_kernel(const __grid_constant__ kernel_arg_t a, const WideScene sc, uint32_t local_rows,
                            uint32_t total_jobs, unsigned long long* counters, float4* samples) {
  wide_render_body<SCHED, STATS, true>(a, sc, local_rows, total_jobs, counters, samples);
}

// ---- resolve of a split multi-sample frame: the pixel loop's additions, in its order ---------------------
__global__ void __launch_bounds__(256) rt_resolve_kernel(const __grid_constant__ kernel_arg_t a, uint32_t local_rows,
                                                         const float4* __restrict__ samples) {
  const uint32_t W = a.dst_width, x = blockIdx.x * 64u + (threadIdx.x & 63u), ly = blockIdx.y * 4u + (threadIdx.x >> 6);
  if (x >= W || ly >= local_rows) return;
  const size_t pixels = (size_t)local_rows * W, p = (size_t)ly * W + x;
  V3 sum = mk(0.0f, 0.0f, 0.0f);
  for (uint32_t s = 0; s < a.spp; ++s) {
    const float4 r = __ldcs(samples + s * pixels + p);
    sum = sum + mk(r.x, r.y, r.z);
  }
  const V3 c = sum / (float)a.spp;
  *reinterpret_cast<uint32_t*>(a.cbuf_addr + (uint64_t)x * a.cbuf_stride + (uint64_t)ly * a.cbuf_pitch) = pack_argb(c);
}

// ---- explicit ray list (mode 3): job = ray index, closest hit, 16-byte hit record out -------------
template <int SCHED, bool STATS>
__global__ void __launch_bounds__(kWideThreads, SRT_MIN_BLOCKS)
rt_wide_rays_kernel(const float4* __restrict__ rays, float4* __restrict__ hits, const WideScene sc,
                    uint32_t total_jobs, unsigned long long* counters) {
  extern __shared__ uint2 s_stack[];
  uint2 spill[kLocalStack];
  Stack st;
  st.base = (uint32_t)__cvta_generic_to_shared(s_stack + threadIdx.x);
  st.spill = spill;
  st.reset();
  Pool pool = {0u, 0u, false};
  RayState r;
  Hit best;
  Stats stats = {0u, 0u, 0u, 0u, 0u};
  Cursor cu;
  uint32_t my_ray = 0u;
  bool have_job = false;
  cu.clear();
  best.t = 0.0f; best.u = 0.0f; best.v = 0.0f; best.prim = SRT_MISS; best.rec = 0u;
  ray_setup(r, mk(0.0f, 0.0f, 0.0f), mk(0.0f, 0.0f, 1.0f), 0.0f, false);

  auto advance = [&]() {
    if (have_job && !cu.busy()) {
      hits[my_ray] = make_float4(best.t, best.u, best.v, __uint_as_float(best.prim));
      have_job = false;
    }
    const uint32_t job = take_jobs(pool, !have_job, counters + CTR_CURSOR, total_jobs);
    if (job != ~0u) {
      const float4 r0 = __ldg(rays + 2 * (size_t)job), r1 = __ldg(rays + 2 * (size_t)job + 1);
      my_ray = job;
      have_job = true;
      if (STATS) stats.rays += 1;
      ray_setup(r, mk(r0.x, r0.y, r0.z), mk(r1.x, r1.y, r1.z), r0.w, false);
      best.t = r1.w; best.u = 0.0f; best.v = 0.0f; best.prim = SRT_MISS; best.rec = 0u;
      st.reset();
      cu.go(0u);
    }
  };

  run_warp<SCHED, STATS, false>(sc, cu, r, best, st, stats, pool, have_job, advance);  // HBM-resident lists: 16-byte loads
  if (STATS) flush_stats(stats, counters);
}

}  // namespace srt
