// rt_pool.cuh -- ray-pool traversal core: the state of 64 rays per warp lives in shared memory and,
// every iteration, the warp hands the 32 lanes to (up to) 32 rays that all want the SAME kind of step.
//
// Why (DESIGN.md "Warp execution efficiency"): with one ray per lane a warp's lanes want different
// things at any moment -- a node step, a triangle test, or the end-of-ray work -- and whatever the
// loop structure (if-if, while-while, majority vote) the minority lanes idle; ncu showed 9-14
// active threads per instruction for those kernels.  Decoupling rays from lanes turns that
// divergence into a scheduling problem: rays wanting a rare step simply wait in the pool (costing
// no issue slots) while the lanes keep running full node steps on other rays, and the rare steps
// run once a full warp of them has piled up.  This is the north star's "warp-level ballot/shfl ray
// compaction and refill", applied at every step instead of only at ray termination.
//
// Cost: a step loads its ray's state from shared memory (9-10 LDS) and stores back what changed.
#pragma once

#include "rt_wide.cuh"

namespace srt {

constexpr int kPoolThreads = 128;            // 4 warps per CTA, each with its own pool
constexpr int kPoolSlots = 64;               // rays per warp
#ifndef SRT_POOL_STACK
#define SRT_POOL_STACK 8
#endif
constexpr int kPoolStack = SRT_POOL_STACK;   // shared-memory stack entries per ray
constexpr int kPoolSpill = 80;               // further entries per ray, in a global scratch buffer
constexpr uint32_t kEmpty = 0xFFFFFFFEu;     // the slot holds no ray (and never will again)

// per-ray words in shared memory, struct-of-arrays: f[field][slot]
enum {
  F_OX, F_OY, F_OZ, F_DX, F_DY, F_DZ, F_IX, F_IY, F_IZ,  // o, d, 1/d
  F_TMIN, F_BEST_T,
  F_BEST_REC,  // triangle record of the closest hit so far, SRT_MISS if none (prim id, u, v are
               // re-derived from it when needed: they are a pure function of (ray, record))
  F_CUR,   // Node4 child ref the ray is at; kNone = ray ended (result not yet collected); kEmpty
  F_SP,    // stack height
  F_FLAGS, // bit 0: any-hit ray, bit 1: any-hit found
  F_JOB,   // job id (ray index / queue entry) the slot is working on
  F_COUNT
};

struct WarpPool {
  uint32_t f[F_COUNT][kPoolSlots];
  uint2 stack[kPoolStack][kPoolSlots];
  uint2 pick[32];  // this iteration's lane -> (slot, cur) assignment
};

struct PoolCtx {
  WarpPool* wp;
  uint2* spill;  // this warp's spill area: [slot][kPoolSpill]
};

__device__ __forceinline__ uint32_t pool_pop(const PoolCtx& c, uint32_t slot, int& sp, float best_t) {
  while (sp > 0) {
    --sp;
    const uint2 e = sp < kPoolStack ? c.wp->stack[sp][slot] : c.spill[slot * kPoolSpill + (sp - kPoolStack)];
    if (__uint_as_float(e.y) <= best_t * kPad) return e.x;
  }
  return kNone;
}

__device__ __forceinline__ void pool_push(const PoolCtx& c, uint32_t slot, int& sp, uint32_t ref, float tn) {
  const uint2 e = make_uint2(ref, __float_as_uint(tn));
  if (sp < kPoolStack) c.wp->stack[sp][slot] = e;
  else if (sp < kPoolStack + kPoolSpill) c.spill[slot * kPoolSpill + (sp - kPoolStack)] = e;
  else __trap();  // the host checked 3 * depth <= capacity before launching
  ++sp;
}

// ---- NODE step on the ray in `slot` (currently at inner node `cur`) ----------------------------------
template <bool STATS>
__device__ __forceinline__ void pool_node_step(const WideScene& sc, const PoolCtx& c, uint32_t slot, uint32_t cur,
                                               Stats& stats) {
  WarpPool& wp = *c.wp;
  const float ix = __uint_as_float(wp.f[F_IX][slot]), iy = __uint_as_float(wp.f[F_IY][slot]),
              iz = __uint_as_float(wp.f[F_IZ][slot]);
  const float qx = __uint_as_float(wp.f[F_OX][slot]) * ix, qy = __uint_as_float(wp.f[F_OY][slot]) * iy,
              qz = __uint_as_float(wp.f[F_OZ][slot]) * iz;
  const float tmin = __uint_as_float(wp.f[F_TMIN][slot]), best_t = __uint_as_float(wp.f[F_BEST_T][slot]);
  int sp = (int)wp.f[F_SP][slot];
  const uint32_t ox = ix < 0.0f ? 1u : 0u, oy = iy < 0.0f ? 3u : 2u, oz = iz < 0.0f ? 5u : 4u;
  const float4* np = sc.nodes + (size_t)cur * 8;
  const float4 nx = __ldg(np + ox), fx = __ldg(np + (ox ^ 1u));
  const float4 ny = __ldg(np + oy), fy = __ldg(np + (oy ^ 1u));
  const float4 nz = __ldg(np + oz), fz = __ldg(np + (oz ^ 1u));
  const float4 rf = __ldg(np + 6);
  if (STATS) stats.nodes += 1;
#define SRT_CHILD(c, tn, hit)                                                                   \
  const float tn = fmaxf(fmaxf(__fmaf_rn(nx.c, ix, -qx), __fmaf_rn(ny.c, iy, -qy)),            \
                         fmaxf(__fmaf_rn(nz.c, iz, -qz), tmin));                                \
  const bool hit = tn <= fminf(fminf(__fmaf_rn(fx.c, ix, -qx), __fmaf_rn(fy.c, iy, -qy)),      \
                               fminf(__fmaf_rn(fz.c, iz, -qz), best_t)) * kPad;
  SRT_CHILD(x, t0, h0)
  SRT_CHILD(y, t1, h1)
  SRT_CHILD(z, t2, h2)
  SRT_CHILD(w, t3, h3)
#undef SRT_CHILD
  uint32_t next;
  if (!(h0 | h1 | h2 | h3)) {
    next = pool_pop(c, slot, sp, best_t);
  } else {
    const float inf = __int_as_float(0x7f800000);
    const float k0 = h0 ? t0 : inf, k1 = h1 ? t1 : inf, k2 = h2 ? t2 : inf, k3 = h3 ? t3 : inf;
    const float kmin = fminf(fminf(k0, k1), fminf(k2, k3));
    const int sel = k0 == kmin ? 0 : (k1 == kmin ? 1 : (k2 == kmin ? 2 : 3));
    const uint32_t c0 = __float_as_uint(rf.x), c1 = __float_as_uint(rf.y), c2 = __float_as_uint(rf.z),
                   c3 = __float_as_uint(rf.w);
    next = sel == 0 ? c0 : (sel == 1 ? c1 : (sel == 2 ? c2 : c3));
    const bool p0 = h0 && sel != 0, p1 = h1 && sel != 1, p2 = h2 && sel != 2, p3 = h3 && sel != 3;
    if (sp + 3 <= kPoolStack) {  // common case: predicated stores, no branches
      if (p3) { wp.stack[sp][slot] = make_uint2(c3, __float_as_uint(t3)); ++sp; }
      if (p2) { wp.stack[sp][slot] = make_uint2(c2, __float_as_uint(t2)); ++sp; }
      if (p1) { wp.stack[sp][slot] = make_uint2(c1, __float_as_uint(t1)); ++sp; }
      if (p0) { wp.stack[sp][slot] = make_uint2(c0, __float_as_uint(t0)); ++sp; }
    } else {
      if (p3) pool_push(c, slot, sp, c3, t3);
      if (p2) pool_push(c, slot, sp, c2, t2);
      if (p1) pool_push(c, slot, sp, c1, t1);
      if (p0) pool_push(c, slot, sp, c0, t0);
    }
  }
  wp.f[F_CUR][slot] = next;
  wp.f[F_SP][slot] = (uint32_t)sp;
}

// ---- TRIANGLE step on the ray in `slot` (currently at triangle ref `cur`) ---------------------------
template <bool STATS>
__device__ __forceinline__ void pool_tri_step(const WideScene& sc, const PoolCtx& c, uint32_t slot, uint32_t cur,
                                              Stats& stats) {
  WarpPool& wp = *c.wp;
  const uint32_t rec = cur & 0x7FFFFFFFu;
  const V3 o = mk(__uint_as_float(wp.f[F_OX][slot]), __uint_as_float(wp.f[F_OY][slot]), __uint_as_float(wp.f[F_OZ][slot]));
  const V3 d = mk(__uint_as_float(wp.f[F_DX][slot]), __uint_as_float(wp.f[F_DY][slot]), __uint_as_float(wp.f[F_DZ][slot]));
  const float tmin = __uint_as_float(wp.f[F_TMIN][slot]);
  float best_t = __uint_as_float(wp.f[F_BEST_T][slot]);
  const uint32_t flags = wp.f[F_FLAGS][slot];
  int sp = (int)wp.f[F_SP][slot];
  if (STATS) stats.tris += 1;
  const float4* tp = sc.tris + 3 * (size_t)rec;
  const float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
  float t, u, v;
  uint32_t next;
  if (intersect_tri(r0, r1, r2, o, d, t, u, v) && t > tmin) {
    if (flags & 1u) {  // any-hit: the first accepted triangle inside (tmin, tmax) ends the ray
      if (t < best_t) {
        wp.f[F_FLAGS][slot] = flags | 2u;
        wp.f[F_CUR][slot] = kNone;
        wp.f[F_SP][slot] = 0u;
        return;
      }
    } else {
      bool take = t < best_t;
      if (t == best_t) {  // tie: lowest prim id wins (rare: the previous winner's id is fetched again)
        const uint32_t best_rec = wp.f[F_BEST_REC][slot];
        take = best_rec != SRT_MISS && __float_as_uint(r0.w) < __float_as_uint(__ldg(sc.tris + 3 * (size_t)best_rec).w);
      }
      if (take) {
        best_t = t;
        wp.f[F_BEST_T][slot] = __float_as_uint(t);
        wp.f[F_BEST_REC][slot] = rec;
      }
    }
  }
  next = pool_pop(c, slot, sp, best_t);
  wp.f[F_CUR][slot] = next;
  wp.f[F_SP][slot] = (uint32_t)sp;
}

// store a fresh ray into `slot`
__device__ __forceinline__ void pool_put_ray(WarpPool& wp, uint32_t slot, V3 o, V3 d, float tmin, float tmax, bool any,
                                             uint32_t job) {
  const V3 id = mk(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
  wp.f[F_OX][slot] = __float_as_uint(o.x); wp.f[F_OY][slot] = __float_as_uint(o.y); wp.f[F_OZ][slot] = __float_as_uint(o.z);
  wp.f[F_DX][slot] = __float_as_uint(d.x); wp.f[F_DY][slot] = __float_as_uint(d.y); wp.f[F_DZ][slot] = __float_as_uint(d.z);
  wp.f[F_IX][slot] = __float_as_uint(id.x); wp.f[F_IY][slot] = __float_as_uint(id.y); wp.f[F_IZ][slot] = __float_as_uint(id.z);
  wp.f[F_TMIN][slot] = __float_as_uint(tmin);
  wp.f[F_BEST_T][slot] = __float_as_uint(tmax);
  wp.f[F_BEST_REC][slot] = SRT_MISS;
  wp.f[F_SP][slot] = 0u;
  wp.f[F_FLAGS][slot] = any ? 1u : 0u;
  wp.f[F_JOB][slot] = job;
  wp.f[F_CUR][slot] = 0u;  // root
}

constexpr uint32_t kVacant = 0xFFFFFFFDu;  // the slot holds no ray but may be given one
constexpr int kPoolFetchAt = 16;           // run the fetch step once this many slots want it
constexpr int kPoolTriAt = 24;             // run the triangle step once this many rays want it

// ---- the pool scheduler -----------------------------------------------------------------------------
// collect(slot): the ray in `slot` has ended -- consume its result (hit record, path update).
// spawn(slot, job): start job `job` in `slot`; returns false if the job produced no ray.
// Both run inside the fetch step on up to 32 lanes at once.
template <bool STATS, class Collect, class Spawn>
__device__ __forceinline__ void pool_run(const WideScene& sc, const PoolCtx& c, uint32_t total_jobs,
                                         unsigned long long* cursor, Stats& stats, Collect collect, Spawn spawn) {
  WarpPool& wp = *c.wp;
  const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  const int fetch_at = sc.tune & 0xff, tri_at = (sc.tune >> 8) & 0xff;
  Pool pool = {0u, 0u, false};
  wp.f[F_CUR][lane] = kVacant;
  wp.f[F_CUR][lane + 32] = kVacant;
  __syncwarp();
  for (;;) {
    const uint32_t ca = wp.f[F_CUR][lane], cb = wp.f[F_CUR][lane + 32];
    const bool na = (int32_t)ca >= 0, nb = (int32_t)cb >= 0;
    const bool ta = (int32_t)ca < 0 && ca < kVacant, tb = (int32_t)cb < 0 && cb < kVacant;
    const bool fa = ca == kNone || ca == kVacant, fb = cb == kNone || cb == kVacant;
    const uint32_t mna = __ballot_sync(0xffffffffu, na), mnb = __ballot_sync(0xffffffffu, nb);
    const uint32_t mta = __ballot_sync(0xffffffffu, ta), mtb = __ballot_sync(0xffffffffu, tb);
    const uint32_t mfa = __ballot_sync(0xffffffffu, fa), mfb = __ballot_sync(0xffffffffu, fb);
    const int cn = __popc(mna) + __popc(mnb), ct = __popc(mta) + __popc(mtb), cf = __popc(mfa) + __popc(mfb);
    if (cn + ct + cf == 0) break;
    int kind;  // 0 node, 1 triangle, 2 fetch
    uint32_t ma, mb;
    if (cf >= fetch_at || cn + ct == 0) { kind = 2; ma = mfa; mb = mfb; }
    else if (ct >= tri_at || cn == 0) { kind = 1; ma = mta; mb = mtb; }
    else { kind = 0; ma = mna; mb = mnb; }
    // compaction: the k-th ray of the chosen kind goes to lane k
    const uint32_t ra = __popc(ma & lt), rb = __popc(ma) + __popc(mb & lt);
    if (((ma >> lane) & 1u) && ra < 32u) wp.pick[ra] = make_uint2(lane, ca);
    if (((mb >> lane) & 1u) && rb < 32u) wp.pick[rb] = make_uint2(lane + 32u, cb);
    __syncwarp();
    const uint32_t n = min(32u, (uint32_t)(__popc(ma) + __popc(mb)));
    const bool mine = lane < n;
    const uint2 pk = mine ? wp.pick[lane] : make_uint2(0u, kEmpty);
    if (kind == 0) {
      if (mine) pool_node_step<STATS>(sc, c, pk.x, pk.y, stats);
    } else if (kind == 1) {
      if (mine) pool_tri_step<STATS>(sc, c, pk.x, pk.y, stats);
    } else {
      if (mine && pk.y == kNone) collect(pk.x);
      const uint32_t job = take_jobs(pool, mine, cursor, total_jobs);
      if (mine) {
        if (job == ~0u) wp.f[F_CUR][pk.x] = pool.dry ? kEmpty : kVacant;
        else if (!spawn(pk.x, job)) wp.f[F_CUR][pk.x] = kVacant;
      }
    }
    __syncwarp();
  }
}

// ---- explicit ray list (mode 3) on the pool core -------------------------------------------------------
template <bool STATS>
__global__ void __launch_bounds__(kPoolThreads)
rt_pool_rays_kernel(const float4* __restrict__ rays, float4* __restrict__ hits, const WideScene sc, uint2* spill,
                    uint32_t total_jobs, unsigned long long* counters) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  WarpPool* pools = reinterpret_cast<WarpPool*>(s_raw);
  const uint32_t warp_in_cta = threadIdx.x >> 5;
  const size_t warp_gid = (size_t)blockIdx.x * (kPoolThreads / 32) + warp_in_cta;
  PoolCtx c = {pools + warp_in_cta, spill + warp_gid * (size_t)(kPoolSlots * kPoolSpill)};
  Stats stats = {0u, 0u, 0u, 0u, 0u};
  WarpPool& wp = *c.wp;
  auto collect = [&](uint32_t slot) {
    const uint32_t rec = wp.f[F_BEST_REC][slot];
    float4 h = make_float4(__uint_as_float(wp.f[F_BEST_T][slot]), 0.0f, 0.0f, __uint_as_float(SRT_MISS));
    if (rec != SRT_MISS) {  // u, v, prim id of the winning triangle: the same test on the same operands
      const float4* tp = sc.tris + 3 * (size_t)rec;
      const float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
      const V3 o = mk(__uint_as_float(wp.f[F_OX][slot]), __uint_as_float(wp.f[F_OY][slot]), __uint_as_float(wp.f[F_OZ][slot]));
      const V3 d = mk(__uint_as_float(wp.f[F_DX][slot]), __uint_as_float(wp.f[F_DY][slot]), __uint_as_float(wp.f[F_DZ][slot]));
      float t;
      intersect_tri(r0, r1, r2, o, d, t, h.y, h.z);
      h.w = r0.w;
    }
    hits[wp.f[F_JOB][slot]] = h;
  };
  auto spawn = [&](uint32_t slot, uint32_t job) -> bool {
    const float4 r0 = __ldg(rays + 2 * (size_t)job), r1 = __ldg(rays + 2 * (size_t)job + 1);
    if (STATS) stats.rays += 1;
    pool_put_ray(wp, slot, mk(r0.x, r0.y, r0.z), mk(r1.x, r1.y, r1.z), r0.w, r1.w, false, job);
    return true;
  };
  pool_run<STATS>(sc, c, total_jobs, counters + CTR_CURSOR, stats, collect, spawn);
  if (STATS) flush_stats(stats, counters);
}

}  // namespace srt
