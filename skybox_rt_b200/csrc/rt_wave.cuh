// rt_wave.cuh -- the wavefront form of the RT path: traversal kernels and shading passes as separate launches.
//
// north_star: "Traversal and intersection are hand-written sm_100a kernels using persistent threads ... and
// warp-level ballot/shfl ray compaction and refill to fight divergence, with a coalesced shading and
// framebuffer-write pass."  Round 1 ran the whole path (primary -> shadow -> bounce -> bounce shadow, with the
// shading in between) inside ONE persistent kernel; ncu showed its shading code running at 2-5 of 32 lanes and
// its 80 registers holding occupancy at 24 warps per SM (profiles/r01_tri_vote_*).  Here every stage is its own
// launch on the device stream, chained through ray queues in HBM / L2, with no host round trip:
//
//   K1 trace   <camera rays>            -> hit1[job] = (t, record)              persistent, dynamic fetch
//   K2 shade1  one thread per job       -> path state, shadow-ray queue, bounce-ray queue (warp-aggregated append)
//   K3 trace   <shadow + bounce queues> -> occluded ? ndl1 = 0 ; hit2[slot]     one launch for both ray classes
//   K4 shade2  one thread per bounce    -> path state, second shadow-ray queue
//   K5 trace   <shadow queue 2>         -> occluded ? ndl2 = 0
//   K6 combine one thread per job       -> radiance, sample sum, A8R8G8B8 pixel (coalesced in 8x4 tile order)
//
// The trace kernel does nothing but traverse: ~30 fewer registers than the megakernel, no path state machine,
// and every shading instruction runs on full warps.  A multi-sample frame runs as ONE pass over pixels x samples
// jobs (job = sample * plane_jobs + pixel job, so 32 consecutive jobs are one sample of one 8x4 tile): six to
// twelve launches per frame whatever spp is, and trace kernels whose queues are spp times longer (shorter tails).
// K6 runs one thread per pixel and adds the pixel's samples in sample order -- the per-pixel loop's additions in
// its order, which keeps the frame bit-identical to the oracle's.  Frames whose pixels x samples exceed the queue
// budget run as several passes over sample ranges with a per-pixel running sum in between.
//
// All arithmetic that reaches a pixel is the shared no-FMA code of rt_device.cuh; which kernel evaluates it
// (and with how many lanes) cannot change a bit.  Rays are appended to queues in arrival order, which varies from
// run to run; every ray carries its job id and every result is stored by job / slot, so the frame does not.
#pragma once

#include "rt_wide.cuh"

namespace srt {

#ifndef SRT_WAVE_MIN_BLOCKS
#define SRT_WAVE_MIN_BLOCKS 4
#endif

// Device-side control block of one wavefront pass (zeroed before the first sample plane, and again by K6).
struct WaveCtl {
  uint32_t n_shadow1, n_bounce, n_shadow2, pad;
  unsigned long long cursor[4];  // job cursors of K1, K3, K5 (+ spare)
};

// Internal ray record, 32 B: (o.xyz, job | d.xyz, tmax); tmin = 0 for every generated ray.
struct WaveBufs {
  uint32_t job_base; // first job of this chain (a frame may run as several independent chains, rt_pipeline.cu)
  WaveCtl* ctl;
  uint2* hit1;      // [jobs]  (t, triangle record) of the primary ray; record = SRT_MISS when nothing was hit
  float4* path;     // [jobs]  (albedo1 bits, n.l at hit 1 or 0, albedo2 bits | kBounceMissed, n.l at hit 2 or 0)
  float4* q_shadow; // [jobs]  shadow rays of the stage in flight (K2 -> K3, then K4 -> K5)
  float4* q_bounce; // [2 * jobs] bounce rays (K2 -> K3, read again by K4)
  uint2* hit2;      // [jobs]  (t, record) per bounce-queue slot
  float4* sum;      // [plane_jobs]  running sample sum (only when a frame needs more than one pass)
};
constexpr float kBounceMissed = -1.0f;  // path.w of a path whose bounce ray left the scene (n.l values stored are >= 0)

// job -> pixel of the local strip: 32 consecutive jobs = one 8x4 tile, tiles in row-major order
__device__ __forceinline__ bool job_pixel(uint32_t job, uint32_t W, uint32_t rows, uint32_t tiles_x, uint32_t& x, uint32_t& ly) {
  const uint32_t tile = job >> 5, l = job & 31u;
  x = (tile % tiles_x) * 8u + (l & 7u);
  ly = (tile / tiles_x) * 4u + (l >> 3);
  return x < W && ly < rows;
}

__device__ __forceinline__ V3 wave_camera_dir(const kernel_arg_t& a, const Camera& cam, uint32_t seed_hash, uint32_t x,
                                              uint32_t y, uint32_t s) {
  float jx = 0.5f, jy = 0.5f;
  if (a.spp > 1) { jx = rnd01(seed_hash, y * a.dst_width + x, s, 0); jy = rnd01(seed_hash, y * a.dst_width + x, s, 1); }
  return camera_dir(cam, x, y, jx, jy);
}

// warp-aggregated queue append: returns the slot of every lane with `want`
__device__ __forceinline__ uint32_t queue_slot(uint32_t* counter, bool want) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t mask = __ballot_sync(0xffffffffu, want);
  uint32_t base = 0;
  if (mask == 0u) return 0;
  if (lane == (uint32_t)(__ffs(mask) - 1)) base = atomicAdd(counter, (uint32_t)__popc(mask));
  base = __shfl_sync(0xffffffffu, base, __ffs(mask) - 1);
  return base + __popc(mask & ((1u << lane) - 1u));
}

// ---- the trace kernel ----------------------------------------------------------------------------------
// STAGE 0: camera rays of sample plane s (job = pixel in tile order)       -> hit1[job]
// STAGE 1: jobs [0, n_shadow1) = shadow rays of hit 1 (any hit)            -> path[job].y = 0 when occluded
//          jobs [n_shadow1, n_shadow1 + n_bounce) = bounce rays (closest)  -> hit2[slot]
// STAGE 2: shadow rays of hit 2 (any hit)                                  -> path[job].w = 0 when occluded
// STAGE 3: stage 1 with K4 and K5 folded in (the default): a lane whose bounce ray ends shades the bounce hit itself
//          (the arithmetic of rt_wave_shade2_kernel), writes the path state and goes on with that path's second
//          shadow ray instead of taking a new job.  Removes a latency-bound shading pass and a trace launch that ran
//          at 30 % warps active (its queue is short and ends in a long tail); the bounce rays are the back half of
//          the launch's job list, so the in-lane shading runs when most advancing lanes of a warp need it.
// Persistent: grid = #SM x resident CTAs, 32 jobs per fetch from ctl->cursor[STAGE], lanes refilled by ballot /
// prefix popc when fewer than `tune` lanes of the warp are still traversing (run_warp, rt_wide.cuh).
template <int STAGE, int SCHED, bool STATS>
__global__ void __launch_bounds__(kWideThreads, SRT_WAVE_MIN_BLOCKS)
rt_wave_trace_kernel(const __grid_constant__ kernel_arg_t a, const WideScene sc, const WaveBufs wb, uint32_t rows,
                     uint32_t cam_jobs, uint32_t plane_jobs, uint32_t s_base, unsigned long long* counters) {
  extern __shared__ uint2 s_stack[];
  uint2 spill[kLocalStack];
  Stack st;
  st.base = (uint32_t)__cvta_generic_to_shared(s_stack + threadIdx.x);
  st.spill = spill;
  st.reset();
  const uint32_t W = a.dst_width, tiles_x = (W + 7u) / 8u;
  const uint32_t seed_hash = mix32(a.seed);
  const Camera cam = make_camera(a);
  constexpr bool kStage1 = STAGE == 1 || STAGE == 3, kFused = STAGE == 3;
  const uint32_t n_first = kStage1 ? wb.ctl->n_shadow1 : 0u;
  const uint32_t total_jobs = STAGE == 0 ? cam_jobs : (kStage1 ? n_first + wb.ctl->n_bounce : wb.ctl->n_shadow2);
  uint32_t n_second = 0u;  // fused: second shadow rays started by this lane
  bool second = false;     // fused: the any-hit ray in flight is the path's second shadow ray
  Pool pool = {0u, 0u, false};
  RayState r;
  Hit best;
  Stats stats = {0u, 0u, 0u, 0u, 0u};
  Cursor cu;
  uint32_t my_job = 0u;   // STAGE 0: job; STAGE 1 / 2: queue index (shadow) or n_first + slot (bounce)
  uint32_t my_path = 0u;  // job id the ray belongs to (from the ray record)
  bool have_job = false;
  cu.clear();
  best.t = 0.0f; best.u = 0.0f; best.v = 0.0f; best.prim = SRT_MISS; best.rec = SRT_MISS;
  ray_setup(r, cam.eye, cam.fwd, 0.0f, false);

  auto advance = [&]() {
    if (have_job && !cu.busy()) {
      have_job = false;
      if (r.any) {
        if (r.found) reinterpret_cast<float*>(wb.path + my_path)[(kStage1 && !second) ? 1 : 3] = 0.0f;
      } else if (kFused) {
        // the bounce hit, shaded here: rt_wave_shade2_kernel's statements on this lane's ray
        float* p = reinterpret_cast<float*>(wb.path + my_path);
        if (best.prim == SRT_MISS) {
          p[3] = kBounceMissed;
        } else {
          const float4 r1 = __ldg(sc.tris + 3 * (size_t)best.rec + 1), r2 = __ldg(sc.tris + 3 * (size_t)best.rec + 2);
          const Surface sf = surface_at(a, r1, r2, r.o, r.d, best.t);
          const bool lit = sf.ndl > 0.0f;
          p[2] = r1.w;
          p[3] = lit ? sf.ndl : 0.0f;
          if (lit && a.mode >= SRT_MODE_SHADOW) {
            ++n_second;
            second = true;
            have_job = true;  // the lane stays on this path
            if (STATS) stats.rays += 1;
            ray_setup(r, sf.po, sf.ld, 0.0f, true);
            best.t = sf.dist; best.prim = SRT_MISS; best.rec = SRT_MISS;
            st.reset();
            cu.go(0u);
          }
        }
      } else {
        const uint2 h = make_uint2(__float_as_uint(best.t), best.prim == SRT_MISS ? SRT_MISS : best.rec);
        if (STAGE == 0) wb.hit1[my_job] = h;
        else wb.hit2[my_job - n_first] = h;
      }
    }
    const uint32_t job = take_jobs(pool, !have_job, wb.ctl->cursor + (kStage1 ? 1 : STAGE), total_jobs);
    if (job != ~0u) {
      my_job = job;
      if (STAGE == 0) {
        uint32_t x, ly;
        my_job = job + wb.job_base;  // hit1 / path are indexed by the pass-wide job id = sample * plane_jobs + pixel job
        const uint32_t sp = my_job / plane_jobs;
        if (job_pixel(my_job - sp * plane_jobs, W, rows, tiles_x, x, ly)) {
          have_job = true;
          if (STATS) stats.rays += 1;
          ray_setup(r, cam.eye, wave_camera_dir(a, cam, seed_hash, x, global_row(a, ly), s_base + sp), 0.0f, false);
          best.t = kTMax; best.prim = SRT_MISS; best.rec = SRT_MISS;
          st.reset();
          cu.go(0u);
        }
      } else {
        const bool shadow = STAGE == 2 || job < n_first;
        const float4* q = shadow ? wb.q_shadow + 2 * (size_t)job : wb.q_bounce + 2 * (size_t)(job - n_first);
        const float4 r0 = __ldcs(q), r1 = __ldcs(q + 1);
        my_path = __float_as_uint(r0.w);
        have_job = true;
        if (STATS) stats.rays += 1;
        second = false;
        ray_setup(r, mk(r0.x, r0.y, r0.z), mk(r1.x, r1.y, r1.z), 0.0f, shadow);
        best.t = r1.w; best.prim = SRT_MISS; best.rec = SRT_MISS;
        st.reset();
        cu.go(0u);
      }
    }
  };
  // camera rays are coherent (ALU-bound): 16-byte loads; shadow / bounce rays (L1-bound): 32-byte loads
  run_warp<SCHED, STATS, STAGE != 0>(sc, cu, r, best, st, stats, pool, have_job, advance);
  if (kFused) {  // the combine adds ctl->n_shadow2 to the shadow-ray counter
    n_second = __reduce_add_sync(0xffffffffu, n_second);
    if ((threadIdx.x & 31u) == 0u && n_second) atomicAdd(&wb.ctl->n_shadow2, n_second);
  }
  if (STATS) flush_stats(stats, counters);
}

// ---- K2: shading at the primary hit, one thread per job --------------------------------------------------
// Writes the path state and appends this path's shadow ray and bounce ray.  The bounce ray does not depend on the
// shadow ray's answer, so both go out together and K3 traces them in one launch.
__global__ void __launch_bounds__(256)
rt_wave_shade1_kernel(const __grid_constant__ kernel_arg_t a, const float4* __restrict__ tris, const WaveBufs wb,
                      uint32_t rows, uint32_t cam_jobs, uint32_t plane_jobs, uint32_t s_base) {
  const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x, job = wb.job_base + local;
  const uint32_t W = a.dst_width, tiles_x = (W + 7u) / 8u;
  const uint32_t sp = job / plane_jobs, s = s_base + sp;
  uint32_t x = 0, ly = 0;
  const bool live = local < cam_jobs && job_pixel(job - sp * plane_jobs, W, rows, tiles_x, x, ly);
  bool want_shadow = false, want_bounce = false;
  float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0, b0 = s0, b1 = s0;
  if (live) {
    const uint2 h = wb.hit1[job];
    if (h.y != SRT_MISS) {
      const uint32_t seed_hash = mix32(a.seed);
      const Camera cam = make_camera(a);
      const uint32_t y = global_row(a, ly);
      const V3 d = wave_camera_dir(a, cam, seed_hash, x, y, s);
      const float4 r1 = __ldg(tris + 3 * (size_t)h.y + 1), r2 = __ldg(tris + 3 * (size_t)h.y + 2);
      const Surface sf = surface_at(a, r1, r2, cam.eye, d, __uint_as_float(h.x));
      const bool lit = sf.ndl > 0.0f;
      wb.path[job] = make_float4(r1.w, lit ? sf.ndl : 0.0f, 0.0f, 0.0f);
      if (lit && a.mode >= SRT_MODE_SHADOW) {
        want_shadow = true;
        s0 = make_float4(sf.po.x, sf.po.y, sf.po.z, __uint_as_float(job));
        s1 = make_float4(sf.ld.x, sf.ld.y, sf.ld.z, sf.dist);
      }
      if (a.mode >= SRT_MODE_BOUNCE) {
        const V3 bd = bounce_dir(seed_hash, y * W + x, s, sf.n);
        want_bounce = true;
        b0 = make_float4(sf.po.x, sf.po.y, sf.po.z, __uint_as_float(job));
        b1 = make_float4(bd.x, bd.y, bd.z, kTMax);
      }
    }
  }
  const uint32_t ss = queue_slot(&wb.ctl->n_shadow1, want_shadow);
  if (want_shadow) { wb.q_shadow[2 * (size_t)ss] = s0; wb.q_shadow[2 * (size_t)ss + 1] = s1; }
#ifdef SRT_NO_BOUNCE_SORT
  const uint32_t bs = queue_slot(&wb.ctl->n_bounce, want_bounce);
#else
  // Bounce rays leave neighbouring pixels in unrelated directions.  The CTA's (up to 256) rays -- eight adjacent
  // 8x4 tiles, so their origins are close -- are appended to the queue grouped by direction octant: the 32
  // consecutive rays a trace warp fetches then share the near/far child order of every node and stay together
  // for more of the tree.  A counting sort in shared memory, one global atomic per CTA; which slot a ray gets
  // cannot change its result (results are stored by slot / job).
  __shared__ uint32_t s_cnt[8], s_off[8];
  if (threadIdx.x < 8) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  const uint32_t key = (b1.x < 0.0f ? 1u : 0u) | (b1.y < 0.0f ? 2u : 0u) | (b1.z < 0.0f ? 4u : 0u);
  uint32_t pos = 0;
  if (want_bounce) pos = atomicAdd(&s_cnt[key], 1u);
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t total = 0;
    for (int k = 0; k < 8; ++k) { s_off[k] = total; total += s_cnt[k]; }
    const uint32_t base = total ? atomicAdd(&wb.ctl->n_bounce, total) : 0u;
    for (int k = 0; k < 8; ++k) s_off[k] += base;
  }
  __syncthreads();
  const uint32_t bs = s_off[key] + pos;
#endif
  if (want_bounce) { wb.q_bounce[2 * (size_t)bs] = b0; wb.q_bounce[2 * (size_t)bs + 1] = b1; }
}

// ---- K4: shading at the bounce hit, one thread per bounce-queue slot ------------------------------------------
__global__ void __launch_bounds__(256)
rt_wave_shade2_kernel(const __grid_constant__ kernel_arg_t a, const float4* __restrict__ tris, const WaveBufs wb) {
  const uint32_t n = wb.ctl->n_bounce;
  for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {  // warp-uniform trip count
    const uint32_t slot = base + threadIdx.x;
    bool want_shadow = false;
    float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
    if (slot < n) {
      const float4 q0 = wb.q_bounce[2 * (size_t)slot], q1 = wb.q_bounce[2 * (size_t)slot + 1];
      const uint32_t job = __float_as_uint(q0.w);
      const uint2 h = wb.hit2[slot];
      float* p = reinterpret_cast<float*>(wb.path + job);
      if (h.y == SRT_MISS) {
        p[3] = kBounceMissed;
      } else {
        const float4 r1 = __ldg(tris + 3 * (size_t)h.y + 1), r2 = __ldg(tris + 3 * (size_t)h.y + 2);
        const Surface sf = surface_at(a, r1, r2, mk(q0.x, q0.y, q0.z), mk(q1.x, q1.y, q1.z), __uint_as_float(h.x));
        const bool lit = sf.ndl > 0.0f;
        p[2] = r1.w;
        p[3] = lit ? sf.ndl : 0.0f;
        if (lit && a.mode >= SRT_MODE_SHADOW) {
          want_shadow = true;
          s0 = make_float4(sf.po.x, sf.po.y, sf.po.z, __uint_as_float(job));
          s1 = make_float4(sf.ld.x, sf.ld.y, sf.ld.z, sf.dist);
        }
      }
    }
    const uint32_t ss = queue_slot(&wb.ctl->n_shadow2, want_shadow);
    if (want_shadow) { wb.q_shadow[2 * (size_t)ss] = s0; wb.q_shadow[2 * (size_t)ss + 1] = s1; }
  }
}

// ---- K6: combine, one thread per PIXEL: the radiance expression of each of its samples, the sample sum, the pixel ----
// Direct term k = ambient + n.l * visibility, written as the megakernel and the oracle branch it (visibility is
// 0 or 1, an occluded or unlit hit has n.l stored as 0).  The pass holds samples [s_first, s_first + s_count) of
// every pixel at job = (s - s_first) * plane_jobs + pixel job; they are added in sample order onto the running sum
// of the earlier passes (wb.sum, only when s_first > 0) and the pixel is written when the last sample is in.
// Also closes the pass: adds the queue sizes of all chains to the ray counters and zeroes their control blocks.
__global__ void __launch_bounds__(256)
rt_wave_combine_kernel(const __grid_constant__ kernel_arg_t a, const WaveBufs wb, uint32_t rows, uint32_t plane_jobs,
                       uint32_t s_first, uint32_t s_count, WaveCtl* ctl, uint32_t chains, unsigned long long* counters) {
  const uint32_t pj = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t W = a.dst_width, tiles_x = (W + 7u) / 8u;
  uint32_t x = 0, ly = 0;
  if (pj < plane_jobs && job_pixel(pj, W, rows, tiles_x, x, ly)) {
    V3 sum = mk(0.0f, 0.0f, 0.0f);
    if (s_first > 0) { const float4 q = wb.sum[pj]; sum = mk(q.x, q.y, q.z); }
    for (uint32_t i = 0; i < s_count; ++i) {
      const uint32_t job = i * plane_jobs + pj;
      V3 rad = ld3(a.sky);
      const uint2 h = wb.hit1[job];
      if (h.y != SRT_MISS) {
        const float4 p = wb.path[job];
        const V3 alb = unpack_albedo(__float_as_uint(p.x));
        const float k1 = p.y > 0.0f ? a.ambient + p.y : a.ambient;
        rad = alb * k1;
        if (a.mode >= SRT_MODE_BOUNCE) {
          V3 c2 = ld3(a.sky);
          if (p.w != kBounceMissed) {
            const float k2 = p.w > 0.0f ? a.ambient + p.w : a.ambient;
            c2 = unpack_albedo(__float_as_uint(p.z)) * k2;
          }
          rad.x = rad.x + (alb.x * 0.5f) * c2.x;
          rad.y = rad.y + (alb.y * 0.5f) * c2.y;
          rad.z = rad.z + (alb.z * 0.5f) * c2.z;
        }
      }
      sum = sum + rad;
    }
    if (s_first + s_count < a.spp) {
      wb.sum[pj] = make_float4(sum.x, sum.y, sum.z, 0.0f);
    } else {
      const V3 c = sum / (float)a.spp;
      *reinterpret_cast<uint32_t*>(a.cbuf_addr + (uint64_t)x * a.cbuf_stride + (uint64_t)ly * a.cbuf_pitch) = pack_argb(c);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < chains) {
    WaveCtl* c = ctl + threadIdx.x;
    atomicAdd(counters + CTR_RAYS_SHADOW, (unsigned long long)c->n_shadow1 + c->n_shadow2);
    atomicAdd(counters + CTR_RAYS_BOUNCE, (unsigned long long)c->n_bounce);
    c->n_shadow1 = c->n_bounce = c->n_shadow2 = 0u;
    c->cursor[0] = c->cursor[1] = c->cursor[2] = 0ull;
  }
}

}  // namespace srt
