// rt_baseline.cuh -- variant 1: the baseline kernel every optimisation is A/B-measured against.
//
// One thread per pixel, the whole path (primary -> shadow -> bounce -> bounce shadow) in one
// kernel, traversing the builder's BVH2 exactly as uploaded (32-byte nodes, two 16-byte loads
// each), per-thread stack in local memory.  No repack, no persistence, no compaction.
// SURVEY.md section 7.1 step 3 ("first CUDA kernel, parity-checked").
#pragma once

#include "rt_device.cuh"

namespace srt {

constexpr int kBaseStack = 64;

// Conservative slab test: boxes inflated by dl, exit distance widened by kPad (DESIGN.md
// "Conservative box test").
__device__ __forceinline__ bool slab2(const float4 lo, const float4 hi, float dl, V3 o, V3 id, float tmin,
                                      float tmax, float& tn) {
  const float x0 = ((lo.x - dl) - o.x) * id.x, x1 = ((hi.x + dl) - o.x) * id.x;
  const float y0 = ((lo.y - dl) - o.y) * id.y, y1 = ((hi.y + dl) - o.y) * id.y;
  const float z0 = ((lo.z - dl) - o.z) * id.z, z1 = ((hi.z + dl) - o.z) * id.z;
  tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
  const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax));
  return tn <= tf * kPad;
}

// nodes as float4 pairs: [2n] = (bmin.xyz, left_first), [2n+1] = (bmax.xyz, count)

template <bool ANY, bool STATS>
__device__ __forceinline__ bool bvh2_trace(const float4* __restrict__ nodes, const float4* __restrict__ tris, V3 o,
                                           V3 d, float tmin, float tmax, Hit& best, Stats& st) {
  uint32_t stack[kBaseStack];
  int sp = 0;
  const V3 id = mk(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
  best.t = tmax; best.u = 0.0f; best.v = 0.0f; best.prim = SRT_MISS; best.rec = 0;
  if (STATS) st.rays += 1;
  float tn, dl;
  {
    const float4 lo = __ldg(nodes), hi = __ldg(nodes + 1);
    dl = scene_delta(lo, hi);
    if (STATS) st.nodes += 1;
    if (!slab2(lo, hi, dl, o, id, tmin, tmax, tn)) return false;
  }
  uint32_t cur = 0;
  for (;;) {
    const float4 lo = __ldg(nodes + 2 * (size_t)cur), hi = __ldg(nodes + 2 * (size_t)cur + 1);
    const uint32_t left_first = __float_as_uint(lo.w), count = __float_as_uint(hi.w);
    bool descend = false;
    if (STATS) st.nodes += 1;
    if (count > 0) {
      for (uint32_t i = 0; i < count; ++i) {
        if (STATS) st.tris += 1;
        if (ANY) {
          const float4 r0 = __ldg(tris + 3 * (size_t)(left_first + i));
          const float4 r1 = __ldg(tris + 3 * (size_t)(left_first + i) + 1);
          const float4 r2 = __ldg(tris + 3 * (size_t)(left_first + i) + 2);
          float t, u, v;
          if (intersect_tri(r0, r1, r2, o, d, t, u, v) && t > tmin && t < tmax) return true;
        } else {
          consider(tris, left_first + i, o, d, tmin, best);
        }
      }
    } else {
      const uint32_t l = left_first, r = left_first + 1;
      const float4 llo = __ldg(nodes + 2 * (size_t)l), lhi = __ldg(nodes + 2 * (size_t)l + 1);
      const float4 rlo = __ldg(nodes + 2 * (size_t)r), rhi = __ldg(nodes + 2 * (size_t)r + 1);
      float t0, t1;
      if (STATS) st.nodes += 2;
      const bool h0 = slab2(llo, lhi, dl, o, id, tmin, best.t, t0);
      const bool h1 = slab2(rlo, rhi, dl, o, id, tmin, best.t, t1);
      if (h0 && h1) {
        const bool swap = t1 < t0;
        cur = swap ? r : l;
        if (sp < kBaseStack) stack[sp++] = swap ? l : r;
        descend = true;
      } else if (h0) { cur = l; descend = true; }
      else if (h1) { cur = r; descend = true; }
    }
    if (!descend) {
      if (sp == 0) break;
      cur = stack[--sp];
    }
  }
  return best.prim != SRT_MISS;
}

// Lambert term at a hit (ambient + n.l * visibility) -- the oracle's direct_light().
template <bool STATS>
__device__ __forceinline__ float base_direct(const kernel_arg_t& a, const float4* nodes, const float4* tris,
                                             uint32_t rec, V3 o, V3 d, float t, Surface& s, uint32_t& n_shadow,
                                             Stats& st) {
  const float4 r1 = __ldg(tris + 3 * (size_t)rec + 1), r2 = __ldg(tris + 3 * (size_t)rec + 2);
  s = surface_at(a, r1, r2, o, d, t);
  if (!(s.ndl > 0.0f)) return a.ambient;
  float vis = 1.0f;
  if (a.mode >= SRT_MODE_SHADOW) {
    Hit dummy;
    ++n_shadow;
    if (bvh2_trace<true, STATS>(nodes, tris, s.po, s.ld, 0.0f, s.dist, dummy, st)) vis = 0.0f;
  }
  return a.ambient + s.ndl * vis;
}

template <bool STATS>
__global__ void __launch_bounds__(256) rt_baseline_kernel(const __grid_constant__ kernel_arg_t a,
                                                         uint32_t local_rows, unsigned long long* counters) {
  const float4* nodes = reinterpret_cast<const float4*>(a.node_addr);
  const float4* tris = reinterpret_cast<const float4*>(a.tri_addr);
  const uint32_t W = a.dst_width;
  const uint32_t tiles_x = (W + 7) / 8;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const uint32_t x = (warp % tiles_x) * 8 + (lane & 7);
  const uint32_t ly = (warp / tiles_x) * 4 + (lane >> 3);
  uint32_t n_shadow = 0, n_bounce = 0;
  Stats st = {0, 0, 0, 0, 0};
  if (x < W && ly < local_rows) {
    const uint32_t y = global_row(a, ly);
    const uint32_t pix = y * W + x;
    const uint32_t seed_hash = mix32(a.seed);
    const Camera cam = make_camera(a);
    V3 sum = mk(0.0f, 0.0f, 0.0f);
    for (uint32_t s = 0; s < a.spp; ++s) {
      float jx = 0.5f, jy = 0.5f;
      if (a.spp > 1) { jx = rnd01(seed_hash, pix, s, 0); jy = rnd01(seed_hash, pix, s, 1); }
      const V3 d = camera_dir(cam, x, y, jx, jy);
      Hit h;
      V3 rad = ld3(a.sky);
      if (bvh2_trace<false, STATS>(nodes, tris, cam.eye, d, 0.0f, kTMax, h, st)) {
        Surface sf;
        const float k = base_direct<STATS>(a, nodes, tris, h.rec, cam.eye, d, h.t, sf, n_shadow, st);
        const V3 alb = unpack_albedo(__float_as_uint(__ldg(tris + 3 * (size_t)h.rec + 1).w));
        rad = alb * k;
        if (a.mode >= SRT_MODE_BOUNCE) {
          const V3 bd = bounce_dir(seed_hash, pix, s, sf.n);
          Hit h2;
          V3 c2 = ld3(a.sky);
          ++n_bounce;
          if (bvh2_trace<false, STATS>(nodes, tris, sf.po, bd, 0.0f, kTMax, h2, st)) {
            Surface sf2;
            const float k2 = base_direct<STATS>(a, nodes, tris, h2.rec, sf.po, bd, h2.t, sf2, n_shadow, st);
            c2 = unpack_albedo(__float_as_uint(__ldg(tris + 3 * (size_t)h2.rec + 1).w)) * k2;
          }
          rad.x = rad.x + (alb.x * 0.5f) * c2.x;
          rad.y = rad.y + (alb.y * 0.5f) * c2.y;
          rad.z = rad.z + (alb.z * 0.5f) * c2.z;
        }
      }
      sum = sum + rad;
    }
    sum = sum / (float)a.spp;
    uint32_t* dst = reinterpret_cast<uint32_t*>(a.cbuf_addr + (uint64_t)x * a.cbuf_stride + (uint64_t)ly * a.cbuf_pitch);
    *dst = pack_argb(sum);
  }
  // ray counters: warp reduce, one atomic per warp and class
  n_shadow = __reduce_add_sync(0xffffffffu, n_shadow);
  n_bounce = __reduce_add_sync(0xffffffffu, n_bounce);
  if (lane == 0) {
    if (n_shadow) atomicAdd(counters + CTR_RAYS_SHADOW, (unsigned long long)n_shadow);
    if (n_bounce) atomicAdd(counters + CTR_RAYS_BOUNCE, (unsigned long long)n_bounce);
  }
  if (STATS) flush_stats(st, counters);
}

// Explicit ray list (mode 3): one thread per ray.
template <bool STATS>
__global__ void __launch_bounds__(256) rt_baseline_rays_kernel(const __grid_constant__ kernel_arg_t a,
                                                              unsigned long long* counters) {
  const float4* nodes = reinterpret_cast<const float4*>(a.node_addr);
  const float4* tris = reinterpret_cast<const float4*>(a.tri_addr);
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  Stats st = {0, 0, 0, 0, 0};
  if (i < a.num_rays) {
    const float4* rp = reinterpret_cast<const float4*>(a.ray_addr) + 2 * i;
    const float4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
    Hit h;
    bvh2_trace<false, STATS>(nodes, tris, mk(r0.x, r0.y, r0.z), mk(r1.x, r1.y, r1.z), r0.w, r1.w, h, st);
    reinterpret_cast<float4*>(a.hit_addr)[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.prim));
  }
  if (STATS) flush_stats(st, counters);
}

}  // namespace srt
