// rt_device.cuh -- device-side arithmetic of the RT path (sm_100a).
//
// Bit-exactness contract (DESIGN.md "Arithmetic"): every value that can reach a hit record or a
// pixel is computed in IEEE binary32, round-to-nearest, with NO fused multiply-add, in the
// expression order written here.  The translation unit is compiled with
//   -fmad=false -prec-div=true -prec-sqrt=true -ftz=false
// and the few places where an FMA is wanted (conservative box tests only, whose results never
// reach an output) say so explicitly with __fmaf_rn.
//
// The RT stages are ABSENT from the reference; the conventions
// kept from it are the framebuffer ones: A8R8G8B8 pixels at cbuf + x*stride + y*pitch
// (tests/regression/tex/kernel.cpp:99-101) and float->byte by truncation
// (tests/regression/draw3d/kernel.cpp:94-98).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "skybox_rt.h"

namespace srt {

constexpr float kTMax = 1e30f;
constexpr float kPad = 1.00012207031250f;  // 1 + 2^-13, slack of the conservative slab test

struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ V3 ld3(const float* p) { return mk(p[0], p[1], p[2]); }

// ---- counter-based RNG: same integer hash as the scene generator ---------------------------
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du;
  x ^= x >> 15; x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
__device__ __forceinline__ float rnd01(uint32_t seed_hash, uint32_t a, uint32_t b, uint32_t c) {
  // seed_hash = mix32(seed), hoisted by the caller
  uint32_t h = mix32(c + mix32(b + mix32(a + seed_hash)));
  return (float)(h >> 8) * (1.0f / 16777216.0f);
}

// ---- ray / triangle: Moeller-Trumbore on the builder's precomputed (v0, e1, e2) ------------
// r0 = (v0.xyz, prim_id), r1 = (e1.xyz, albedo), r2 = (e2.xyz, pad): three 16-byte loads.
__device__ __forceinline__ bool intersect_tri(const float4 r0, const float4 r1, const float4 r2, V3 o, V3 d,
                                              float& t, float& u, float& v) {
  const V3 v0 = mk(r0.x, r0.y, r0.z), e1 = mk(r1.x, r1.y, r1.z), e2 = mk(r2.x, r2.y, r2.z);
  const V3 p = cross(d, e2);
  const float det = dot(e1, p);
  if (det == 0.0f) return false;
  const float inv = 1.0f / det;
  const V3 tv = o - v0;
  const float uu = dot(tv, p) * inv;
  if (!(uu >= 0.0f) || uu > 1.0f) return false;
  const V3 q = cross(tv, e1);
  const float vv = dot(d, q) * inv;
  if (!(vv >= 0.0f) || uu + vv > 1.0f) return false;
  t = dot(e2, q) * inv;
  u = uu;
  v = vv;
  return true;
}

// ---- ray / box --------------------------------------------------------------------------------
__device__ __forceinline__ float safe_inv(float d) {
  const float a = fabsf(d);
  const float dd = a < 1e-20f ? (d < 0.0f ? -1e-20f : 1e-20f) : d;
  return 1.0f / dd;
}

// Box inflation: 2^-18 * largest |coordinate| of the root box (lo/hi = root bmin/bmax).
__device__ __forceinline__ float scene_delta(const float4 lo, const float4 hi) {
  float m = fmaxf(fmaxf(fabsf(lo.x), fabsf(lo.y)), fabsf(lo.z));
  m = fmaxf(m, fmaxf(fmaxf(fabsf(hi.x), fabsf(hi.y)), fabsf(hi.z)));
  return m * (1.0f / 262144.0f);
}

struct Hit {
  float t, u, v;
  uint32_t prim;  // soup index, SRT_MISS if none
  uint32_t rec;   // triangle-record index of the hit (device-internal)
};

// closest hit = min over (t, prim_id) of accepted triangles with tmin < t < tmax
__device__ __forceinline__ void consider(const float4* __restrict__ tris, uint32_t rec, V3 o, V3 d, float tmin,
                                         Hit& best) {
  const float4 r0 = __ldg(tris + 3 * (size_t)rec);
  const float4 r1 = __ldg(tris + 3 * (size_t)rec + 1);
  const float4 r2 = __ldg(tris + 3 * (size_t)rec + 2);
  float t, u, v;
  if (!intersect_tri(r0, r1, r2, o, d, t, u, v)) return;
  if (!(t > tmin)) return;
  const uint32_t prim = __float_as_uint(r0.w);
  if (t < best.t || (t == best.t && best.prim != SRT_MISS && prim < best.prim)) {
    best.t = t; best.u = u; best.v = v; best.prim = prim; best.rec = rec;
  }
}

__device__ __forceinline__ uint32_t to_byte(float c) {
  if (!(c > 0.0f)) c = 0.0f;
  if (c > 1.0f) c = 1.0f;
  return (uint32_t)(c * 255.0f);
}
__device__ __forceinline__ V3 unpack_albedo(uint32_t c) {
  return mk((float)((c >> 16) & 255u) * (1.0f / 255.0f), (float)((c >> 8) & 255u) * (1.0f / 255.0f),
            (float)(c & 255u) * (1.0f / 255.0f));
}
__device__ __forceinline__ uint32_t pack_argb(V3 c) {
  return 0xff000000u | (to_byte(c.x) << 16) | (to_byte(c.y) << 8) | to_byte(c.z);
}

// Camera ray through pixel (x,y), sub-pixel offset (jx,jy).  Row 0 is the bottom of the image.
struct Camera {
  V3 eye, fwd, right, up;
  float inv_w, inv_h, sx, sy;
};
__device__ __forceinline__ Camera make_camera(const kernel_arg_t& a) {
  Camera c;
  c.eye = ld3(a.cam_pos); c.fwd = ld3(a.cam_fwd); c.right = ld3(a.cam_right); c.up = ld3(a.cam_up);
  c.inv_w = 1.0f / (float)a.dst_width;
  c.inv_h = 1.0f / (float)a.dst_height;
  const float aspect = (float)a.dst_width / (float)a.dst_height;
  c.sx = a.tan_half_fov * aspect;
  c.sy = a.tan_half_fov;
  return c;
}
__device__ __forceinline__ V3 camera_dir(const Camera& c, uint32_t x, uint32_t y, float jx, float jy) {
  const float fx = (((float)x + jx) * c.inv_w) * 2.0f - 1.0f;
  const float fy = (((float)y + jy) * c.inv_h) * 2.0f - 1.0f;
  const V3 d = (c.fwd + c.right * (fx * c.sx)) + c.up * (fy * c.sy);
  return d / sqrtf(dot(d, d));
}

// Geometry of a hit needed by shading: oriented unit normal, offset point, light direction.
struct Surface {
  V3 n, po, ld;
  float dist, ndl;
};
__device__ __forceinline__ Surface surface_at(const kernel_arg_t& a, const float4 r1, const float4 r2, V3 o, V3 d,
                                              float t) {
  Surface s;
  V3 n = cross(mk(r1.x, r1.y, r1.z), mk(r2.x, r2.y, r2.z));
  n = n / sqrtf(dot(n, n));
  if (dot(n, d) > 0.0f) n = mk(-n.x, -n.y, -n.z);
  const V3 p = o + d * t;
  s.po = p + n * a.ray_eps;
  const V3 l = ld3(a.light_pos) - s.po;
  s.dist = sqrtf(dot(l, l));
  s.ld = l / s.dist;
  s.ndl = dot(n, s.ld);
  s.n = n;
  return s;
}

// Bounce direction for (pixel, sample): normalize(n + uniform point on the unit sphere).
__device__ __forceinline__ V3 bounce_dir(uint32_t seed_hash, uint32_t pix, uint32_t s, V3 n) {
  V3 sd = n;
  for (uint32_t tries = 0; tries < 16; ++tries) {
    const float x = 2.0f * rnd01(seed_hash, pix, s, 2 + 3 * tries) - 1.0f;
    const float y = 2.0f * rnd01(seed_hash, pix, s, 3 + 3 * tries) - 1.0f;
    const float z = 2.0f * rnd01(seed_hash, pix, s, 4 + 3 * tries) - 1.0f;
    const float r2 = (x * x + y * y) + z * z;
    if (r2 <= 1.0f && r2 > 1e-4f) {
      sd = mk(x, y, z) / sqrtf(r2);
      break;
    }
  }
  V3 bd = n + sd;
  const float bl = sqrtf(dot(bd, bd));
  if (bl < 1e-3f) bd = n; else bd = bd / bl;
  return bd;
}

struct Stats {
  uint32_t nodes, tris;  // node records fetched / triangles tested by this thread
  uint32_t rays;         // rays started by this thread
  uint32_t wnode, wtri;  // warp-level node / triangle steps this thread led (lowest active lane): nodes / wnode =
                         // lanes per node step, the execution efficiency of the traversal loop measured in place
};
// call inside a (possibly divergent) step: one lane of the active set counts the step for the warp
__device__ __forceinline__ bool step_leader() { return (threadIdx.x & 31u) == (uint32_t)(__ffs(__activemask()) - 1); }

// Per-thread traversal statistics -> device counters (warp reduce, one atomic per warp).  All 32
// lanes of the warp must call this.  Counter slots: vx_internal.h CTR_NODES = 2, CTR_TRIS = 3, CTR_PER_SM = 8.
template <class S>
__device__ __forceinline__ void flush_stats(const S& st, unsigned long long* counters) {
  const uint32_t n = __reduce_add_sync(0xffffffffu, st.nodes);
  const uint32_t t = __reduce_add_sync(0xffffffffu, st.tris);
  const uint32_t r = __reduce_add_sync(0xffffffffu, st.rays);
  const uint32_t wn = __reduce_add_sync(0xffffffffu, st.wnode), wt = __reduce_add_sync(0xffffffffu, st.wtri);
  if ((threadIdx.x & 31) == 0) {
    if (n) atomicAdd(counters + 2, (unsigned long long)n);
    if (t) atomicAdd(counters + 3, (unsigned long long)t);
    if (wn) atomicAdd(counters + 5, (unsigned long long)wn);
    if (wt) atomicAdd(counters + 6, (unsigned long long)wt);
    uint32_t smid;
    asm("mov.u32 %0, %%smid;" : "=r"(smid));
    if (r) atomicAdd(counters + 8 + (smid & 255u), (unsigned long long)r);
  }
}

// Row y of the frame -> row of this rank's strip buffer (or -1 if another rank owns it).
__device__ __forceinline__ int local_row(const kernel_arg_t& a, uint32_t y) {
  if (a.num_ranks <= 1) return (int)y;
  const uint32_t blk = y / a.rows_per_block;
  if (blk % a.num_ranks != a.rank_id) return -1;
  return (int)((blk / a.num_ranks) * a.rows_per_block + y % a.rows_per_block);
}
// Inverse: local row -> frame row.
__device__ __forceinline__ uint32_t global_row(const kernel_arg_t& a, uint32_t ly) {
  if (a.num_ranks <= 1) return ly;
  const uint32_t lb = ly / a.rows_per_block;
  return (lb * a.num_ranks + a.rank_id) * a.rows_per_block + ly % a.rows_per_block;
}

}  // namespace srt
