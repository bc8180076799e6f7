// raster_tile.cuh -- the raster hot path on sm_100a: one CTA per 32x32 tile, one thread per 2x2 quad.
//
// What one launch computes is what the reference's draw3d device kernel computes for one draw call
// (tests/regression/draw3d/kernel.cpp:232-299 with gpu_sw.h): for every binned tile, for every
// primitive of the tile in submission order: hierarchical coverage (sim/common/graphics.cpp:715-842),
// barycentric gradients + fixed-point attribute interpolation (kernel.cpp:37-59), texture sampling
// (graphics.cpp:35-314), texture environment (kernel.cpp:61-71), depth/stencil test
// (graphics.cpp:564-596), blending (graphics.cpp:407-524, :622-636) and the masked colour / depth
// writes (gpu_sw.h:103-168).
//
// B200 mapping.  The reference walks a tile's primitives sequentially and, per primitive, recurses
// 32 -> 16 -> 8 -> 4 -> 2 pixels before calling a scalar quad shader that reads and writes the
// framebuffer per fragment.  Here a tile's 256 quads are the 256 threads of a CTA: every thread owns
// its four pixels for the whole draw, so the colour and depth/stencil words live in REGISTERS across
// the primitive loop (one optional load at the start, one store of the touched pixels at the end;
// the order of fragments per pixel is the submission order by construction, no atomics).  Primitive
// records (120 B) are staged through shared memory in chunks by the whole CTA.  The reference's
// hierarchical rejection is evaluated per thread for its own ancestors (5 levels x 3 edges of integer
// multiply-add), which yields bit-identical coverage including the int32 wrap-around behaviour of the
// reference's fixed-point type.  Warps cover 16x8-pixel blocks, so coverage divergence follows
// triangle edges only.
//
// Arithmetic notes (parity): the gradient maths is fp32 with IEEE division and no FMA (this library is
// built -fmad=false -prec-div=true); float->int conversions mimic x86 cvttss2si (out of range / NaN ->
// 0x80000000) because that is what the reference's static_cast does on its host and simx builds.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "skybox_draw3d.h"

namespace srt {

struct RasterParams {
  // raster unit
  const uint8_t* tilebuf;
  const uint8_t* primbuf;
  uint32_t num_tiles, tile_format, prim_stride;
  uint32_t prim_limit;      // whole primitive records the primitive buffer holds: a larger id is refused
  uint64_t tilebuf_bytes;   // bytes of the tile buffer from `tilebuf` on: a primitive list that leaves them is refused
  uint32_t* fault;          // device view of the page-locked fault word the host reads after the run (RASTER_FAULT_*)
  uint32_t scissor_left, scissor_right, scissor_top, scissor_bottom;
  // output merger
  uint8_t* cbuf;
  uint8_t* zbuf;
  uint32_t cbuf_pitch, zbuf_pitch;
  uint32_t cbuf_writemask;  // expanded to byte lanes
  uint32_t stencil_writemask;
  uint32_t depth_func, s_func, s_zpass, s_zfail, s_fail, s_mask, s_ref;
  uint32_t bl_mode_rgb, bl_mode_a, bl_src_rgb, bl_src_a, bl_dst_rgb, bl_dst_a, bl_const, logic_op;
  // texture unit, stage 0, level 0
  const uint8_t* texbuf;
  uint32_t tex_log_width, tex_log_height, tex_format, tex_filter, tex_wrapu, tex_wrapv;
  // switches
  uint8_t depth_writemask, color_read, color_write, om_depth_enabled, om_stencil_enabled, blend_enabled;
  uint8_t sh_depth, sh_color, sh_tex, sh_modulate;  // kernel_arg_t booleans
  unsigned long long* counters;                      // RCTR_* slots of raster_pipeline.cu (STATS instantiations only)
};

enum { RASTER_FAULT_LIST = 1, RASTER_FAULT_PRIM = 2 };  // bits of *RasterParams::fault
constexpr int kRasterThreads = 256;
constexpr int kPrimChunk = 48;  // primitives staged per round: 48 * 120 B = 5760 B of shared memory

__device__ __forceinline__ int32_t x86_cvtt(float f) {
  return (f >= -2147483648.0f && f < 2147483648.0f) ? (int32_t)f : (int32_t)0x80000000;
}

// ---------------------------------------------------------------- texture unit (graphics.cpp:35-314)
constexpr int kTexFrac = VX_TEX_FXD_FRAC;
constexpr int32_t kTexMask = (1 << kTexFrac) - 1;

__device__ __forceinline__ int32_t tex_wrap(int32_t d, uint32_t wrap) {
  int32_t ret;
  if (wrap == VX_TEX_WRAP_REPEAT) {
    ret = d;
  } else if (wrap == VX_TEX_WRAP_MIRROR) {
    ret = d ^ ((int32_t)((uint32_t)d << (31 - kTexFrac)) >> 31);
  } else {  // clamp
    ret = d & -(int32_t)(d >= 0);
    ret |= ((kTexMask - ret) >> 31);
  }
  return ret & kTexMask;
}

__device__ __forceinline__ void unpack8888(uint32_t format, uint32_t texel, uint32_t& lo, uint32_t& hi) {
  uint32_t r, g, b, a;
  switch (format) {
    default:
    case VX_TEX_FORMAT_A8R8G8B8:
      lo = texel & 0x00ff00ff;
      hi = (texel >> 8) & 0x00ff00ff;
      return;
    case VX_TEX_FORMAT_R5G6B5: {
      // the reference widens by bit replication, c8 = (c5 << 3) | (c5 >> 2) and (c6 << 2) | (c6 >> 4)
      // (graphics.cpp:84-89); that equals (c5 * 33) >> 2 and (c6 * 65) >> 4, done here on r and b at once
      const uint32_t rb = ((texel & 0xf800u) << 5) | (texel & 0x1fu);  // r5 in the high half-word, b5 in the low
      lo = ((rb * 33u) >> 2) & 0x00ff00ffu;
      hi = 0x00ff0000u | ((((texel >> 5) & 0x3fu) * 65u) >> 4);
      return;
    }
    case VX_TEX_FORMAT_A1R5G5B5:
      r = ((texel >> 7) & 0xf8) | ((texel >> 12) & 0x07);
      g = ((texel >> 2) & 0xf8) | ((texel >> 7) & 0x07);
      b = ((texel << 3) & 0xf8) | ((texel >> 2) & 0x07);
      a = (uint32_t)(((int32_t)(texel << 16)) >> 31) & 0xff;
      break;
    case VX_TEX_FORMAT_A4R4G4B4:
      r = ((texel >> 4) & 0xf0) | ((texel >> 8) & 0x0f);
      g = ((texel >> 0) & 0xf0) | ((texel >> 4) & 0x0f);
      b = ((texel << 4) & 0xf0) | ((texel >> 0) & 0x0f);
      a = ((texel >> 8) & 0xf0) | ((texel >> 12) & 0x0f);
      break;
    case VX_TEX_FORMAT_A8L8:
      r = g = b = texel & 0xff;
      a = (texel >> 8) & 0xff;
      break;
    case VX_TEX_FORMAT_L8:
      r = g = b = texel & 0xff;
      a = 0xff;
      break;
    case VX_TEX_FORMAT_A8:
      r = g = b = 0xff;
      a = texel & 0xff;
      break;
  }
  lo = (r << 16) + b;
  hi = (a << 16) + g;
}

__device__ __forceinline__ uint32_t lerp8888(uint32_t a, uint32_t b, uint32_t f) {  // graphics.h:82-86
  uint32_t p = a * (0xff - f) + b * f + 0x00800080;
  uint32_t q = (p >> 8) & 0x00ff00ff;
  return ((p + q) >> 8) & 0x00ff00ff;
}

__device__ __forceinline__ uint32_t fetch_texel(const uint8_t* base, uint32_t offset, uint32_t format) {
  if (format == VX_TEX_FORMAT_A8R8G8B8) return __ldg(reinterpret_cast<const uint32_t*>(base) + offset);
  if (format <= VX_TEX_FORMAT_A8L8) return __ldg(reinterpret_cast<const uint16_t*>(base) + offset);
  return __ldg(base + offset);
}

// fmt / filter / wrapu / wrapv are passed separately so that a specialised kernel can hand in constants
__device__ __forceinline__ uint32_t tex_sample(const RasterParams& P, int32_t u, int32_t v, uint32_t fmt, uint32_t filter,
                                               uint32_t wrapu, uint32_t wrapv) {
  const uint32_t lw = P.tex_log_width, lh = P.tex_log_height;
  if (filter == VX_TEX_FILTER_BILINEAR) {
    const int32_t dx = (1 << (kTexFrac - 1)) >> lw, dy = (1 << (kTexFrac - 1)) >> lh;
    const uint32_t u0 = tex_wrap(u - dx, wrapu), u1 = tex_wrap(u + dx, wrapu);
    const uint32_t v0 = tex_wrap(v - dy, wrapv), v1 = tex_wrap(v + dy, wrapv);
    const uint32_t shift_u = kTexFrac - lw, shift_v = kTexFrac - lh;
    const uint32_t x0s = (u0 << 8) >> shift_u, y0s = (v0 << 8) >> shift_v;
    const uint32_t x0 = x0s >> 8, y0 = y0s >> 8, x1 = u1 >> shift_u, y1 = v1 >> shift_v;
    const uint32_t alpha = x0s & 0xff, beta = y0s & 0xff;
    const uint32_t t00 = fetch_texel(P.texbuf, x0 + (y0 << lw), fmt);
    const uint32_t t01 = fetch_texel(P.texbuf, x1 + (y0 << lw), fmt);
    const uint32_t t10 = fetch_texel(P.texbuf, x0 + (y1 << lw), fmt);
    const uint32_t t11 = fetch_texel(P.texbuf, x1 + (y1 << lw), fmt);
    uint32_t c0l, c0h, c1l, c1h, c2l, c2h, c3l, c3h;
    unpack8888(fmt, t00, c0l, c0h);
    unpack8888(fmt, t01, c1l, c1h);
    unpack8888(fmt, t10, c2l, c2h);
    unpack8888(fmt, t11, c3l, c3h);
    const uint32_t cl = lerp8888(lerp8888(c0l, c1l, alpha), lerp8888(c2l, c3l, alpha), beta);
    const uint32_t ch = lerp8888(lerp8888(c0h, c1h, alpha), lerp8888(c2h, c3h, alpha), beta);
    return (ch << 8) | cl;
  }
  const uint32_t uu = tex_wrap(u, wrapu), vv = tex_wrap(v, wrapv);
  const uint32_t x = uu >> (kTexFrac - lw), y = vv >> (kTexFrac - lh);
  uint32_t cl, ch;
  unpack8888(fmt, fetch_texel(P.texbuf, x + (y << lw), fmt), cl, ch);
  return (ch << 8) | cl;
}

// ---------------------------------------------------------------- depth / stencil (graphics.cpp:320-364, :564-596)
__device__ __forceinline__ bool do_compare(uint32_t func, uint32_t a, uint32_t b) {
  switch (func) {
    default:
    case VX_OM_DEPTH_FUNC_NEVER: return false;
    case VX_OM_DEPTH_FUNC_LESS: return a < b;
    case VX_OM_DEPTH_FUNC_EQUAL: return a == b;
    case VX_OM_DEPTH_FUNC_LEQUAL: return a <= b;
    case VX_OM_DEPTH_FUNC_GREATER: return a > b;
    case VX_OM_DEPTH_FUNC_NOTEQUAL: return a != b;
    case VX_OM_DEPTH_FUNC_GEQUAL: return a >= b;
    case VX_OM_DEPTH_FUNC_ALWAYS: return true;
  }
}

__device__ __forceinline__ uint32_t do_stencil_op(uint32_t op, uint32_t ref, uint32_t val) {
  switch (op) {
    default:
    case VX_OM_STENCIL_OP_KEEP: return val;
    case VX_OM_STENCIL_OP_ZERO: return 0;
    case VX_OM_STENCIL_OP_REPLACE: return ref;
    case VX_OM_STENCIL_OP_INCR: return (val < 0xff) ? (val + 1) : val;
    case VX_OM_STENCIL_OP_DECR: return (val > 0) ? (val - 1) : val;
    case VX_OM_STENCIL_OP_INVERT: return ~val;
    case VX_OM_STENCIL_OP_INCR_WRAP: return (val + 1) & 0xff;
    case VX_OM_STENCIL_OP_DECR_WRAP: return (val - 1) & 0xff;
  }
}

__device__ __forceinline__ bool ds_test(const RasterParams& P, uint32_t depth, uint32_t ds_val, uint32_t& ds_result) {
  const uint32_t depth_val = ds_val & VX_OM_DEPTH_MASK, stencil_val = ds_val >> VX_OM_DEPTH_BITS;
  const uint32_t depth_ref = depth & VX_OM_DEPTH_MASK;
  bool passed = do_compare(P.s_func, P.s_ref & P.s_mask, stencil_val & P.s_mask);
  uint32_t op;
  if (passed) {
    passed = do_compare(P.depth_func, depth_ref, depth_val);
    op = passed ? P.s_zpass : P.s_zfail;
  } else {
    op = P.s_fail;
  }
  ds_result = (do_stencil_op(op, P.s_ref, stencil_val) << VX_OM_DEPTH_BITS) | depth_ref;
  return passed;
}

// ---------------------------------------------------------------- blender (graphics.cpp:366-524, :622-636)
// Colours are handled as packed A8R8G8B8 words; ch(c, k) extracts channel k (0 = b .. 3 = a).
__device__ __forceinline__ int ch(uint32_t c, int k) { return (int)((c >> (8 * k)) & 0xff); }

__device__ __forceinline__ uint32_t blend_factor(uint32_t func, uint32_t src, uint32_t dst, uint32_t cst) {
  const uint32_t sa = src >> 24, da = dst >> 24, ca = cst >> 24;
  switch (func) {
    default:
    case VX_OM_BLEND_FUNC_ZERO: return 0;
    case VX_OM_BLEND_FUNC_ONE: return 0xffffffffu;
    case VX_OM_BLEND_FUNC_SRC_RGB: return src;
    case VX_OM_BLEND_FUNC_ONE_MINUS_SRC_RGB: return ~src;  // 0xff - x per byte lane
    case VX_OM_BLEND_FUNC_DST_RGB: return dst;
    case VX_OM_BLEND_FUNC_ONE_MINUS_DST_RGB: return ~dst;
    case VX_OM_BLEND_FUNC_SRC_A: return sa * 0x01010101u;
    case VX_OM_BLEND_FUNC_ONE_MINUS_SRC_A: return (0xff - sa) * 0x01010101u;
    case VX_OM_BLEND_FUNC_DST_A: return da * 0x01010101u;
    case VX_OM_BLEND_FUNC_ONE_MINUS_DST_A: return (0xff - da) * 0x01010101u;
    case VX_OM_BLEND_FUNC_CONST_RGB: return cst;
    case VX_OM_BLEND_FUNC_ONE_MINUS_CONST_RGB: return ~cst;
    case VX_OM_BLEND_FUNC_CONST_A: return ca * 0x01010101u;
    case VX_OM_BLEND_FUNC_ONE_MINUS_CONST_A: return ~cst;  // sic: the reference complements every channel (graphics.cpp:466-472)
    case VX_OM_BLEND_FUNC_ALPHA_SAT: {
      const uint32_t f = min(sa, 0xff - da);
      return 0xff000000u | (f * 0x00010101u);
    }
  }
}

__device__ __forceinline__ int div255(int x) { return (x + (x >> 8)) >> 8; }

__device__ __forceinline__ uint32_t blend_mode(uint32_t mode, uint32_t logic_op, uint32_t src, uint32_t dst, uint32_t s, uint32_t d) {
  uint32_t out = 0;
  switch (mode) {
    default:
    case VX_OM_BLEND_MODE_ADD:
#pragma unroll
      for (int k = 0; k < 4; ++k) out |= (uint32_t)div255(min(ch(src, k) * ch(s, k) + ch(dst, k) * ch(d, k) + 0x80, 0xFF00)) << (8 * k);
      return out;
    case VX_OM_BLEND_MODE_SUB:
#pragma unroll
      for (int k = 0; k < 4; ++k) out |= ((uint32_t)div255(max(ch(src, k) * ch(s, k) - ch(dst, k) * ch(d, k) + 0x80, 0)) & 0xff) << (8 * k);
      return out;
    case VX_OM_BLEND_MODE_REV_SUB:
#pragma unroll
      for (int k = 0; k < 4; ++k) out |= ((uint32_t)div255(max(ch(dst, k) * ch(d, k) - ch(src, k) * ch(s, k) + 0x80, 0)) & 0xff) << (8 * k);
      return out;
    case VX_OM_BLEND_MODE_MIN: return __vminu4(src, dst);
    case VX_OM_BLEND_MODE_MAX: return __vmaxu4(src, dst);
    case VX_OM_BLEND_MODE_LOGICOP:
      switch (logic_op) {
        default:
        case 0: return 0;
        case 1: return src & dst;
        case 2: return src & ~dst;
        case 3: return src;
        case 4: return ~src & dst;
        case 5: return dst;
        case 6: return src ^ dst;
        case 7: return src | dst;
        case 8: return ~(src | dst);
        case 9: return ~(src ^ dst);
        case 10: return ~dst;
        case 11: return src | ~dst;
        case 12: return ~src;
        case 13: return ~src | dst;
        case 14: return ~(src & dst);
        case 15: return 0xffffffffu;
      }
  }
}

struct BlendState { uint32_t mode_rgb, mode_a, src_rgb, src_a, dst_rgb, dst_a; };
__device__ __forceinline__ uint32_t blend(const RasterParams& P, const BlendState& B, uint32_t src, uint32_t dst) {
  const uint32_t s_rgb = blend_factor(B.src_rgb, src, dst, P.bl_const), s_a = blend_factor(B.src_a, src, dst, P.bl_const);
  const uint32_t d_rgb = blend_factor(B.dst_rgb, src, dst, P.bl_const), d_a = blend_factor(B.dst_a, src, dst, P.bl_const);
  const uint32_t rgb = blend_mode(B.mode_rgb, P.logic_op, src, dst, s_rgb, d_rgb);
  const uint32_t a = blend_mode(B.mode_a, P.logic_op, src, dst, s_a, d_a);
  return (a & 0xff000000u) | (rgb & 0x00ffffffu);
}

// ---------------------------------------------------------------- shader helpers (kernel.cpp:46-71)
__device__ __forceinline__ int32_t imadd24(int32_t a, int32_t b, int32_t c) {
  // ((int64)a * b) >> 24 truncated to 32 bits, + c
  const uint32_t lo = (uint32_t)a * (uint32_t)b;
  const uint32_t hi = (uint32_t)__mulhi(a, b);
  return (int32_t)(((hi << 8) | (lo >> 24)) + (uint32_t)c);
}
__device__ __forceinline__ int32_t interp(const int32_t* p, int32_t dx, int32_t dy) {  // p = {x, y, z} of one attribute
  return imadd24(p[1], dy, imadd24(p[0], dx, p[2]));
}
__device__ __forceinline__ uint32_t mul_u8(int32_t fx, uint32_t k) { return (((uint32_t)fx * k) >> 24) & 0xff; }

// ---------------------------------------------------------------- the tile kernel
// Thread -> quad: warp w covers the 16x8-pixel block at ((w & 1) * 16, (w >> 1) * 8) of the tile, lane l the
// quad (l & 7, l >> 3) of the block.  A warp's framebuffer rows are then 64 contiguous bytes.
//
// Specialisation.  All fixed-function state is uniform over a launch; the generic kernel walks it through
// switch statements per fragment (format decode, blend factors, compare functions).  TEXK / BLENDK select
// instantiations in which the commonest states are compile-time constants so those switches fold away:
//   TEXK   0 no texture | 1 any state | 2 A8R8G8B8 | 3 R5G6B5 | 4 A8L8  (2..4: bilinear, repeat/repeat)
//   BLENDK 0 blending off | 1 any state | 2 SRC_A / ONE_MINUS_SRC_A, ADD (glBlendFunc(SRC_ALPHA, ONE_MINUS_SRC_ALPHA))
// raster_pipeline.cu picks the instantiation from the decoded state; results are identical by construction
// (same functions, constant arguments).
//
// PPT = pixels per thread.  4: one thread per 2x2 quad, 256 threads per tile (best throughput: the coverage
// tests are shared by four pixels).  1: one thread per pixel, 1024 threads per tile, for draws with fewer tiles
// than the GPU has room for -- a tile's primitive list is a serial chain, and with 32 warps instead of 8
// walking it the chain of the heaviest tile (the draw's critical path) is ~4x shorter.
template <bool STATS, int TEXK, int BLENDK, int PPT>
__global__ void __launch_bounds__(PPT == 4 ? kRasterThreads : 1024) raster_tile_kernel(const RasterParams P) {
  constexpr uint32_t NT = PPT == 4 ? kRasterThreads : 1024;
  const uint32_t tex_fmt = TEXK == 2 ? VX_TEX_FORMAT_A8R8G8B8 : TEXK == 3 ? VX_TEX_FORMAT_R5G6B5 : TEXK == 4 ? VX_TEX_FORMAT_A8L8 : P.tex_format;
  const uint32_t tex_filter = TEXK >= 2 ? VX_TEX_FILTER_BILINEAR : P.tex_filter;
  const uint32_t tex_wrapu = TEXK >= 2 ? VX_TEX_WRAP_REPEAT : P.tex_wrapu, tex_wrapv = TEXK >= 2 ? VX_TEX_WRAP_REPEAT : P.tex_wrapv;
  const bool sh_tex = TEXK != 0, blend_enabled = BLENDK != 0;
  BlendState B;
  if (BLENDK == 2) {
    B.mode_rgb = B.mode_a = VX_OM_BLEND_MODE_ADD;
    B.src_rgb = B.src_a = VX_OM_BLEND_FUNC_SRC_A;
    B.dst_rgb = B.dst_a = VX_OM_BLEND_FUNC_ONE_MINUS_SRC_A;
  } else {
    B.mode_rgb = P.bl_mode_rgb; B.mode_a = P.bl_mode_a; B.src_rgb = P.bl_src_rgb; B.src_a = P.bl_src_a;
    B.dst_rgb = P.bl_dst_rgb; B.dst_a = P.bl_dst_a;
  }
  __shared__ int32_t s_prims[kPrimChunk][30];  // rast_prim_t = 30 words
  __shared__ uint32_t s_mask[kPrimChunk];      // per staged primitive: which 8x8 blocks of the tile survive levels 5..3
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // tile header (gpu_sw.h:38-45)
  uint32_t tile_x, tile_y, pids_count;
  const uint32_t* pids;
  if (P.tile_format == SRT_TILEBUF_WIDE) {
    const srt_tile_header_wide_t* h = reinterpret_cast<const srt_tile_header_wide_t*>(P.tilebuf) + blockIdx.x;
    const uint4 hv = __ldg(reinterpret_cast<const uint4*>(h));
    tile_x = hv.x & 0xffff; tile_y = hv.x >> 16; pids_count = hv.y;
    pids = reinterpret_cast<const uint32_t*>(reinterpret_cast<const srt_tile_header_wide_t*>(P.tilebuf) + P.num_tiles) + hv.z;
  } else {
    const rast_tile_header_t* h = reinterpret_cast<const rast_tile_header_t*>(P.tilebuf) + blockIdx.x;
    const uint2 hv = __ldg(reinterpret_cast<const uint2*>(h));
    tile_x = hv.x & 0xffff; tile_y = hv.x >> 16; pids_count = hv.y >> 16;
    pids = reinterpret_cast<const uint32_t*>(h + 1) + (hv.y & 0xffff);
  }
  // a malformed tile buffer must not become a device fault: a list that leaves the buffer is skipped and reported
  if ((uint64_t)(reinterpret_cast<const uint8_t*>(pids) - P.tilebuf) + 4ull * pids_count > P.tilebuf_bytes) {
    if (tid == 0) atomicOr(P.fault, (uint32_t)RASTER_FAULT_LIST);
    pids_count = 0;
  }
  const uint32_t X0 = tile_x << RASTER_TILE_LOGSIZE, Y0 = tile_y << RASTER_TILE_LOGSIZE;
  // PPT 4: warp = 16x8 pixels (8x4 quads); PPT 1: warp = 8x4 pixels.  Either way inside one 16x16 block.
  const uint32_t lx = PPT == 4 ? ((warp & 1) << 4) + ((lane & 7) << 1) : ((warp & 3) << 3) + (lane & 7);
  const uint32_t ly = PPT == 4 ? ((warp >> 1) << 3) + ((lane >> 3) << 1) : ((warp >> 2) << 2) + (lane >> 3);
  const uint32_t px = X0 + lx, py = Y0 + ly;  // this thread's pixel / the top-left pixel of its quad

  // scissor per pixel (graphics.cpp:817-818); the app sets it to the render target
  uint32_t inside = 0;
#pragma unroll
  for (int p = 0; p < PPT; ++p) {
    const uint32_t x = px + (p & 1), y = py + (p >> 1);
    if (x >= P.scissor_left && x < P.scissor_right && y >= P.scissor_top && y < P.scissor_bottom) inside |= 1u << p;
  }

  // framebuffer words of the quad, resident in registers for the whole draw (gpu_sw.h:125-140 reads)
  uint32_t col[4] = {0, 0, 0, 0}, ds[4] = {0, 0, 0, 0};
  uint32_t col_dirty = 0, ds_dirty = 0;
  const bool need_ds = P.om_depth_enabled || P.om_stencil_enabled;
  const bool need_col = P.color_write && (P.color_read || blend_enabled);
#pragma unroll
  for (int p = 0; p < PPT; ++p) {
    if (inside & (1u << p)) {
      const size_t x = px + (p & 1), y = py + (p >> 1);
      if (need_ds) ds[p] = *reinterpret_cast<const uint32_t*>(P.zbuf + y * P.zbuf_pitch + x * 4);
      if (need_col) col[p] = *reinterpret_cast<const uint32_t*>(P.cbuf + y * P.cbuf_pitch + x * 4);
    }
  }

  const uint32_t my_bit = 1u << (((ly >> 3) << 2) | (lx >> 3));         // the 8x8 block of this thread's quad / pixel
  const uint32_t warp_bits = __reduce_or_sync(0xffffffffu, my_bit);    // one block (PPT 1) or two (PPT 4)
  unsigned long long n_frag = 0, n_tested = 0;
  uint32_t n_om_rd = 0;
  if (STATS) n_om_rd = (uint32_t)__popc(inside) * ((need_ds ? 1u : 0u) + (need_col ? 1u : 0u));

  for (uint32_t base = 0; base < pids_count; base += kPrimChunk) {
    const uint32_t n = min((uint32_t)kPrimChunk, pids_count - base);
    __syncthreads();  // previous chunk fully consumed
    for (uint32_t w = tid; w < n * 30; w += NT) {
      const uint32_t k = w / 30, j = w - k * 30;
      uint32_t pid = __ldg(pids + base + k);
      if (pid >= P.prim_limit) {  // an id past the primitive buffer: record 0 stands in and the run is reported
        if (j == 0) atomicOr(P.fault, (uint32_t)RASTER_FAULT_PRIM);
        pid = 0;
      }
      s_prims[k][j] = __ldg(reinterpret_cast<const int32_t*>(P.primbuf + (size_t)pid * P.prim_stride) + j);
    }
    __syncthreads();

    // Block masks: the reference's renderTile recursion rejects a block when value(origin) + (extent << level) < 0 on
    // any edge (graphics.cpp:749-753), at level 5 (the tile), 4 (16x16) and 3 (8x8) before it reaches the quads.
    // Those tests depend on the block only, so sixteen threads per primitive evaluate them once for the sixteen 8x8
    // blocks of the tile -- same int32 wrap-around expressions -- and every warp then drops a primitive that misses
    // its blocks with one shared-memory word instead of redoing the edge arithmetic.
    for (uint32_t w0 = 0; w0 < n * 16u; w0 += NT) {
      const uint32_t w = w0 + tid;
      bool pass = false;
      if (w < n * 16u) {
        const int32_t* pr = s_prims[w >> 4];
        const uint32_t bx = w & 3u, by = (w >> 2) & 3u;
        uint32_t rej = 0;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const uint32_t a = (uint32_t)pr[3 * i], b = (uint32_t)pr[3 * i + 1], c = (uint32_t)pr[3 * i + 2];
          const int32_t sa = (int32_t)a, sb = (int32_t)b;
          const uint32_t ext = (sb >= 0) ? ((sa >= 0) ? a + b : b) : ((sa >= 0) ? a : 0u);  // CalcEdgeExtents (graphics.cpp:644-647)
          rej |= (a * X0 + b * Y0 + c + (ext << 5)) >> 31;
          rej |= (a * (X0 + ((bx >> 1) << 4)) + b * (Y0 + ((by >> 1) << 4)) + c + (ext << 4)) >> 31;
          rej |= (a * (X0 + (bx << 3)) + b * (Y0 + (by << 3)) + c + (ext << 3)) >> 31;
        }
        pass = rej == 0;
      }
      const uint32_t bal = __ballot_sync(0xffffffffu, pass);
      if (w < n * 16u && (lane & 15u) == 0u) s_mask[w >> 4] = (bal >> (lane & 16u)) & 0xffffu;
    }
    __syncthreads();

    for (uint32_t k = 0; k < n; ++k) {
      const uint32_t blocks = s_mask[k];
      if (STATS) ++n_tested;
      if (!(blocks & warp_bits)) continue;  // warp-uniform: the primitive misses every block this warp covers
      if (!(blocks & my_bit)) continue;     // levels 5, 4, 3 of this thread's own ancestors
      const int32_t* pr = s_prims[k];
      // edges e[i] = (a, b, c) at pr[3i .. 3i+2]; int32 wrap-around arithmetic throughout.  The remaining levels of
      // the recursion (4x4 block, the quad: graphics.cpp:805-809) differ per thread.
      uint32_t reject = 0;
      int32_t eq[3];  // edge values at the quad's top-left pixel
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const uint32_t a = (uint32_t)pr[3 * i], b = (uint32_t)pr[3 * i + 1], c = (uint32_t)pr[3 * i + 2];
        const int32_t sa = (int32_t)a, sb = (int32_t)b;
        const uint32_t ext = (sb >= 0) ? ((sa >= 0) ? a + b : b) : ((sa >= 0) ? a : 0u);
#pragma unroll
        for (int L = RASTER_TILE_LOGSIZE - 3; L >= 1; --L) {
          const uint32_t ox = X0 + (lx & ~((1u << L) - 1)), oy = Y0 + (ly & ~((1u << L) - 1));
          reject |= (a * ox + b * oy + c + (ext << L)) >> 31;
        }
        eq[i] = (int32_t)(a * px + b * py + c);
      }
      if (reject) continue;
      // PREPARE_QUAD (graphics.cpp:811-830)
      uint32_t mask = 0;
      int32_t bc[4][3];
#pragma unroll
      for (int p = 0; p < PPT; ++p) {
        uint32_t neg = 0;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          bc[p][i] = (int32_t)((uint32_t)eq[i] + ((p & 1) ? (uint32_t)pr[3 * i] : 0u) + ((p >> 1) ? (uint32_t)pr[3 * i + 1] : 0u));
          neg |= (uint32_t)bc[p][i] >> 31;
        }
        if (!neg) mask |= 1u << p;
      }
      mask &= inside;
      if (!mask) continue;

#pragma unroll
      for (int p = 0; p < PPT; ++p) {
        if (!(mask & (1u << p))) continue;
        if (STATS) ++n_frag;
        // GRADIENTS_SW (kernel.cpp:37-44)
        const float F0 = (float)bc[p][0] / 16777216.0f, F1 = (float)bc[p][1] / 16777216.0f, F2 = (float)bc[p][2] / 16777216.0f;
        const float r = 1.0f / (F0 + F1 + F2);
        const int32_t dx = x86_cvtt((r * F0) * 16777216.0f), dy = x86_cvtt((r * F1) * 16777216.0f);
        // attributes at pr[9 ..]: z r g b a u v, each {d(v0-v2), d(v1-v2), v2}
        int32_t z = 0;
        uint32_t color;
        if (P.sh_depth) z = interp(pr + 9, dx, dy);
        int32_t cr = 1 << 24, cg = 1 << 24, cb = 1 << 24, ca = 1 << 24;
        if (P.sh_color) {
          cr = interp(pr + 12, dx, dy); cg = interp(pr + 15, dx, dy);
          cb = interp(pr + 18, dx, dy); ca = interp(pr + 21, dx, dy);
        }
        if (sh_tex) {
          const int32_t u = interp(pr + 24, dx, dy), v = interp(pr + 27, dx, dy);
          const uint32_t t = tex_sample(P, u >> 1, v >> 1, tex_fmt, tex_filter, tex_wrapu, tex_wrapv);  // 8.24 -> 9.23 (kernel.cpp:160-164)
          color = P.sh_modulate ? ((mul_u8(ca, t >> 24) << 24) | (mul_u8(cr, (t >> 16) & 0xff) << 16) |
                                   (mul_u8(cg, (t >> 8) & 0xff) << 8) | mul_u8(cb, t & 0xff))
                                : t;
        } else {
          color = (mul_u8(ca, 255) << 24) | (mul_u8(cr, 255) << 16) | (mul_u8(cg, 255) << 8) | mul_u8(cb, 255);
        }
        // OutputMerger::write (gpu_sw.h:103-168) on the register copies
        uint32_t ds_new = 0;
        const bool ds_passed = !need_ds || ds_test(P, (uint32_t)z, ds[p], ds_new);
        if (blend_enabled && ds_passed) color = blend(P, B, color, col[p]);
        const uint32_t ds_wm = ((P.om_depth_enabled && ds_passed && P.depth_writemask) ? VX_OM_DEPTH_MASK : 0u) |
                               (P.om_stencil_enabled ? (P.stencil_writemask << VX_OM_DEPTH_BITS) : 0u);
        if (ds_wm) {
          ds[p] = (ds[p] & ~ds_wm) | (ds_new & ds_wm);
          ds_dirty |= 1u << p;
        }
        if (P.color_write && ds_passed) {
          col[p] = (col[p] & ~P.cbuf_writemask) | (color & P.cbuf_writemask);
          col_dirty |= 1u << p;
        }
      }
    }
  }

  // write back what was touched
#pragma unroll
  for (int p = 0; p < PPT; ++p) {
    const size_t x = px + (p & 1), y = py + (p >> 1);
    if (ds_dirty & (1u << p)) *reinterpret_cast<uint32_t*>(P.zbuf + y * P.zbuf_pitch + x * 4) = ds[p];
    if (col_dirty & (1u << p)) *reinterpret_cast<uint32_t*>(P.cbuf + y * P.cbuf_pitch + x * 4) = col[p];
  }

  if (STATS) {  // slots: raster_pipeline.cu RCTR_*
    uint32_t n_om_wr = (uint32_t)(__popc(ds_dirty) + __popc(col_dirty));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      n_frag += __shfl_xor_sync(0xffffffffu, n_frag, o);
      n_tested += __shfl_xor_sync(0xffffffffu, n_tested, o);
      n_om_rd += __shfl_xor_sync(0xffffffffu, n_om_rd, o);
      n_om_wr += __shfl_xor_sync(0xffffffffu, n_om_wr, o);
    }
    if (lane == 0) {
      uint32_t smid;
      asm("mov.u32 %0, %%smid;" : "=r"(smid));
      atomicAdd(P.counters + 0, n_frag);
      atomicAdd(P.counters + 1, n_tested);
      if (sh_tex) atomicAdd(P.counters + 2, n_frag * (tex_filter ? 4ull : 1ull));  // texels fetched
      atomicAdd(P.counters + 3, (unsigned long long)n_om_rd);                       // framebuffer words read
      atomicAdd(P.counters + 4, (unsigned long long)n_om_wr);                       // framebuffer words written
      if (warp == 0) atomicAdd(P.counters + 5, 1ull + 2ull * pids_count);           // tile header + pid + primitive records
      atomicAdd(P.counters + 8 + (smid & 255u), n_frag);                            // fragments per SM
    }
  }
}

}  // namespace srt
