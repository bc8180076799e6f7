// raster_pipeline.cu -- host side of the raster pipeline behind vx_start (pipeline SRT_PIPELINE_DRAW3D).
//
// Decodes what the reference's device code decodes at kernel start -- GpuSW::configure
// (tests/regression/draw3d/gpu_sw.h:215-220) -> Rasterizer::configure (gpu_sw.h:25-32,
// sim/common/graphics.cpp:708-713), OutputMerger::configure (gpu_sw.h:80-101), DepthTencil::configure
// (graphics.cpp:533-562), Blender::configure (graphics.cpp:603-620), TextureSampler::read's DCR reads
// (graphics.cpp:253-262) -- into one RasterParams block passed by value to the tile kernel, and
// enqueues the launch on the device stream.  Nothing here touches pixels.
//
// State source per unit, as in the reference: a unit run in software (kernel_arg_t.sw_rast / sw_tex /
// sw_om) reads the DCR mirror inside kernel_arg_t; otherwise the unit is "hardware" and reads the
// device DCR file written through vx_dcr_write (sim/simx/raster_unit.cpp, tex_unit.cpp, om_unit.cpp).
// Both end in the same CUDA kernel.
//
// Addresses: the DCRs carry (address >> 6) in 32 bits (draw3d/main.cpp:214-224); CUDA device pointers
// need 48, so a block address is resolved against the live allocations of the device.
#include <stdio.h>
#include <string.h>

#include "raster_tile.cuh"
#include "vx_internal.h"

namespace srt {

// counter slots written by the STATS instantiations of raster_tile_kernel
enum { RCTR_FRAGMENTS = 0, RCTR_TESTS, RCTR_TEXELS, RCTR_OM_READS, RCTR_OM_WRITES, RCTR_RASTER_READS, RCTR_PER_SM = 8,
       RCTR_COUNT = 8 + kMaxSms };

struct RasterState {
  unsigned long long* d_counters = nullptr;  // RCTR_COUNT u64
  bool counters_live = false;                // the last launch counted
  // RASTER_FAULT_* bits set by a tile kernel that met a primitive list or id outside its buffer: one page-locked,
  // device-mapped word, read by the host once the stream is idle (raster_take_fault)
  volatile uint32_t* h_fault = nullptr;
  uint32_t* d_fault = nullptr;
};

namespace {

// Finds the allocation whose 64-byte block address matches `block` in its low 32 bits.
uint8_t* resolve_block(vx_device* dev, uint32_t block, const char* what, uint64_t* size_left,
                       const vx_buffer** owner = nullptr) {
  uint8_t* found = nullptr;
  int matches = 0;
  for (vx_buffer* b = dev->buffers; b; b = b->next) {
    const uint64_t lo = b->reported_addr, hi = b->reported_addr + b->size;
    // candidate address with the buffer's own upper bits
    uint64_t cand = (lo & ~((1ull << 38) - 1)) | ((uint64_t)block << 6);
    if (cand < lo) cand += 1ull << 38;
    if (cand >= lo && cand < hi) {
      found = (uint8_t*)b->dptr + (cand - lo);
      if (size_left) *size_left = hi - cand;
      if (owner) *owner = b;
      ++matches;
    }
  }
  if (matches != 1) {
    fprintf(stderr, "[VXDRV] draw3d: %s block address 0x%x matches %d device buffers\n", what, block, matches);
    return nullptr;
  }
  return found;
}

uint8_t* resolve_addr(vx_device* dev, uint64_t addr, const char* what) {
  for (vx_buffer* b = dev->buffers; b; b = b->next)
    if (addr >= b->reported_addr && addr < b->reported_addr + b->size) return (uint8_t*)b->dptr + (addr - b->reported_addr);
  fprintf(stderr, "[VXDRV] draw3d: %s address 0x%llx is not inside a device buffer\n", what, (unsigned long long)addr);
  return nullptr;
}

uint32_t dev_dcr(vx_device* dev, uint32_t addr) {
  auto it = dev->dcrs.find(addr);
  return it == dev->dcrs.end() ? 0u : it->second;
}

}  // namespace

int raster_launch(vx_device* dev, const draw3d_kernel_arg_t& arg) {
  if (!dev->raster) dev->raster = new RasterState();
  RasterState* st = dev->raster;
  RasterParams P;
  memset(&P, 0, sizeof(P));

  auto ras = [&](uint32_t a) { return arg.sw_rast ? DRAW3D_RASTER_DCR(&arg, a) : dev_dcr(dev, a); };
  auto om = [&](uint32_t a) { return arg.sw_om ? DRAW3D_OM_DCR(&arg, a) : dev_dcr(dev, a); };
  // stage 0 only: draw3d samples stage 0 (kernel.cpp:160-164); device tex DCRs are written with STAGE = 0
  auto tex = [&](uint32_t a) { return arg.sw_tex ? DRAW3D_TEX_DCR(&arg, 0, a) : dev_dcr(dev, a); };

  // ---- raster unit
  P.num_tiles = ras(VX_DCR_RASTER_TILE_COUNT);
  P.prim_stride = ras(VX_DCR_RASTER_PBUF_STRIDE);
  P.tile_format = dev_dcr(dev, VX_DCR_RASTER_EXT_TBUF_FORMAT);
  P.scissor_left = ras(VX_DCR_RASTER_SCISSOR_X) & 0xffff;
  P.scissor_right = ras(VX_DCR_RASTER_SCISSOR_X) >> 16;
  P.scissor_top = ras(VX_DCR_RASTER_SCISSOR_Y) & 0xffff;
  P.scissor_bottom = ras(VX_DCR_RASTER_SCISSOR_Y) >> 16;
  dev->last_launches = 0;
  dev->last_raster_tiles = P.num_tiles;
  st->counters_live = false;
  if (P.num_tiles == 0) return 0;
  if (P.tile_format > SRT_TILEBUF_WIDE || P.prim_stride < sizeof(rast_prim_t) || (P.prim_stride & 3)) {
    fprintf(stderr, "[VXDRV] draw3d: unsupported tile format %u / primitive stride %u\n", P.tile_format, P.prim_stride);
    return -1;
  }
  uint64_t tb_left = 0, pb_left = 0;
  const vx_buffer* tb_owner = nullptr;
  P.tilebuf = resolve_block(dev, ras(VX_DCR_RASTER_TBUF_ADDR), "tile buffer", &tb_left, &tb_owner);
  P.primbuf = resolve_block(dev, ras(VX_DCR_RASTER_PBUF_ADDR), "primitive buffer", &pb_left);
  if (!P.tilebuf || !P.primbuf) return -1;
  // The kernel is built for 32 x 32-pixel tiles (RASTER_TILE_LOGSIZE, like the reference's raster unit) and neither
  // tile-buffer format carries the size it was binned with.  What is known is checked: the size vx_cuda_draw3d_binning
  // recorded on the buffer it returned, else the size the app declared in VX_DCR_RASTER_EXT_TILE_LOGSIZE (0 = undeclared).
  const uint32_t binned = tb_owner->tile_logsize ? tb_owner->tile_logsize : dev_dcr(dev, VX_DCR_RASTER_EXT_TILE_LOGSIZE);
  if (binned != 0 && binned != RASTER_TILE_LOGSIZE) {
    fprintf(stderr, "[VXDRV] draw3d: the tile buffer was binned with 2^%u-pixel tiles; this back-end renders 2^%u only\n",
            binned, (unsigned)RASTER_TILE_LOGSIZE);
    return -1;
  }
  P.tilebuf_bytes = tb_left;
  P.prim_limit = (uint32_t)(pb_left / P.prim_stride > 0xffffffffull ? 0xffffffffull : pb_left / P.prim_stride);
  if (P.prim_limit == 0) {
    fprintf(stderr, "[VXDRV] draw3d: the primitive buffer holds no whole record of %u bytes\n", P.prim_stride);
    return -1;
  }
  if (!st->h_fault) {
    uint32_t* h = nullptr;
    VX_CUDA_OK(cudaHostAlloc((void**)&h, sizeof(uint32_t), cudaHostAllocMapped));
    *h = 0;
    if (cudaHostGetDevicePointer((void**)&st->d_fault, h, 0) != cudaSuccess) {
      cudaFreeHost(h);
      cudaGetLastError();
      return -1;
    }
    st->h_fault = h;
  }
  P.fault = st->d_fault;
  const uint64_t hdr = P.tile_format == SRT_TILEBUF_WIDE ? sizeof(srt_tile_header_wide_t) : sizeof(rast_tile_header_t);
  if ((uint64_t)P.num_tiles * hdr > tb_left) {
    fprintf(stderr, "[VXDRV] draw3d: %u tile headers do not fit the tile buffer\n", P.num_tiles);
    return -1;
  }
  // the shader reads attributes through kernel_arg_t.prim_addr (kernel.cpp:241-242), the raster unit reads
  // edges through PBUF_ADDR (gpu_sw.h:50-56); one staged copy serves both, so they must be the same buffer
  if (arg.prim_addr != 0) {
    uint8_t* pa = resolve_addr(dev, arg.prim_addr, "prim_addr");
    if (!pa) return -1;
    if (pa != P.primbuf) {
      fprintf(stderr, "[VXDRV] draw3d: kernel_arg.prim_addr and VX_DCR_RASTER_PBUF_ADDR name different buffers\n");
      return -1;
    }
  }

  // ---- output merger
  const uint32_t wm = om(VX_DCR_OM_CBUF_WRITEMASK) & 0xf;
  P.cbuf_writemask = (((wm >> 0) & 1) * 0x000000ffu) | (((wm >> 1) & 1) * 0x0000ff00u) | (((wm >> 2) & 1) * 0x00ff0000u) |
                     (((wm >> 3) & 1) * 0xff000000u);
  P.color_read = (wm != 0xf);
  P.color_write = (wm != 0x0);
  P.cbuf_pitch = om(VX_DCR_OM_CBUF_PITCH);
  P.zbuf_pitch = om(VX_DCR_OM_ZBUF_PITCH);
  P.depth_writemask = om(VX_DCR_OM_DEPTH_WRITEMASK) & 0x1;
  P.stencil_writemask = om(VX_DCR_OM_STENCIL_WRITEMASK) & 0xffff;  // front face: draw3d passes face 0
  P.depth_func = om(VX_DCR_OM_DEPTH_FUNC);
  P.s_func = om(VX_DCR_OM_STENCIL_FUNC) & 0xffff;
  P.s_zpass = om(VX_DCR_OM_STENCIL_ZPASS) & 0xffff;
  P.s_zfail = om(VX_DCR_OM_STENCIL_ZFAIL) & 0xffff;
  P.s_fail = om(VX_DCR_OM_STENCIL_FAIL) & 0xffff;
  P.s_ref = om(VX_DCR_OM_STENCIL_REF) & 0xffff;
  P.s_mask = om(VX_DCR_OM_STENCIL_MASK) & 0xffff;
  P.om_depth_enabled = !((P.depth_func == VX_OM_DEPTH_FUNC_ALWAYS) && !P.depth_writemask);
  P.om_stencil_enabled = !((P.s_func == VX_OM_DEPTH_FUNC_ALWAYS) && (P.s_zpass == VX_OM_STENCIL_OP_KEEP) &&
                           (P.s_zfail == VX_OM_STENCIL_OP_KEEP));
  P.bl_mode_rgb = om(VX_DCR_OM_BLEND_MODE) & 0xffff;
  P.bl_mode_a = om(VX_DCR_OM_BLEND_MODE) >> 16;
  P.bl_src_rgb = (om(VX_DCR_OM_BLEND_FUNC) >> 0) & 0xff;
  P.bl_src_a = (om(VX_DCR_OM_BLEND_FUNC) >> 8) & 0xff;
  P.bl_dst_rgb = (om(VX_DCR_OM_BLEND_FUNC) >> 16) & 0xff;
  P.bl_dst_a = (om(VX_DCR_OM_BLEND_FUNC) >> 24) & 0xff;
  P.bl_const = om(VX_DCR_OM_BLEND_CONST);
  P.logic_op = om(VX_DCR_OM_LOGIC_OP);
  P.blend_enabled = !((P.bl_mode_rgb == VX_OM_BLEND_MODE_ADD) && (P.bl_mode_a == VX_OM_BLEND_MODE_ADD) &&
                      (P.bl_src_rgb == VX_OM_BLEND_FUNC_ONE) && (P.bl_src_a == VX_OM_BLEND_FUNC_ONE) &&
                      (P.bl_dst_rgb == VX_OM_BLEND_FUNC_ZERO) && (P.bl_dst_a == VX_OM_BLEND_FUNC_ZERO));
  if (P.depth_func > 7 || P.s_func > 7 || P.s_zpass > 7 || P.s_zfail > 7 || P.s_fail > 7 || P.bl_mode_rgb > 5 ||
      P.bl_mode_a > 5 || P.bl_src_rgb > 14 || P.bl_src_a > 14 || P.bl_dst_rgb > 14 || P.bl_dst_a > 14 || P.logic_op > 15) {
    fprintf(stderr, "[VXDRV] draw3d: output-merger state out of range\n");  // the reference asserts
    return -1;
  }
  uint64_t cb_left = 0, zb_left = 0;
  if (P.color_write) {
    P.cbuf = resolve_block(dev, om(VX_DCR_OM_CBUF_ADDR), "colour buffer", &cb_left);
    if (!P.cbuf) return -1;
    if (P.scissor_bottom && ((uint64_t)(P.scissor_bottom - 1) * P.cbuf_pitch + (uint64_t)P.scissor_right * 4 > cb_left)) {
      fprintf(stderr, "[VXDRV] draw3d: scissor rectangle exceeds the colour buffer\n");
      return -1;
    }
  }
  if (P.om_depth_enabled || P.om_stencil_enabled) {
    P.zbuf = resolve_block(dev, om(VX_DCR_OM_ZBUF_ADDR), "depth buffer", &zb_left);
    if (!P.zbuf) return -1;
    if (P.scissor_bottom && ((uint64_t)(P.scissor_bottom - 1) * P.zbuf_pitch + (uint64_t)P.scissor_right * 4 > zb_left)) {
      fprintf(stderr, "[VXDRV] draw3d: scissor rectangle exceeds the depth buffer\n");
      return -1;
    }
  }

  // ---- shader switches + texture unit
  P.sh_depth = arg.depth_enabled != 0;
  P.sh_color = arg.color_enabled != 0;
  P.sh_tex = arg.tex_enabled != 0;
  P.sh_modulate = arg.tex_modulate != 0;
  if (P.sh_tex) {
    const uint32_t logdim = tex(VX_DCR_TEX_LOGDIM), wrap = tex(VX_DCR_TEX_WRAP);
    P.tex_log_width = logdim & 0xffff;
    P.tex_log_height = logdim >> 16;
    P.tex_format = tex(VX_DCR_TEX_FORMAT);
    P.tex_filter = tex(VX_DCR_TEX_FILTER);
    P.tex_wrapu = wrap & 0xffff;
    P.tex_wrapv = wrap >> 16;
    if (P.tex_log_width > VX_TEX_LOD_MAX || P.tex_log_height > VX_TEX_LOD_MAX || P.tex_format > VX_TEX_FORMAT_A8 ||
        P.tex_filter > VX_TEX_FILTER_BILINEAR || P.tex_wrapu > VX_TEX_WRAP_MIRROR || P.tex_wrapv > VX_TEX_WRAP_MIRROR) {
      fprintf(stderr, "[VXDRV] draw3d: texture state out of range\n");
      return -1;
    }
    uint64_t tx_left = 0;
    uint8_t* tb = resolve_block(dev, tex(VX_DCR_TEX_ADDR), "texture", &tx_left);
    if (!tb) return -1;
    const uint32_t mip0 = tex(VX_DCR_TEX_MIPOFF(0));
    const uint32_t stride = P.tex_format == VX_TEX_FORMAT_A8R8G8B8 ? 4 : (P.tex_format <= VX_TEX_FORMAT_A8L8 ? 2 : 1);
    if (mip0 + ((uint64_t)stride << (P.tex_log_width + P.tex_log_height)) > tx_left || (mip0 % stride)) {
      fprintf(stderr, "[VXDRV] draw3d: texture level 0 does not fit its buffer\n");
      return -1;
    }
    P.texbuf = tb + mip0;
  }

  // pick the instantiation whose compile-time state equals the decoded state (raster_tile.cuh)
  int texk = 0, blendk = 0;
  if (P.sh_tex) {
    texk = 1;
    if (P.tex_filter == VX_TEX_FILTER_BILINEAR && P.tex_wrapu == VX_TEX_WRAP_REPEAT && P.tex_wrapv == VX_TEX_WRAP_REPEAT) {
      if (P.tex_format == VX_TEX_FORMAT_A8R8G8B8) texk = 2;
      else if (P.tex_format == VX_TEX_FORMAT_R5G6B5) texk = 3;
      else if (P.tex_format == VX_TEX_FORMAT_A8L8) texk = 4;
    }
  }
  if (P.blend_enabled) {
    blendk = 1;
    if (P.bl_mode_rgb == VX_OM_BLEND_MODE_ADD && P.bl_mode_a == VX_OM_BLEND_MODE_ADD && P.bl_src_rgb == VX_OM_BLEND_FUNC_SRC_A &&
        P.bl_src_a == VX_OM_BLEND_FUNC_SRC_A && P.bl_dst_rgb == VX_OM_BLEND_FUNC_ONE_MINUS_SRC_A &&
        P.bl_dst_a == VX_OM_BLEND_FUNC_ONE_MINUS_SRC_A)
      blendk = 2;
  }
  if (dev_dcr(dev, VX_DCR_RASTER_EXT_GENERIC) != 0) {  // A/B switch: always the generic instantiation
    texk = texk ? 1 : 0;
    blendk = blendk ? 1 : 0;
  }
  const bool stats = vx_dev_counting(dev);
  if (stats) {
    if (!st->d_counters) VX_CUDA_OK(cudaMalloc(&st->d_counters, RCTR_COUNT * sizeof(unsigned long long)));
    VX_CUDA_OK(cudaMemsetAsync(st->d_counters, 0, RCTR_COUNT * sizeof(unsigned long long), dev->stream));
    P.counters = st->d_counters;
    st->counters_live = true;
  }
  // one thread per pixel when the draw has too few tiles to fill the GPU with quad-per-thread CTAs
  // (4 CTAs of 256 threads fit an SM: below 4 x #SM tiles the GPU is not full); VX_DCR_RASTER_EXT_PPT forces 4 or 1 for A/B runs
  uint32_t ppt = dev_dcr(dev, VX_DCR_RASTER_EXT_PPT);
  if (ppt != 1 && ppt != 4) ppt = (P.num_tiles < 4u * (uint32_t)dev->prop.multiProcessorCount) ? 1 : 4;
  if (stats) ppt = 4;  // the quad-test counter is defined per quad
  using kernel_t = void (*)(const RasterParams);
#define SRT_ROW(S, T, N) {raster_tile_kernel<S, T, 0, N>, raster_tile_kernel<S, T, 1, N>, raster_tile_kernel<S, T, 2, N>}
  static const kernel_t table[3][5][3] = {
      {SRT_ROW(false, 0, 4), SRT_ROW(false, 1, 4), SRT_ROW(false, 2, 4), SRT_ROW(false, 3, 4), SRT_ROW(false, 4, 4)},
      {SRT_ROW(true, 0, 4), SRT_ROW(true, 1, 4), SRT_ROW(true, 2, 4), SRT_ROW(true, 3, 4), SRT_ROW(true, 4, 4)},
      {SRT_ROW(false, 0, 1), SRT_ROW(false, 1, 1), SRT_ROW(false, 2, 1), SRT_ROW(false, 3, 1), SRT_ROW(false, 4, 1)}};
#undef SRT_ROW
  table[stats ? 1 : (ppt == 1 ? 2 : 0)][texk][blendk]<<<P.num_tiles, ppt == 1 ? 1024 : kRasterThreads, 0, dev->stream>>>(P);
  VX_CUDA_OK(cudaGetLastError());
  dev->last_launches = 1;
  return 0;
}

int raster_read_counters(vx_device* dev) {
  RasterState* st = dev->raster;
  dev->last_raster_fragments = dev->last_raster_tests = 0;
  memset(dev->last_raster_cls, 0, sizeof(dev->last_raster_cls));
  memset(dev->last_instrs_sm, 0, sizeof(dev->last_instrs_sm));
  if (!st || !st->counters_live) return 0;
  unsigned long long c[RCTR_COUNT];
  VX_CUDA_OK(cudaMemcpyAsync(c, st->d_counters, sizeof(c), cudaMemcpyDeviceToHost, dev->stream));
  VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
  dev->last_raster_fragments = c[RCTR_FRAGMENTS];
  dev->last_raster_tests = c[RCTR_TESTS];
  dev->last_raster_cls[0] = c[RCTR_TEXELS];
  dev->last_raster_cls[1] = c[RCTR_OM_READS];
  dev->last_raster_cls[2] = c[RCTR_OM_WRITES];
  dev->last_raster_cls[3] = c[RCTR_RASTER_READS];
  for (int i = 0; i < kMaxSms; ++i) dev->last_instrs_sm[i] = c[RCTR_PER_SM + i];
  return 0;
}

// Called when the launch stream has drained: reports and clears what the tile kernels flagged since the last call.
int raster_take_fault(vx_device* dev) {
  RasterState* st = dev->raster;
  if (!st || !st->h_fault) return 0;
  const uint32_t bits = *st->h_fault;
  if (bits == 0) return 0;
  *st->h_fault = 0;
  if (bits & RASTER_FAULT_LIST) fprintf(stderr, "[VXDRV] draw3d: a tile's primitive list lies outside the tile buffer (tile skipped)\n");
  if (bits & RASTER_FAULT_PRIM) fprintf(stderr, "[VXDRV] draw3d: a primitive id lies outside the primitive buffer (record 0 drawn instead)\n");
  return -1;
}

void raster_destroy(vx_device* dev) {
  if (!dev->raster) return;
  if (dev->raster->h_fault) cudaFreeHost(const_cast<uint32_t*>(dev->raster->h_fault));
  cudaFree(dev->raster->d_counters);
  delete dev->raster;
  dev->raster = nullptr;
}

}  // namespace srt
