// ref_draw3d_harness.cpp -- TEST INFRASTRUCTURE ONLY.  Runs the REFERENCE's own raster hot path
// natively on the host cores, compiled from the sources where they lie under /root/reference:
//   sim/common/graphics.cpp, sim/common/gfxutil.cpp, tests/regression/draw3d/kernel.cpp (+gpu_sw.h),
//   third_party/cocogfx/src/{blitter,format}.cpp
// (see oracle/Makefile, target _ref/libref_draw3d.so).  This file is the only non-reference code in
// that library: it replaces the pieces that cannot build here -- the Boost .cgltrace loader (the flat
// SKBXTRC1 blob of tests/golden/make_raster_fixtures.py is read instead), libpng, and the Vortex device
// (a low-address host arena stands in for device memory; vx_spawn_threads becomes a host thread pool).
// The per-draw host flow below follows tests/regression/draw3d/main.cpp:171-389 statement by
// statement, including its quirks (STENCIL_ZPASS written twice :259-260, wrapV taken from addressU
// :304, the filter test :301-302).
//
// Only tests/ and bench.py's cpu_baseline / --impl reference legs may load the resulting library.
#include <assert.h>
#include <stdint.h>
#include <string.h>
#include <sys/mman.h>

#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

#include <bitmanip.h>
#include <gfxutil.h>
#include <graphics.h>
#include <vx_spawn.h>

#include <cocogfx/include/blitter.hpp>

#include "common.h"  // tests/regression/draw3d/common.h : kernel_arg_t

using namespace cocogfx;

extern "C" {
uintptr_t ref_shim_mscratch = 0;
__thread dim3_t blockIdx;
}
static uint32_t g_pool_threads = 1;

// kernel/src/vx_spawn.c:54-81 : one call of kernel_func per task id, blockIdx.x = task id
extern "C" int vx_spawn_threads(uint32_t dimension, const uint32_t* grid_dim, const uint32_t*,
                                vx_kernel_func_cb kernel_func, const void* arg) {
  assert(dimension == 1);
  (void)dimension;
  const uint32_t num_tasks = grid_dim[0];
  std::atomic<uint32_t> next(0);
  auto worker = [&]() {
    for (;;) {
      uint32_t t = next.fetch_add(1);
      if (t >= num_tasks) break;
      blockIdx.x = t;
      kernel_func(const_cast<void*>(arg));
    }
  };
  std::vector<std::thread> pool;
  for (uint32_t i = 1; i < g_pool_threads; ++i) pool.emplace_back(worker);
  worker();
  for (auto& t : pool) t.join();
  return 0;
}

int ref_kernel_main();  // tests/regression/draw3d/kernel.cpp:286 (compiled with -Dmain=ref_kernel_main)

namespace {

// ---- device memory stand-in: DCRs carry addr/64 in 32 bits (gpu_sw.h:28-29,83-89) => below 2^38
constexpr uintptr_t kArenaBase = 0x100000000ull;
constexpr size_t kArenaSize = 3ull << 30;
uint8_t* g_arena = nullptr;
size_t g_arena_used = 0;

uint8_t* arena_alloc(size_t size) {
  if (!g_arena) {
    void* p = mmap((void*)kArenaBase, kArenaSize, PROT_READ | PROT_WRITE,
                   MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE | MAP_FIXED_NOREPLACE, -1, 0);
    if (p == MAP_FAILED || (uintptr_t)p != kArenaBase) return nullptr;
    g_arena = (uint8_t*)p;
  }
  size_t off = (g_arena_used + 63) & ~size_t(63);
  if (off + size > kArenaSize) return nullptr;
  g_arena_used = off + size;
  return g_arena + off;
}

struct Reader {
  const uint8_t* p;
  const uint8_t* end;
  bool ok = true;
  template <typename T> T get() {
    T v{};
    if (p + sizeof(T) > end) { ok = false; return v; }
    memcpy(&v, p, sizeof(T));
    p += sizeof(T);
    return v;
  }
};

bool load_trace(CGLTrace& trace, const uint8_t* blob, uint64_t size) {
  Reader r{blob, blob + size};
  if (size < 16 || memcmp(blob, "SKBXTRC1", 8) != 0) return false;
  r.p += 8;
  uint32_t nd = r.get<uint32_t>(), nt = r.get<uint32_t>();
  for (uint32_t i = 0; i < nt && r.ok; ++i) {
    uint32_t id = r.get<uint32_t>();
    CGLTrace::texture_t t;
    t.format = (ePixelFormat)r.get<uint32_t>();
    t.width = r.get<uint32_t>();
    t.height = r.get<uint32_t>();
    uint32_t sz = r.get<uint32_t>();
    if (r.p + sz > r.end) return false;
    t.pixels.assign(r.p, r.p + sz);
    r.p += (sz + 3) & ~3u;
    trace.textures[id] = t;
  }
  for (uint32_t i = 0; i < nd && r.ok; ++i) {
    CGLTrace::drawcall_t d;
    auto& s = d.states;
    uint32_t f[24];
    for (auto& x : f) x = r.get<uint32_t>();
    s.color_enabled = f[0]; s.color_format = (ePixelFormat)f[1]; s.color_writemask = f[2];
    s.depth_test = f[3]; s.depth_writemask = f[4]; s.depth_format = (ePixelFormat)f[5];
    s.depth_func = (CGLTrace::ecompare)f[6];
    s.stencil_test = f[7]; s.stencil_func = (CGLTrace::ecompare)f[8];
    s.stencil_zpass = (CGLTrace::eStencilOp)f[9]; s.stencil_zfail = (CGLTrace::eStencilOp)f[10];
    s.stencil_fail = (CGLTrace::eStencilOp)f[11];
    s.stencil_ref = f[12]; s.stencil_mask = f[13]; s.stencil_writemask = f[14];
    s.texture_enabled = f[15]; s.texture_envmode = (CGLTrace::eEnvMode)f[16];
    s.texture_minfilter = (CGLTrace::eTexFilter)f[17]; s.texture_magfilter = (CGLTrace::eTexFilter)f[18];
    s.texture_addressU = (CGLTrace::eTexAddress)f[19]; s.texture_addressV = (CGLTrace::eTexAddress)f[20];
    s.blend_enabled = f[21]; s.blend_src = (CGLTrace::eBlendOp)f[22]; s.blend_dst = (CGLTrace::eBlendOp)f[23];
    s.texture_envcolor.r = r.get<float>(); s.texture_envcolor.g = r.get<float>();
    s.texture_envcolor.b = r.get<float>(); s.texture_envcolor.a = r.get<float>();
    d.texture_id = r.get<uint32_t>();
    d.viewport.left = r.get<int32_t>(); d.viewport.right = r.get<int32_t>();
    d.viewport.top = r.get<int32_t>(); d.viewport.bottom = r.get<int32_t>();
    d.viewport.near = r.get<float>(); d.viewport.far = r.get<float>();
    uint32_t nv = r.get<uint32_t>(), np = r.get<uint32_t>();
    for (uint32_t k = 0; k < nv && r.ok; ++k) {
      CGLTrace::vertex_t v;
      v.pos.x = r.get<float>(); v.pos.y = r.get<float>(); v.pos.z = r.get<float>(); v.pos.w = r.get<float>();
      v.color.r = r.get<float>(); v.color.g = r.get<float>(); v.color.b = r.get<float>(); v.color.a = r.get<float>();
      v.texcoord.u = r.get<float>(); v.texcoord.v = r.get<float>();
      d.vertices[k] = v;
    }
    d.primitives.resize(np);
    for (uint32_t k = 0; k < np && r.ok; ++k) {
      d.primitives[k].i0 = r.get<uint32_t>();
      d.primitives[k].i1 = r.get<uint32_t>();
      d.primitives[k].i2 = r.get<uint32_t>();
    }
    trace.drawcalls.push_back(std::move(d));
  }
  return r.ok;
}

}  // namespace

#define RASTER_DCR_WRITE(addr, value) kernel_arg.raster_dcrs.write(addr, value)
#define OM_DCR_WRITE(addr, value) kernel_arg.om_dcrs.write(addr, value)
#define TEX_DCR_WRITE(addr, value) kernel_arg.tex_dcrs.write(addr, value)

extern "C" {

uint32_t ref_draw3d_kernel_arg_size() { return (uint32_t)sizeof(kernel_arg_t); }
uint32_t ref_draw3d_prim_size() { return (uint32_t)sizeof(graphics::rast_prim_t); }

// graphics::Binning of one draw call (sim/common/gfxutil.cpp:103-276).  Buffers are caller-sized
// (*_cap bytes); returns the tile count, or -1 when a buffer is too small / the blob is bad.
int64_t ref_draw3d_binning(const uint8_t* blob, uint64_t blob_size, uint32_t draw, uint32_t width,
                           uint32_t height, uint32_t tile_logsize, uint8_t* tilebuf_out,
                           uint64_t* tilebuf_size, uint8_t* primbuf_out, uint64_t* primbuf_size) {
  CGLTrace trace;
  if (!load_trace(trace, blob, blob_size) || draw >= trace.drawcalls.size()) return -1;
  auto& dc = trace.drawcalls[draw];
  std::vector<uint8_t> tilebuf, primbuf;
  auto n = graphics::Binning(tilebuf, primbuf, dc.vertices, dc.primitives, width, height,
                             dc.viewport.near, dc.viewport.far, tile_logsize);
  if (tilebuf.size() > *tilebuf_size || primbuf.size() > *primbuf_size) return -1;
  memcpy(tilebuf_out, tilebuf.data(), tilebuf.size());
  memcpy(primbuf_out, primbuf.data(), primbuf.size());
  *tilebuf_size = tilebuf.size();
  *primbuf_size = primbuf.size();
  return n;
}

// Test hook (no counterpart in main.cpp, which always programs MODE_ADD): the four OM DCR values used for every
// draw whose trace state enables blending -- blend mode, blend func, blend constant, logic op.  ov == NULL clears it.
static uint32_t g_om_override[4];
static bool g_om_override_on = false;
void ref_draw3d_set_om_override(const uint32_t* ov) {
  g_om_override_on = ov != nullptr;
  if (ov) for (int i = 0; i < 4; ++i) g_om_override[i] = ov[i];
}

// Whole-trace render, the flow of draw3d/main.cpp:392-517 + render() :171-389 with -x -y -u.
// cbuf_out / zbuf_out: width*height u32, row 0 = framebuffer row 0 (the saved PNG is this flipped).
// kernel_seconds (optional) = time inside the kernel launches only (main.cpp:349-361 measures the same span).
int ref_draw3d_render(const uint8_t* blob, uint64_t blob_size, uint32_t dst_width, uint32_t dst_height,
                      uint32_t num_threads, uint32_t tile_logsize, uint32_t* cbuf_out, uint32_t* zbuf_out,
                      double* kernel_seconds) {
  CGLTrace trace;
  if (!load_trace(trace, blob, blob_size)) return -1;
  g_pool_threads = num_threads ? num_threads : 1;
  g_arena_used = 0;
  static kernel_arg_t kernel_arg;
  kernel_arg = kernel_arg_t{};
  double ksec = 0.0;

  const uint32_t clear_color = 0xff000000, clear_depth = 0xffffffff;
  const uint32_t zbuf_pitch = dst_width * 4, cbuf_pitch = dst_width * 4;
  const size_t zbuf_size = (size_t)dst_height * zbuf_pitch, cbuf_size = (size_t)dst_height * cbuf_pitch;
  auto* zbuf = (uint32_t*)arena_alloc(zbuf_size);
  auto* cbuf = (uint32_t*)arena_alloc(cbuf_size);
  auto* argbuf = (kernel_arg_t*)arena_alloc(sizeof(kernel_arg_t));
  if (!zbuf || !cbuf || !argbuf) return -2;
  const uint64_t zbuf_addr = (uint64_t)(uintptr_t)zbuf, cbuf_addr = (uint64_t)(uintptr_t)cbuf;
  for (size_t i = 0; i < zbuf_size / 4; ++i) zbuf[i] = clear_depth;
  for (size_t i = 0; i < cbuf_size / 4; ++i) cbuf[i] = clear_color;

  uint32_t num_tasks = 64;  // cores*warps*threads in the reference; only strides the tile loop
  while (num_tasks < 4 * g_pool_threads) num_tasks *= 2;
  kernel_arg.log_num_tasks = log2ceil(num_tasks);
  kernel_arg.sw_tex = true;
  kernel_arg.sw_rast = true;
  kernel_arg.sw_om = true;

  for (uint32_t d = 0, nd = trace.drawcalls.size(); d < nd; ++d) {
    auto& drawcall = trace.drawcalls.at(d);
    auto& states = drawcall.states;
    std::vector<uint8_t> tilebuf, primbuf;
    auto num_tiles = graphics::Binning(tilebuf, primbuf, drawcall.vertices, drawcall.primitives, dst_width,
                                       dst_height, drawcall.viewport.near, drawcall.viewport.far, tile_logsize);
    if (0 == num_tiles) continue;
    auto* tile_dev = arena_alloc(tilebuf.size());
    auto* prim_dev = arena_alloc(primbuf.size());
    if (!tile_dev || !prim_dev) return -2;
    memcpy(tile_dev, tilebuf.data(), tilebuf.size());
    memcpy(prim_dev, primbuf.data(), primbuf.size());
    const uint64_t tilebuf_addr = (uint64_t)(uintptr_t)tile_dev, primbuf_addr = (uint64_t)(uintptr_t)prim_dev;

    RASTER_DCR_WRITE(VX_DCR_RASTER_TBUF_ADDR, tilebuf_addr / 64);
    RASTER_DCR_WRITE(VX_DCR_RASTER_TILE_COUNT, num_tiles);
    RASTER_DCR_WRITE(VX_DCR_RASTER_PBUF_ADDR, primbuf_addr / 64);
    RASTER_DCR_WRITE(VX_DCR_RASTER_PBUF_STRIDE, sizeof(graphics::rast_prim_t));
    RASTER_DCR_WRITE(VX_DCR_RASTER_SCISSOR_X, (dst_width << 16) | 0);
    RASTER_DCR_WRITE(VX_DCR_RASTER_SCISSOR_Y, (dst_height << 16) | 0);

    OM_DCR_WRITE(VX_DCR_OM_CBUF_ADDR, cbuf_addr / 64);
    OM_DCR_WRITE(VX_DCR_OM_CBUF_PITCH, cbuf_pitch);
    OM_DCR_WRITE(VX_DCR_OM_CBUF_WRITEMASK, states.color_writemask);
    if (states.depth_test || states.stencil_test) {
      OM_DCR_WRITE(VX_DCR_OM_ZBUF_ADDR, zbuf_addr / 64);
      OM_DCR_WRITE(VX_DCR_OM_ZBUF_PITCH, zbuf_pitch);
    }
    if (states.depth_test) {
      OM_DCR_WRITE(VX_DCR_OM_DEPTH_FUNC, graphics::toVXCompare(states.depth_func));
      OM_DCR_WRITE(VX_DCR_OM_DEPTH_WRITEMASK, states.depth_writemask);
    } else {
      OM_DCR_WRITE(VX_DCR_OM_DEPTH_FUNC, VX_OM_DEPTH_FUNC_ALWAYS);
      OM_DCR_WRITE(VX_DCR_OM_DEPTH_WRITEMASK, 0);
    }
    if (states.stencil_test) {
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_FUNC, graphics::toVXCompare(states.stencil_func));
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_ZPASS, graphics::toVXStencilOp(states.stencil_zpass));
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_ZPASS, graphics::toVXStencilOp(states.stencil_zfail));  // sic, main.cpp:260
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_FAIL, graphics::toVXStencilOp(states.stencil_fail));
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_REF, states.stencil_ref);
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_MASK, states.stencil_mask);
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_WRITEMASK, states.stencil_writemask);
    } else {
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_FUNC, VX_OM_DEPTH_FUNC_ALWAYS);
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_ZPASS, VX_OM_STENCIL_OP_KEEP);
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_ZPASS, VX_OM_STENCIL_OP_KEEP);
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_FAIL, VX_OM_STENCIL_OP_KEEP);
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_REF, 0);
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_MASK, VX_OM_STENCIL_MASK);
      OM_DCR_WRITE(VX_DCR_OM_STENCIL_WRITEMASK, 0);
    }
    if (states.blend_enabled && g_om_override_on) {  // not in main.cpp: OM state the trace format cannot express
      OM_DCR_WRITE(VX_DCR_OM_BLEND_MODE, g_om_override[0]);
      OM_DCR_WRITE(VX_DCR_OM_BLEND_FUNC, g_om_override[1]);
      OM_DCR_WRITE(VX_DCR_OM_BLEND_CONST, g_om_override[2]);
      OM_DCR_WRITE(VX_DCR_OM_LOGIC_OP, g_om_override[3]);
    } else if (states.blend_enabled) {
      auto blend_src = graphics::toVXBlendFunc(states.blend_src);
      auto blend_dst = graphics::toVXBlendFunc(states.blend_dst);
      OM_DCR_WRITE(VX_DCR_OM_BLEND_MODE, (VX_OM_BLEND_MODE_ADD << 16) | (VX_OM_BLEND_MODE_ADD << 0));
      OM_DCR_WRITE(VX_DCR_OM_BLEND_FUNC, (blend_dst << 24) | (blend_dst << 16) | (blend_src << 8) | (blend_src << 0));
    } else {
      OM_DCR_WRITE(VX_DCR_OM_BLEND_MODE, (VX_OM_BLEND_MODE_ADD << 16) | (VX_OM_BLEND_MODE_ADD << 0));
      OM_DCR_WRITE(VX_DCR_OM_BLEND_FUNC, (VX_OM_BLEND_FUNC_ZERO << 24) | (VX_OM_BLEND_FUNC_ZERO << 16) |
                                             (VX_OM_BLEND_FUNC_ONE << 8) | (VX_OM_BLEND_FUNC_ONE << 0));
    }
    if (states.texture_enabled) {
      std::vector<uint8_t> texbuf;
      std::vector<uint32_t> mip_offsets;
      auto& texture = trace.textures.at(drawcall.texture_id);
      auto tex_bpp = Format::GetInfo(texture.format).BytePerPixel;
      auto tex_pitch = texture.width * tex_bpp;
      if (GenerateMipmaps(texbuf, mip_offsets, texture.pixels.data(), texture.format, texture.width,
                          texture.height, tex_pitch) != 0)
        return -3;
      uint32_t tex_logwidth = log2ceil(texture.width);
      uint32_t tex_logheight = log2ceil(texture.height);
      int tex_format = graphics::toVXFormat(texture.format);
      int tex_filter = (states.texture_magfilter != CGLTrace::FILTER_NEAREST) ||
                       (states.texture_magfilter != CGLTrace::FILTER_NEAREST);
      int tex_wrapU = (states.texture_addressU == CGLTrace::ADDRESS_WRAP);
      int tex_wrapV = (states.texture_addressU == CGLTrace::ADDRESS_WRAP);  // sic, main.cpp:304
      auto* tex_dev = arena_alloc(texbuf.size());
      if (!tex_dev) return -2;
      memcpy(tex_dev, texbuf.data(), texbuf.size());
      const uint64_t texbuf_addr = (uint64_t)(uintptr_t)tex_dev;
      TEX_DCR_WRITE(VX_DCR_TEX_STAGE, 0);
      TEX_DCR_WRITE(VX_DCR_TEX_LOGDIM, (tex_logheight << 16) | tex_logwidth);
      TEX_DCR_WRITE(VX_DCR_TEX_FORMAT, tex_format);
      TEX_DCR_WRITE(VX_DCR_TEX_WRAP, (tex_wrapV << 16) | tex_wrapU);
      TEX_DCR_WRITE(VX_DCR_TEX_FILTER, tex_filter ? VX_TEX_FILTER_BILINEAR : VX_TEX_FILTER_POINT);
      TEX_DCR_WRITE(VX_DCR_TEX_ADDR, texbuf_addr / 64);
      for (uint32_t i = 0; i < mip_offsets.size(); ++i) {
        assert(i < VX_TEX_LOD_MAX);
        TEX_DCR_WRITE(VX_DCR_TEX_MIPOFF(i), mip_offsets.at(i));
      }
    }
    kernel_arg.depth_enabled = states.depth_test;
    kernel_arg.color_enabled = states.color_enabled;
    kernel_arg.tex_enabled = states.texture_enabled;
    kernel_arg.tex_modulate = (states.texture_enabled && states.texture_envmode == CGLTrace::ENVMODE_MODULATE);
    kernel_arg.prim_addr = primbuf_addr;
    if (kernel_arg.tex_modulate && !kernel_arg.color_enabled) kernel_arg.tex_modulate = false;
    if (kernel_arg.tex_enabled && kernel_arg.color_enabled && !kernel_arg.tex_modulate)
      kernel_arg.color_enabled = false;

    *argbuf = kernel_arg;  // vx_upload_bytes
    ref_shim_mscratch = (uintptr_t)argbuf;
    auto t0 = std::chrono::high_resolution_clock::now();
    ref_kernel_main();  // vx_start + vx_ready_wait
    ksec += std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count();
  }
  if (cbuf_out) memcpy(cbuf_out, cbuf, cbuf_size);
  if (zbuf_out) memcpy(zbuf_out, zbuf, zbuf_size);
  if (kernel_seconds) *kernel_seconds = ksec;
  return 0;
}

}  // extern "C"
