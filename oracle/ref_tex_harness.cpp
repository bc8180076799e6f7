// ref_tex_harness.cpp -- TEST INFRASTRUCTURE ONLY.  Runs the REFERENCE's own texture-unit regression app natively on
// the host cores, compiled from the sources where they lie under /root/reference:
//   tests/regression/tex/kernel.cpp (the kernel, software sampler path `use_sw`), sim/common/graphics.cpp
//   (graphics::TextureSampler), third_party/cocogfx/src/blitter.cpp (ConvertImage, GenerateMipmaps)
// (see oracle/Makefile, target _ref/libref_tex.so).  This file is the only non-reference code in that library: it
// stands in for libpng (the caller passes decoded pixels) and for the Vortex device (a low-address host arena for
// the texture and the destination image; vx_spawn_threads runs the tasks in order on the calling thread).  The host
// flow follows tests/regression/tex/main.cpp:153-252 statement by statement.
//
// Used by tests/golden/make_tex_fixtures.py (in the build container) to turn the reference's input PNGs into the
// texture images its app uploads and to render vectors its golden PNGs do not cover, and by tests/test_oracle_tex.py
// when the library is present.  Never loaded by the product.
#include <assert.h>
#include <stdint.h>
#include <string.h>
#include <sys/mman.h>

#include <algorithm>
#include <vector>

#include <bitmanip.h>
#include <graphics.h>
#include <vx_spawn.h>

#include <cocogfx/include/blitter.hpp>
#include <cocogfx/include/format.hpp>

#include "common.h"  // tests/regression/tex/common.h : kernel_arg_t

using namespace cocogfx;

extern "C" {
uintptr_t ref_shim_mscratch = 0;
__thread dim3_t blockIdx;

// kernel/src/vx_spawn.c:54-81 : one call of kernel_func per task id, blockIdx.x = task id
int vx_spawn_threads(uint32_t dimension, const uint32_t* grid_dim, const uint32_t*, vx_kernel_func_cb kernel_func,
                     const void* arg) {
  assert(dimension == 1);
  (void)dimension;
  for (uint32_t t = 0; t < grid_dim[0]; ++t) {
    blockIdx.x = t;
    kernel_func(const_cast<void*>(arg));
  }
  return 0;
}
}

int ref_tex_kernel_main();  // tests/regression/tex/kernel.cpp:106 (compiled with -Dmain=ref_tex_kernel_main)

namespace {

// device memory stand-in: VX_DCR_TEX_ADDR carries addr / 64 in 32 bits (tex/main.cpp:248) => below 2^38
constexpr uintptr_t kArenaBase = 0x200000000ull;
constexpr size_t kArenaSize = 256ull << 20;
uint8_t* g_arena = nullptr;

uint8_t* arena() {
  if (!g_arena) {
    void* p = mmap((void*)kArenaBase, kArenaSize, PROT_READ | PROT_WRITE,
                   MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE | MAP_FIXED_NOREPLACE, -1, 0);
    if (p == MAP_FAILED || (uintptr_t)p != kArenaBase) return nullptr;
    g_arena = (uint8_t*)p;
  }
  return g_arena;
}

bool tex_format(uint32_t format, ePixelFormat* e) {  // tex/main.cpp:77-92
  switch (format) {
    case VX_TEX_FORMAT_A8R8G8B8: *e = FORMAT_A8R8G8B8; return true;
    case VX_TEX_FORMAT_R5G6B5: *e = FORMAT_R5G6B5; return true;
    case VX_TEX_FORMAT_A1R5G5B5: *e = FORMAT_A1R5G5B5; return true;
    case VX_TEX_FORMAT_A4R4G4B4: *e = FORMAT_A4R4G4B4; return true;
    case VX_TEX_FORMAT_A8L8: *e = FORMAT_A8L8; return true;
    case VX_TEX_FORMAT_L8: *e = FORMAT_L8; return true;
    case VX_TEX_FORMAT_A8: *e = FORMAT_A8; return true;
  }
  return false;
}

}  // namespace

extern "C" {

// LoadImage's conversion (imageutil.cpp:62-84: bytes per pixel of the decoded file -> source format) followed by
// GenerateMipmaps (tex/main.cpp:165-178).  `pixels` = the decoded image in the loader's byte order (BGR / BGRA,
// png.cpp:78), `channels` = 1 / 3 / 4.  out / capacity: the mip chain; mip_offsets: up to max_mips entries.
// Returns the size of the chain in bytes, or -1.
int64_t ref_tex_make_texture(const uint8_t* pixels, uint32_t channels, uint32_t width, uint32_t height, uint32_t format,
                             uint8_t* out, uint64_t capacity, uint32_t* mip_offsets, uint32_t max_mips,
                             uint32_t* num_mips) {
  ePixelFormat eformat;
  if (!pixels || !out || !mip_offsets || !num_mips || !tex_format(format, &eformat)) return -1;
  ePixelFormat img_format;
  switch (channels) {
    case 1: img_format = FORMAT_A8; break;
    case 3: img_format = FORMAT_R8G8B8; break;
    case 4: img_format = FORMAT_A8R8G8B8; break;
    default: return -1;
  }
  std::vector<uint8_t> staging(pixels, pixels + (size_t)width * height * channels);
  if (img_format != eformat) {
    std::vector<uint8_t> converted;
    if (ConvertImage(converted, eformat, staging.data(), img_format, width, height, width * channels) != 0) return -1;
    staging.swap(converted);
  }
  const uint32_t bpp = Format::GetInfo(eformat).BytePerPixel;
  std::vector<uint8_t> chain;
  std::vector<uint32_t> offsets;
  if (GenerateMipmaps(chain, offsets, staging.data(), eformat, width, height, width * bpp) != 0) return -1;
  if (chain.size() > capacity || offsets.size() > max_mips) return -1;
  memcpy(out, chain.data(), chain.size());
  for (size_t i = 0; i < offsets.size(); ++i) mip_offsets[i] = offsets[i];
  *num_mips = (uint32_t)offsets.size();
  return (int64_t)chain.size();
}

// tex/main.cpp:180-252 + the kernel: sample the texture into a dst_width x dst_height A8R8G8B8 image (row 0 first,
// as the app saves it, :147).  wrap / filter are the app's -w / -g values; num_tasks its cores x warps x threads.
int ref_tex_render(const uint8_t* texture, uint64_t texture_size, const uint32_t* mip_offsets, uint32_t num_mips,
                   uint32_t src_width, uint32_t src_height, uint32_t format, uint32_t wrap, uint32_t filter,
                   uint32_t dst_width, uint32_t dst_height, uint32_t num_tasks, uint32_t* out) {
  uint8_t* base = arena();
  if (!base || !texture || !out || !ispow2(src_width) || !ispow2(src_height) || dst_width == 0 || dst_height == 0)
    return -1;
  const uint64_t dst_size = 4ull * dst_width * dst_height;
  const uint64_t dst_off = (texture_size + 63) & ~63ull;
  if (dst_off + dst_size > kArenaSize) return -1;
  uint8_t* src = base;
  uint8_t* dst = base + dst_off;
  memcpy(src, texture, texture_size);
  memset(dst, 0, dst_size);

  kernel_arg_t kernel_arg = {};
  kernel_arg.use_sw = true;  // -z: graphics::TextureSampler instead of the vx_tex instruction
  kernel_arg.num_tasks = std::min<uint32_t>(num_tasks ? num_tasks : 1, dst_height);
  kernel_arg.dst_width = dst_width;
  kernel_arg.dst_height = dst_height;
  kernel_arg.dst_stride = 4;
  kernel_arg.dst_pitch = 4 * dst_width;
  kernel_arg.dst_addr = (uint64_t)(uintptr_t)dst;
#define TEX_DCR_WRITE(addr, value) kernel_arg.dcrs.write(addr, value)  // the kernel_arg half of the app's macro (:122-124)
  TEX_DCR_WRITE(VX_DCR_TEX_STAGE, 0);
  TEX_DCR_WRITE(VX_DCR_TEX_LOGDIM, (log2ceil(src_height) << 16) | log2ceil(src_width));
  TEX_DCR_WRITE(VX_DCR_TEX_FORMAT, format);
  TEX_DCR_WRITE(VX_DCR_TEX_WRAP, (wrap << 16) | wrap);
  TEX_DCR_WRITE(VX_DCR_TEX_FILTER, (filter ? VX_TEX_FILTER_BILINEAR : VX_TEX_FILTER_POINT));
  TEX_DCR_WRITE(VX_DCR_TEX_ADDR, (uint32_t)((uintptr_t)src / 64));
  for (uint32_t i = 0; i < num_mips; ++i) {
    if (i >= VX_TEX_LOD_MAX) return -1;
    TEX_DCR_WRITE(VX_DCR_TEX_MIPOFF(i), mip_offsets[i]);
  }
#undef TEX_DCR_WRITE
  ref_shim_mscratch = (uintptr_t)&kernel_arg;
  if (ref_tex_kernel_main() != 0) return -1;
  memcpy(out, dst, dst_size);
  return 0;
}

}  // extern "C"
