/*
 * draw3d_flow_check.c -- CPU test of the draw3d app's host flows (apps/draw3d/draw3d_app.c) against a recording
 * stand-in for the vx_* back-end.  TEST INFRASTRUCTURE: it links the app's host code and the host pre-pass unchanged
 * and replaces only libvortex.so, so that the flows that need no pixel -- which buffers are uploaded with which bytes,
 * which DCRs hold what at every vx_start, in which order -- run where there is no GPU, and under ThreadSanitizer.
 *
 * What it checks: SRT_DRAW3D_AUTO (host pre-pass on the app's thread pool, the frame packed into one arena upload,
 * draws started back to back) hands every vx_start exactly what the reference's serial flow hands it: the same unit
 * state (DCR file and kernel_arg_t mirrors, addresses aside), the same tile headers and primitive lists, the same
 * primitive records, the same texels.  Frames are repeated so that the pool is reused and the arena recycled.
 *
 *   draw3d_flow_check <trace.trc.gz> <size> <frames>      exit code 0 = equal; prints one line per flow
 *
 * Built and run by tests/test_draw3d_flows.py (plain, and with -fsanitize=thread where the tool chain has it).
 */
#define _DEFAULT_SOURCE
#define _POSIX_C_SOURCE 200809L
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>

#include "../../skybox_rt_b200/apps/draw3d/draw3d_app.h"
#include "skybox_draw3d.h"
#include "vortex_cuda.h"

/* ---- the recording back-end ------------------------------------------------------------------------------------------ */
typedef struct fake_buffer {
  uint8_t* mem;
  uint64_t size;
  struct fake_buffer* next;
} fake_buffer_t;

typedef struct {
  uint32_t dcr[VX_DCR_OM_STATE_END + 1]; /* the reference's DCR file: base, tex, raster, om ranges */
  uint32_t ext[16];                      /* back-end DCRs 0x800.. */
  fake_buffer_t* buffers;
  pthread_t owner; /* every back-end call must come from the thread that opened the app: the pool only bins */
  /* log of launches */
  uint64_t* hashes;
  uint32_t num_launches, cap_launches;
  uint64_t uploads, upload_bytes;
} fake_device_t;

static fake_device_t g_dev;

static void check_thread(const char* what) {
  if (!pthread_equal(pthread_self(), g_dev.owner)) {
    fprintf(stderr, "FAILED: %s called from a worker thread\n", what);
    exit(3);
  }
}

static uint64_t fnv(uint64_t h, const void* p, uint64_t n) {
  const uint8_t* b = (const uint8_t*)p;
  uint64_t i;
  for (i = 0; i < n; ++i) h = (h ^ b[i]) * 0x100000001b3ull;
  return h;
}

/* block address (addr / 64 in 32 bits) -> pointer, against the live buffers (raster_pipeline.cu: resolve_block) */
static const uint8_t* resolve_block(uint32_t block, uint64_t* left) {
  fake_buffer_t* b;
  for (b = g_dev.buffers; b; b = b->next) {
    const uint64_t lo = (uint64_t)(uintptr_t)b->mem, hi = lo + b->size;
    uint64_t cand = (lo & ~((1ull << 38) - 1)) | ((uint64_t)block << 6);
    if (cand < lo) cand += 1ull << 38;
    if (cand >= lo && cand < hi) { *left = hi - cand; return (const uint8_t*)(uintptr_t)cand; }
  }
  return NULL;
}

int vx_dev_caps(vx_device_h d, uint32_t id, uint64_t* v) {
  (void)d;
  switch (id) {
    case VX_CAPS_ISA_FLAGS: *v = VX_ISA_EXT_RASTER | VX_ISA_EXT_TEX | VX_ISA_EXT_OM; return 0;
    case VX_CAPS_NUM_CORES: *v = 148; return 0;
    case VX_CAPS_NUM_WARPS: *v = 64; return 0;
    case VX_CAPS_NUM_THREADS: *v = 32; return 0;
  }
  return -1;
}

/* "Device memory": one reserved range handed out upwards and never reused within a flow, so that a 32-bit block address
 * (address / 64) names one buffer only, as CUDA device addresses do for the real back-end (2^38-byte window). */
#define ARENA_BYTES (32ull << 30)
static uint8_t* g_arena;
static uint64_t g_arena_used;

int vx_mem_alloc(vx_device_h d, uint64_t size, int flags, vx_buffer_h* out) {
  fake_buffer_t* b = (fake_buffer_t*)calloc(1, sizeof(*b));
  (void)d; (void)flags;
  check_thread("vx_mem_alloc");
  if (!g_arena) {
    void* p = mmap(NULL, ARENA_BYTES, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (p == MAP_FAILED) { free(b); return -1; }
    g_arena = (uint8_t*)p;
  }
  if (!b || g_arena_used + size + 256 > ARENA_BYTES) { free(b); return -1; }
  b->mem = g_arena + g_arena_used;
  g_arena_used = (g_arena_used + size + 255) & ~255ull;
  memset(b->mem, 0xCD, size);
  b->size = size;
  b->next = g_dev.buffers;
  g_dev.buffers = b;
  *out = b;
  return 0;
}

int vx_mem_free(vx_buffer_h h) {
  fake_buffer_t** p;
  if (!h) return 0;
  check_thread("vx_mem_free");
  for (p = &g_dev.buffers; *p; p = &(*p)->next)
    if (*p == (fake_buffer_t*)h) {
      fake_buffer_t* b = *p;
      *p = b->next;
      /* the bytes stay where they are: addresses are never handed out twice within a flow */
      free(b);
      return 0;
    }
  fprintf(stderr, "FAILED: vx_mem_free of an unknown buffer\n");
  exit(3);
}

int vx_mem_address(vx_buffer_h h, uint64_t* a) { *a = (uint64_t)(uintptr_t)((fake_buffer_t*)h)->mem; return 0; }

int vx_copy_to_dev(vx_buffer_h h, const void* src, uint64_t off, uint64_t size) {
  fake_buffer_t* b = (fake_buffer_t*)h;
  check_thread("vx_copy_to_dev");
  if (off + size > b->size) return -1;
  memcpy(b->mem + off, src, size);
  g_dev.uploads += 1;
  g_dev.upload_bytes += size;
  return 0;
}

int vx_copy_from_dev(void* dst, vx_buffer_h h, uint64_t off, uint64_t size) {
  fake_buffer_t* b = (fake_buffer_t*)h;
  if (off + size > b->size) return -1;
  memcpy(dst, b->mem + off, size);
  return 0;
}

int vx_upload_bytes(vx_device_h d, const void* src, uint64_t size, vx_buffer_h* out) {
  if (vx_mem_alloc(d, size, VX_MEM_READ, out) != 0) return -1;
  return vx_copy_to_dev(*out, src, 0, size);
}
int vx_upload_kernel_bytes(vx_device_h d, const void* src, uint64_t size, vx_buffer_h* out) { return vx_upload_bytes(d, src, size, out); }

int vx_dcr_write(vx_device_h d, uint32_t addr, uint32_t v) {
  (void)d;
  check_thread("vx_dcr_write");
  if (addr <= VX_DCR_OM_STATE_END) g_dev.dcr[addr] = v;
  else if (addr >= 0x800 && addr < 0x810) g_dev.ext[addr - 0x800] = v;
  else return -1;
  return 0;
}

int vx_ready_wait(vx_device_h d, uint64_t t) { (void)d; (void)t; return 0; }
int vx_cuda_fill32(vx_buffer_h h, uint64_t off, uint64_t size, uint32_t v) {
  fake_buffer_t* b = (fake_buffer_t*)h;
  uint64_t i;
  if (off + size > b->size) return -1;
  for (i = 0; i < size / 4; ++i) memcpy(b->mem + off + 4 * i, &v, 4);
  return 0;
}
int vx_cuda_host_register(vx_device_h d, void* p, uint64_t s) { (void)d; (void)p; (void)s; return -1; }
int vx_cuda_host_unregister(vx_device_h d, void* p) { (void)d; (void)p; return 0; }
int vx_cuda_draw3d_binning(vx_device_h a, vx_buffer_h b, uint32_t c, vx_buffer_h d, uint32_t e, float f, float g, uint32_t h,
                           uint32_t i, uint32_t j, uint32_t k, uint32_t l, uint32_t m, vx_buffer_h* n, vx_buffer_h* o,
                           srt_binning_info_t* p) {
  (void)a; (void)b; (void)c; (void)d; (void)e; (void)f; (void)g; (void)h; (void)i; (void)j; (void)k; (void)l; (void)m; (void)n; (void)o; (void)p;
  fprintf(stderr, "FAILED: device binning is not part of this check\n");
  exit(3);
}
int vx_cuda_copy_from_dev_async(void* dst, vx_buffer_h h, uint64_t off, uint64_t size, uint64_t* ticket) {
  *ticket = 0;
  return vx_copy_from_dev(dst, h, off, size);
}
int vx_cuda_copy_wait(vx_device_h d, uint64_t t) { (void)d; (void)t; return 0; }

/* vx_start: one hash of everything the tile kernel would consume, with device addresses left out */
int vx_start(vx_device_h d, vx_buffer_h kernel, vx_buffer_h args) {
  const fake_buffer_t* ab = (const fake_buffer_t*)args;
  draw3d_kernel_arg_t arg;
  uint32_t dcr[VX_DCR_OM_STATE_END + 1];
  uint64_t h = 0xcbf29ce484222325ull, tb_left = 0, pb_left = 0, tx_left = 0;
  const uint8_t *tb, *pb;
  uint32_t t, num_tiles, stride, wide;
  (void)d; (void)kernel;
  check_thread("vx_start");
  if (ab->size < sizeof(arg)) return -1;
  memcpy(&arg, ab->mem, sizeof(arg));
  memcpy(dcr, g_dev.dcr, sizeof(dcr));
  /* unit state: the device DCR file and the mirrors inside kernel_arg_t must agree, then addresses are blanked */
  for (t = VX_DCR_RASTER_STATE_BEGIN; t < VX_DCR_RASTER_STATE_END; ++t)
    if (dcr[t] != DRAW3D_RASTER_DCR(&arg, t)) { fprintf(stderr, "FAILED: raster DCR 0x%x differs from its mirror\n", t); exit(3); }
  for (t = VX_DCR_OM_STATE_BEGIN; t < VX_DCR_OM_STATE_END; ++t)
    if (dcr[t] != DRAW3D_OM_DCR(&arg, t)) { fprintf(stderr, "FAILED: OM DCR 0x%x differs from its mirror\n", t); exit(3); }
  num_tiles = dcr[VX_DCR_RASTER_TILE_COUNT];
  stride = dcr[VX_DCR_RASTER_PBUF_STRIDE];
  wide = g_dev.ext[4] == SRT_TILEBUF_WIDE;
  tb = resolve_block(dcr[VX_DCR_RASTER_TBUF_ADDR], &tb_left);
  pb = resolve_block(dcr[VX_DCR_RASTER_PBUF_ADDR], &pb_left);
  if (!tb || !pb || (uint64_t)(uintptr_t)pb != arg.prim_addr) { fprintf(stderr, "FAILED: tile / primitive buffer address\n"); exit(3); }
  /* tile headers, primitive lists and the records they name */
  for (t = 0; t < num_tiles; ++t) {
    const uint32_t* pids;
    uint32_t xy, count, k;
    if (wide) {
      const srt_tile_header_wide_t* hd = (const srt_tile_header_wide_t*)tb + t;
      uint32_t w[4];
      memcpy(w, hd, 16);
      xy = w[0]; count = w[1];
      pids = (const uint32_t*)((const srt_tile_header_wide_t*)tb + num_tiles) + w[2];
    } else {
      uint32_t w[2];
      memcpy(w, tb + 8ull * t, 8);
      xy = w[0]; count = w[1] >> 16;
      pids = (const uint32_t*)(tb + 8ull * (t + 1)) + (w[1] & 0xffff);
    }
    if ((uint64_t)((const uint8_t*)(pids + count) - tb) > tb_left) { fprintf(stderr, "FAILED: primitive list leaves the tile buffer\n"); exit(3); }
    h = fnv(h, &xy, 4);
    h = fnv(h, &count, 4);
    for (k = 0; k < count; ++k) {
      if ((uint64_t)(pids[k] + 1) * stride > pb_left) { fprintf(stderr, "FAILED: primitive id past the primitive buffer\n"); exit(3); }
      h = fnv(h, &pids[k], 4);
      h = fnv(h, pb + (uint64_t)pids[k] * stride, stride);
    }
  }
  if (arg.tex_enabled) {
    const uint32_t logdim = DRAW3D_TEX_DCR(&arg, 0, VX_DCR_TEX_LOGDIM), fmt = DRAW3D_TEX_DCR(&arg, 0, VX_DCR_TEX_FORMAT);
    const uint32_t bpp = fmt == VX_TEX_FORMAT_A8R8G8B8 ? 4 : (fmt <= VX_TEX_FORMAT_A8L8 ? 2 : 1);
    const uint64_t bytes = (uint64_t)bpp << ((logdim & 0xffff) + (logdim >> 16));
    const uint8_t* tx = resolve_block(DRAW3D_TEX_DCR(&arg, 0, VX_DCR_TEX_ADDR), &tx_left);
    if (!tx || bytes > tx_left) { fprintf(stderr, "FAILED: texture address\n"); exit(3); }
    for (t = VX_DCR_TEX_LOGDIM; t <= VX_DCR_TEX_MIPOFF(0); ++t) {
      if (g_dev.dcr[t] != DRAW3D_TEX_DCR(&arg, 0, t)) { fprintf(stderr, "FAILED: tex DCR 0x%x differs from its mirror\n", t); exit(3); }
      h = fnv(h, &g_dev.dcr[t], 4);
    }
    h = fnv(h, tx, bytes);
  }
  dcr[VX_DCR_RASTER_TBUF_ADDR] = dcr[VX_DCR_RASTER_PBUF_ADDR] = dcr[VX_DCR_OM_CBUF_ADDR] = dcr[VX_DCR_OM_ZBUF_ADDR] = 0;
  h = fnv(h, &dcr[VX_DCR_RASTER_STATE_BEGIN], 4ull * (VX_DCR_OM_STATE_END - VX_DCR_RASTER_STATE_BEGIN));
  {
    const uint8_t sw[7] = {arg.depth_enabled, arg.color_enabled, arg.tex_enabled, arg.tex_modulate, arg.sw_tex, arg.sw_rast, arg.sw_om};
    h = fnv(h, sw, sizeof(sw));
    h = fnv(h, &arg.log_num_tasks, 4);
  }
  if (g_dev.num_launches == g_dev.cap_launches) {
    g_dev.cap_launches = g_dev.cap_launches ? 2 * g_dev.cap_launches : 256;
    g_dev.hashes = (uint64_t*)realloc(g_dev.hashes, g_dev.cap_launches * sizeof(uint64_t));
    if (!g_dev.hashes) exit(3);
  }
  g_dev.hashes[g_dev.num_launches++] = h;
  return 0;
}

/* ---- the check --------------------------------------------------------------------------------------------------------- */
static int run_flow(const srt_trace_t* trace, uint32_t size, uint32_t flags, uint32_t frames, uint64_t** hashes, uint32_t* n) {
  srt_draw3d_t* app = NULL;
  uint32_t f;
  memset(&g_dev, 0, sizeof(g_dev));
  g_arena_used = 0;
  g_dev.owner = pthread_self();
  if (srt_draw3d_open((vx_device_h)&g_dev, size, size, flags, 0, 1, &app) != 0) return -1;
  for (f = 0; f < frames; ++f) {
    if (srt_draw3d_clear(app) != 0) return -1;
    if (f % 3 == 2 && trace->num_drawcalls > 2) { /* a frame drawn in two ranges, as the per-draw measurements do */
      if (srt_draw3d_render(app, trace, 0, trace->num_drawcalls / 2) != 0) return -1;
      if (srt_draw3d_render(app, trace, trace->num_drawcalls / 2 + 1, 0xffffffffu) != 0) return -1;
    } else if (srt_draw3d_render(app, trace, 0, 0xffffffffu) != 0) {
      return -1;
    }
  }
  printf("flags 0x%02x: %u launches over %u frames, %llu uploads, %llu bytes\n", flags, g_dev.num_launches, frames,
         (unsigned long long)g_dev.uploads, (unsigned long long)g_dev.upload_bytes);
  srt_draw3d_close(app);
  *hashes = g_dev.hashes;
  *n = g_dev.num_launches;
  if (g_dev.buffers) { fprintf(stderr, "FAILED: the app left device buffers behind\n"); return -1; }
  return 0;
}

int main(int argc, char** argv) {
  srt_trace_t trace;
  uint64_t *serial = NULL, *autoh = NULL;
  uint32_t ns = 0, na = 0, i, size, frames;
  const uint32_t sw = SRT_DRAW3D_SW_RAST | SRT_DRAW3D_SW_OM | SRT_DRAW3D_SW_TEX;
  if (argc < 4) { fprintf(stderr, "usage: %s trace.trc.gz size frames\n", argv[0]); return 2; }
  size = (uint32_t)atoi(argv[2]);
  frames = (uint32_t)atoi(argv[3]);
  if (srt_trace_load(argv[1], &trace) != 0) { fprintf(stderr, "cannot load %s\n", argv[1]); return 2; }
  if (run_flow(&trace, size, sw | SRT_DRAW3D_WIDE_TILES, frames, &serial, &ns) != 0) { fprintf(stderr, "FAILED: serial flow\n"); return 1; }
  if (run_flow(&trace, size, sw | SRT_DRAW3D_WIDE_TILES | SRT_DRAW3D_AUTO, frames, &autoh, &na) != 0) { fprintf(stderr, "FAILED: AUTO flow\n"); return 1; }
  if (ns != na || ns == 0) { fprintf(stderr, "FAILED: %u launches in the serial flow, %u under AUTO\n", ns, na); return 1; }
  for (i = 0; i < ns; ++i)
    if (serial[i] != autoh[i]) { fprintf(stderr, "FAILED: launch %u differs\n", i); return 1; }
  printf("PASSED! %u launches identical\n", ns);
  free(serial);
  free(autoh);
  srt_trace_free(&trace);
  return 0;
}
