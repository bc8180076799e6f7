/*
 * draw3d_app.c -- host side of the draw3d regression app (plain C over the vx_* ABI).
 * See draw3d_app.h; line references are to the reference's tests/regression/draw3d/main.cpp.
 */
#define _POSIX_C_SOURCE 200809L
#include "draw3d_app.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "skybox_rt.h"
#include "vortex_cuda.h"

#define CHECK(expr)                                                \
  do {                                                             \
    int _ret = (expr);                                             \
    if (0 != _ret) {                                               \
      printf("Error: '%s' returned %d!\n", #expr, (int)_ret);      \
      return _ret;                                                 \
    }                                                              \
  } while (0)

typedef struct {
  vx_buffer_h tile_buffer, prim_buffer, tex_buffer, args_buffer;
  draw3d_kernel_arg_t arg; /* as uploaded */
  uint32_t num_tiles;
} draw_slot_t;

/* SRT_DRAW3D_AUTO: the host pre-pass of a frame's draws runs on a few worker threads while the calling thread uploads and
 * starts the draws in submission order as their bins arrive (the draws of a frame are independent inputs; each bin is
 * the same srt_draw3d_binning_rank call the serial flow makes, so the buffers are identical). */
#define BIN_WORKERS 3
enum { BIN_SKIP = 0, BIN_QUEUED = 1, BIN_TAKEN = 2, BIN_DONE = 3 };
typedef struct {
  pthread_t th[BIN_WORKERS];
  int num_threads;
  pthread_mutex_t mu;
  pthread_cond_t cv_work, cv_done;
  /* the job: draws [0, n) of one trace; state[d] says which of them are binned on the host */
  const srt_drawcall_t* draws;
  uint32_t n, next, busy;
  uint32_t width, height, format, rank, num_ranks;
  srt_binning_t* out;
  int* rc;
  uint8_t* state;
  uint64_t epoch;
  int quit;
} bin_pool_t;

struct srt_draw3d {
  bin_pool_t* pool;
  vx_device_h device;
  uint32_t width, height, flags, rank, num_ranks;
  uint32_t zbuf_pitch, cbuf_pitch;
  uint64_t zbuf_size, cbuf_size, zbuf_addr, cbuf_addr;
  vx_buffer_h krnl_buffer, depth_buffer, color_buffer;
  /* render(): one set of per-draw buffers, replaced draw by draw like the reference's globals */
  vx_buffer_h tile_buffer, prim_buffer, tex_buffer, args_buffer;
  draw3d_kernel_arg_t kernel_arg; /* persists across draws (main.cpp:77) */
  /* prepare()/submit(): one slot per draw call, resident */
  draw_slot_t* slots;
  uint32_t num_slots;
  /* SRT_DRAW3D_AUTO frames: every host-binned buffer and texture of a frame packed into one device arena (one upload) */
  vx_buffer_h arena_buffer;
  uint64_t arena_capacity, arena_addr;
  uint8_t* pack_host;
  uint64_t pack_capacity;
  /* clear values kept in one staging buffer per target for the life of the app (the reference builds a
   * fresh std::vector per run, main.cpp:480-489); page-locked when the back-end offers it */
  uint32_t *clear_z, *clear_c;
  int clear_z_pinned, clear_c_pinned;
  srt_draw3d_stats_t stats;
  srt_om_override_t om_override;
  int om_override_on;
};

static bin_pool_t* bin_pool_create(void);
static void bin_pool_destroy(bin_pool_t* bp);

static double now_ms(void) {
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return 1e3 * (double)t.tv_sec + 1e-6 * (double)t.tv_nsec;
}

static uint32_t log2ceil_u32(uint32_t v) { /* sim/common/bitmanip.h:31-33 */
  uint32_t n = 0;
  while ((1ull << n) < v) ++n;
  return n;
}

/* graphics::toVXCompare / toVXStencilOp / toVXBlendFunc (sim/common/gfxutil.cpp:298-350) */
static int to_vx_compare(uint32_t c, uint32_t* out) {
  static const uint32_t m[8] = {VX_OM_DEPTH_FUNC_NEVER, VX_OM_DEPTH_FUNC_LESS, VX_OM_DEPTH_FUNC_EQUAL,
                                VX_OM_DEPTH_FUNC_LEQUAL, VX_OM_DEPTH_FUNC_GREATER, VX_OM_DEPTH_FUNC_NOTEQUAL,
                                VX_OM_DEPTH_FUNC_GEQUAL, VX_OM_DEPTH_FUNC_ALWAYS};
  if (c >= 8) { printf("Error: invalid compare function: %u\n", c); return -1; }
  *out = m[c];
  return 0;
}
static int to_vx_stencil_op(uint32_t o, uint32_t* out) {
  static const uint32_t m[6] = {VX_OM_STENCIL_OP_KEEP, VX_OM_STENCIL_OP_REPLACE, VX_OM_STENCIL_OP_INCR,
                                VX_OM_STENCIL_OP_DECR, VX_OM_STENCIL_OP_ZERO, VX_OM_STENCIL_OP_INVERT};
  if (o >= 6) { printf("Error: invalid stencil operation: %u\n", o); return -1; }
  *out = m[o];
  return 0;
}
static int to_vx_blend_func(uint32_t o, uint32_t* out) {
  static const uint32_t m[11] = {VX_OM_BLEND_FUNC_ZERO, VX_OM_BLEND_FUNC_ONE, VX_OM_BLEND_FUNC_SRC_RGB,
                                 VX_OM_BLEND_FUNC_ONE_MINUS_SRC_RGB, VX_OM_BLEND_FUNC_SRC_A,
                                 VX_OM_BLEND_FUNC_ONE_MINUS_SRC_A, VX_OM_BLEND_FUNC_DST_A,
                                 VX_OM_BLEND_FUNC_ONE_MINUS_DST_A, VX_OM_BLEND_FUNC_DST_RGB,
                                 VX_OM_BLEND_FUNC_ONE_MINUS_DST_RGB, VX_OM_BLEND_FUNC_ALPHA_SAT};
  if (o >= 11) { printf("Error: invalid blend function: %u\n", o); return -1; }
  *out = m[o];
  return 0;
}

/* RASTER_DCR_WRITE / OM_DCR_WRITE / TEX_DCR_WRITE (main.cpp:148-159): device DCR + kernel_arg mirror.
 * `live` = 0 records the mirror only (prepare() computes the arguments ahead of submission). */
typedef struct { srt_draw3d_t* r; draw3d_kernel_arg_t* arg; int live; int err; } dcr_ctx_t;
static void ras_w(dcr_ctx_t* c, uint32_t addr, uint32_t v) {
  if (c->live && vx_dcr_write(c->r->device, addr, v) != 0) c->err = -1;
  DRAW3D_RASTER_DCR(c->arg, addr) = v;
}
static void om_w(dcr_ctx_t* c, uint32_t addr, uint32_t v) {
  if (c->live && vx_dcr_write(c->r->device, addr, v) != 0) c->err = -1;
  DRAW3D_OM_DCR(c->arg, addr) = v;
}
static void tex_w(dcr_ctx_t* c, uint32_t addr, uint32_t v) {
  if (c->live && vx_dcr_write(c->r->device, addr, v) != 0) c->err = -1;
  if (addr == VX_DCR_TEX_STAGE) c->arg->tex_stage = v; /* TexDCRS::write, graphics.h:167-176 */
  else DRAW3D_TEX_DCR(c->arg, c->arg->tex_stage, addr) = v;
}

/* main.cpp:214-346: every DCR write of one draw call and the kernel_arg_t switches, in the
 * reference's order and with its quirks (ZPASS written twice :259-260; wrapV from addressU :304). */
static int program_draw(srt_draw3d_t* r, draw3d_kernel_arg_t* arg, int live, const srt_states_t* st,
                        const srt_texture_t* tex, uint64_t tilebuf_addr, uint64_t primbuf_addr,
                        uint64_t texbuf_addr, uint32_t num_tiles) {
  dcr_ctx_t c = {r, arg, live, 0};
  uint32_t v0, v1, v2, v3;
  ras_w(&c, VX_DCR_RASTER_TBUF_ADDR, (uint32_t)(tilebuf_addr / 64));
  ras_w(&c, VX_DCR_RASTER_TILE_COUNT, num_tiles);
  ras_w(&c, VX_DCR_RASTER_PBUF_ADDR, (uint32_t)(primbuf_addr / 64));
  ras_w(&c, VX_DCR_RASTER_PBUF_STRIDE, (uint32_t)sizeof(rast_prim_t));
  ras_w(&c, VX_DCR_RASTER_SCISSOR_X, (r->width << 16) | 0);
  ras_w(&c, VX_DCR_RASTER_SCISSOR_Y, (r->height << 16) | 0);

  om_w(&c, VX_DCR_OM_CBUF_ADDR, (uint32_t)(r->cbuf_addr / 64));
  om_w(&c, VX_DCR_OM_CBUF_PITCH, r->cbuf_pitch);
  om_w(&c, VX_DCR_OM_CBUF_WRITEMASK, st->color_writemask);
  if (st->depth_test || st->stencil_test) {
    om_w(&c, VX_DCR_OM_ZBUF_ADDR, (uint32_t)(r->zbuf_addr / 64));
    om_w(&c, VX_DCR_OM_ZBUF_PITCH, r->zbuf_pitch);
  }
  if (st->depth_test) {
    if (to_vx_compare(st->depth_func, &v0)) return -1;
    om_w(&c, VX_DCR_OM_DEPTH_FUNC, v0);
    om_w(&c, VX_DCR_OM_DEPTH_WRITEMASK, st->depth_writemask);
  } else {
    om_w(&c, VX_DCR_OM_DEPTH_FUNC, VX_OM_DEPTH_FUNC_ALWAYS);
    om_w(&c, VX_DCR_OM_DEPTH_WRITEMASK, 0);
  }
  if (st->stencil_test) {
    if (to_vx_compare(st->stencil_func, &v0) || to_vx_stencil_op(st->stencil_zpass, &v1) ||
        to_vx_stencil_op(st->stencil_zfail, &v2) || to_vx_stencil_op(st->stencil_fail, &v3))
      return -1;
    om_w(&c, VX_DCR_OM_STENCIL_FUNC, v0);
    om_w(&c, VX_DCR_OM_STENCIL_ZPASS, v1);
    om_w(&c, VX_DCR_OM_STENCIL_ZPASS, v2); /* sic */
    om_w(&c, VX_DCR_OM_STENCIL_FAIL, v3);
    om_w(&c, VX_DCR_OM_STENCIL_REF, st->stencil_ref);
    om_w(&c, VX_DCR_OM_STENCIL_MASK, st->stencil_mask);
    om_w(&c, VX_DCR_OM_STENCIL_WRITEMASK, st->stencil_writemask);
  } else {
    om_w(&c, VX_DCR_OM_STENCIL_FUNC, VX_OM_DEPTH_FUNC_ALWAYS);
    om_w(&c, VX_DCR_OM_STENCIL_ZPASS, VX_OM_STENCIL_OP_KEEP);
    om_w(&c, VX_DCR_OM_STENCIL_ZPASS, VX_OM_STENCIL_OP_KEEP);
    om_w(&c, VX_DCR_OM_STENCIL_FAIL, VX_OM_STENCIL_OP_KEEP);
    om_w(&c, VX_DCR_OM_STENCIL_REF, 0);
    om_w(&c, VX_DCR_OM_STENCIL_MASK, VX_OM_STENCIL_MASK);
    om_w(&c, VX_DCR_OM_STENCIL_WRITEMASK, 0);
  }
  om_w(&c, VX_DCR_OM_BLEND_MODE, (VX_OM_BLEND_MODE_ADD << 16) | (VX_OM_BLEND_MODE_ADD << 0));
  if (st->blend_enabled && r->om_override_on) { /* state beyond the trace format (srt_draw3d_set_om_override) */
    om_w(&c, VX_DCR_OM_BLEND_MODE, r->om_override.blend_mode);
    om_w(&c, VX_DCR_OM_BLEND_FUNC, r->om_override.blend_func);
    om_w(&c, VX_DCR_OM_BLEND_CONST, r->om_override.blend_const);
    om_w(&c, VX_DCR_OM_LOGIC_OP, r->om_override.logic_op);
  } else if (st->blend_enabled) {
    if (to_vx_blend_func(st->blend_src, &v0) || to_vx_blend_func(st->blend_dst, &v1)) return -1;
    om_w(&c, VX_DCR_OM_BLEND_FUNC, (v1 << 24) | (v1 << 16) | (v0 << 8) | (v0 << 0));
  } else {
    om_w(&c, VX_DCR_OM_BLEND_FUNC, (VX_OM_BLEND_FUNC_ZERO << 24) | (VX_OM_BLEND_FUNC_ZERO << 16) |
                                       (VX_OM_BLEND_FUNC_ONE << 8) | (VX_OM_BLEND_FUNC_ONE << 0));
  }
  if (st->texture_enabled) {
    uint32_t fmt = srt_format_to_vx(tex->format);
    int tex_filter = (st->texture_magfilter != SRT_FILTER_NEAREST) || (st->texture_magfilter != SRT_FILTER_NEAREST);
    uint32_t wrapU = (st->texture_addressU == SRT_ADDRESS_WRAP);
    uint32_t wrapV = (st->texture_addressU == SRT_ADDRESS_WRAP); /* sic */
    if (fmt == 0xffffffffu) { printf("Error: invalid format: %u\n", tex->format); return -1; }
    tex_w(&c, VX_DCR_TEX_STAGE, 0);
    tex_w(&c, VX_DCR_TEX_LOGDIM, (log2ceil_u32(tex->height) << 16) | log2ceil_u32(tex->width));
    tex_w(&c, VX_DCR_TEX_FORMAT, fmt);
    tex_w(&c, VX_DCR_TEX_WRAP, (wrapV << 16) | wrapU);
    tex_w(&c, VX_DCR_TEX_FILTER, tex_filter ? VX_TEX_FILTER_BILINEAR : VX_TEX_FILTER_POINT);
    tex_w(&c, VX_DCR_TEX_ADDR, (uint32_t)(texbuf_addr / 64));
    /* level 0 only: the draw3d kernel samples lod 0 and nothing else (kernel.cpp:160-164), so the
     * box-filtered chain of GenerateMipmaps (main.cpp:293) is not generated or uploaded */
    tex_w(&c, VX_DCR_TEX_MIPOFF(0), 0);
  }
  arg->depth_enabled = st->depth_test != 0;
  arg->color_enabled = st->color_enabled != 0;
  arg->tex_enabled = st->texture_enabled != 0;
  arg->tex_modulate = (st->texture_enabled && st->texture_envmode == SRT_ENVMODE_MODULATE);
  arg->prim_addr = primbuf_addr;
  if (arg->tex_modulate && !arg->color_enabled) arg->tex_modulate = 0;
  if (arg->tex_enabled && arg->color_enabled && !arg->tex_modulate) arg->color_enabled = 0;
  return c.err;
}

int srt_draw3d_open(vx_device_h device, uint32_t width, uint32_t height, uint32_t flags, uint32_t rank,
                    uint32_t num_ranks, srt_draw3d_t** out) {
  srt_draw3d_t* r;
  srt_kernel_image_t image;
  uint64_t isa_flags, num_cores, num_warps, num_threads;
  if (!device || !out || width == 0 || height == 0 || width > 32768 || height > 32768) return -1;
  if (num_ranks == 0) num_ranks = 1;
  if (rank >= num_ranks) return -1;
  /* main.cpp:400-418: the three graphics extensions must be there */
  CHECK(vx_dev_caps(device, VX_CAPS_ISA_FLAGS, &isa_flags));
  if (!(isa_flags & VX_ISA_EXT_RASTER)) { printf("RASTER ISA extensions are needed!\n"); return -1; }
  if (!(isa_flags & VX_ISA_EXT_TEX)) { printf("TEX ISA extensions are needed!\n"); return -1; }
  if (!(isa_flags & VX_ISA_EXT_OM)) { printf("OM ISA extensions are needed!\n"); return -1; }
  CHECK(vx_dev_caps(device, VX_CAPS_NUM_CORES, &num_cores));
  CHECK(vx_dev_caps(device, VX_CAPS_NUM_WARPS, &num_warps));
  CHECK(vx_dev_caps(device, VX_CAPS_NUM_THREADS, &num_threads));
  r = (srt_draw3d_t*)calloc(1, sizeof(*r));
  if (!r) return -1;
  r->device = device; r->width = width; r->height = height; r->flags = flags;
  r->rank = rank; r->num_ranks = num_ranks;
  /* main.cpp:461-467 (cbuf_size there is width*pitch, a square-target shortcut; height*pitch here) */
  r->zbuf_pitch = width * 4; r->zbuf_size = (uint64_t)height * r->zbuf_pitch;
  r->cbuf_pitch = width * 4; r->cbuf_size = (uint64_t)height * r->cbuf_pitch;
  /* "upload program" (main.cpp:457-459): the kernel image names the built-in pipeline */
  image.min_vma = 0x80000000ull; image.max_vma = 0x80000000ull + 64;
  image.magic = SRT_KERNEL_MAGIC; image.pipeline = SRT_PIPELINE_DRAW3D; image.version = 1;
  if (vx_upload_kernel_bytes(device, &image, sizeof(image), &r->krnl_buffer) != 0 ||
      vx_mem_alloc(device, r->zbuf_size, VX_MEM_READ_WRITE, &r->depth_buffer) != 0 ||
      vx_mem_address(r->depth_buffer, &r->zbuf_addr) != 0 ||
      vx_mem_alloc(device, r->cbuf_size, VX_MEM_READ_WRITE, &r->color_buffer) != 0 ||
      vx_mem_address(r->color_buffer, &r->cbuf_addr) != 0 ||
      vx_dcr_write(device, VX_DCR_RASTER_EXT_TBUF_FORMAT,
                   (flags & SRT_DRAW3D_WIDE_TILES) ? SRT_TILEBUF_WIDE : SRT_TILEBUF_REFERENCE) != 0 ||
      vx_dcr_write(device, VX_DCR_RASTER_EXT_TILE_LOGSIZE, RASTER_TILE_LOGSIZE) != 0) {
    srt_draw3d_close(r);
    return -1;
  }
  {
    uint64_t i;
    if (posix_memalign((void**)&r->clear_z, 4096, (r->zbuf_size + 4095) & ~4095ull) != 0) r->clear_z = NULL;
    if (posix_memalign((void**)&r->clear_c, 4096, (r->cbuf_size + 4095) & ~4095ull) != 0) r->clear_c = NULL;
    if (!r->clear_z || !r->clear_c) { srt_draw3d_close(r); return -1; }
    for (i = 0; i < r->zbuf_size / 4; ++i) r->clear_z[i] = 0xffffffffu; /* clear_depth, main.cpp:46 */
    for (i = 0; i < r->cbuf_size / 4; ++i) r->clear_c[i] = 0xff000000u; /* clear_color, main.cpp:45 */
    r->clear_z_pinned = vx_cuda_host_register(device, r->clear_z, r->zbuf_size) == 0;
    r->clear_c_pinned = vx_cuda_host_register(device, r->clear_c, r->cbuf_size) == 0;
  }
  /* main.cpp:493-496 */
  r->kernel_arg.log_num_tasks = log2ceil_u32((uint32_t)(num_cores * num_warps * num_threads));
  r->kernel_arg.sw_tex = (flags & SRT_DRAW3D_SW_TEX) != 0;
  r->kernel_arg.sw_rast = (flags & SRT_DRAW3D_SW_RAST) != 0;
  r->kernel_arg.sw_om = (flags & SRT_DRAW3D_SW_OM) != 0;
  if (flags & SRT_DRAW3D_AUTO) r->pool = bin_pool_create(); /* NULL (out of memory): the serial pre-pass */
  *out = r;
  return 0;
}

static void free_slots(srt_draw3d_t* r) {
  uint32_t i;
  for (i = 0; i < r->num_slots; ++i) {
    vx_mem_free(r->slots[i].tile_buffer);
    vx_mem_free(r->slots[i].prim_buffer);
    vx_mem_free(r->slots[i].tex_buffer);
    vx_mem_free(r->slots[i].args_buffer);
  }
  free(r->slots);
  r->slots = NULL;
  r->num_slots = 0;
}

void srt_draw3d_set_om_override(srt_draw3d_t* r, const srt_om_override_t* ov) {
  if (!r) return;
  r->om_override_on = ov != NULL;
  if (ov) r->om_override = *ov;
}

void srt_draw3d_close(srt_draw3d_t* r) { /* cleanup(), main.cpp:137-146 (the device stays the caller's) */
  if (!r) return;
  bin_pool_destroy(r->pool);
  free_slots(r);
  if (r->clear_z_pinned) vx_cuda_host_unregister(r->device, r->clear_z);
  if (r->clear_c_pinned) vx_cuda_host_unregister(r->device, r->clear_c);
  free(r->clear_z);
  free(r->clear_c);
  vx_mem_free(r->depth_buffer);
  vx_mem_free(r->color_buffer);
  vx_mem_free(r->arena_buffer);
  free(r->pack_host);
  vx_mem_free(r->tex_buffer);
  vx_mem_free(r->tile_buffer);
  vx_mem_free(r->prim_buffer);
  vx_mem_free(r->krnl_buffer);
  vx_mem_free(r->args_buffer);
  free(r);
}

int srt_draw3d_clear(srt_draw3d_t* r) { /* main.cpp:478-490 */
  int rc;
  if ((r->flags & SRT_DRAW3D_DEVICE_CLEAR) || ((r->flags & SRT_DRAW3D_AUTO) && (uint64_t)r->width * r->height >= (1u << 20))) {
    rc = vx_cuda_fill32(r->depth_buffer, 0, r->zbuf_size, 0xffffffffu);
    if (rc == 0) rc = vx_cuda_fill32(r->color_buffer, 0, r->cbuf_size, 0xff000000u);
    r->stats.h2d_bytes = 0;
    return rc;
  }
  rc = vx_copy_to_dev(r->depth_buffer, r->clear_z, 0, r->zbuf_size);
  if (rc == 0) rc = vx_copy_to_dev(r->color_buffer, r->clear_c, 0, r->cbuf_size);
  r->stats.h2d_bytes = r->zbuf_size + r->cbuf_size;
  return rc;
}

/* ---- binning pool ------------------------------------------------------------------------------------------------ */
/* Takes the next queued draw and bins it; called with the mutex held, returns with it held.  0 when nothing is queued. */
static int bin_take_one(bin_pool_t* bp) {
  while (bp->next < bp->n) {
    const uint32_t d = bp->next++;
    int rc;
    if (bp->state[d] != BIN_QUEUED) continue;
    bp->state[d] = BIN_TAKEN;
    pthread_mutex_unlock(&bp->mu);
    rc = srt_draw3d_binning_rank(&bp->draws[d], bp->width, bp->height, RASTER_TILE_LOGSIZE, bp->format, bp->rank,
                                 bp->num_ranks, &bp->out[d]);
    pthread_mutex_lock(&bp->mu);
    bp->rc[d] = rc;
    bp->state[d] = BIN_DONE;
    pthread_cond_broadcast(&bp->cv_done);
    return 1;
  }
  return 0;
}

static void* bin_worker(void* arg) {
  bin_pool_t* bp = (bin_pool_t*)arg;
  uint64_t seen = 0;
  pthread_mutex_lock(&bp->mu);
  for (;;) {
    while (!bp->quit && bp->epoch == seen) pthread_cond_wait(&bp->cv_work, &bp->mu);
    if (bp->quit) break;
    seen = bp->epoch;
    ++bp->busy;
    while (bin_take_one(bp)) {}
    --bp->busy;
    pthread_cond_broadcast(&bp->cv_done);
  }
  pthread_mutex_unlock(&bp->mu);
  return NULL;
}

static bin_pool_t* bin_pool_create(void) {
  bin_pool_t* bp = (bin_pool_t*)calloc(1, sizeof(*bp));
  int i;
  if (!bp) return NULL;
  pthread_mutex_init(&bp->mu, NULL);
  pthread_cond_init(&bp->cv_work, NULL);
  pthread_cond_init(&bp->cv_done, NULL);
  for (i = 0; i < BIN_WORKERS; ++i) {
    if (pthread_create(&bp->th[bp->num_threads], NULL, bin_worker, bp) != 0) break; /* fewer workers: the caller bins more itself */
    ++bp->num_threads;
  }
  return bp;
}

static void bin_pool_destroy(bin_pool_t* bp) {
  int i;
  if (!bp) return;
  pthread_mutex_lock(&bp->mu);
  bp->quit = 1;
  pthread_cond_broadcast(&bp->cv_work);
  pthread_mutex_unlock(&bp->mu);
  for (i = 0; i < bp->num_threads; ++i) pthread_join(bp->th[i], NULL);
  pthread_mutex_destroy(&bp->mu);
  pthread_cond_destroy(&bp->cv_work);
  pthread_cond_destroy(&bp->cv_done);
  free(bp);
}

/* Queues the host-binned draws of [start, end] and wakes the workers.  0, or -1 when out of memory (nothing queued). */
static int bin_job_begin(srt_draw3d_t* r, const srt_trace_t* trace, uint32_t start_draw, uint32_t end_draw) {
  bin_pool_t* bp = r->pool;
  const uint32_t n = trace->num_drawcalls;
  uint32_t d;
  srt_binning_t* out = (srt_binning_t*)calloc(n ? n : 1, sizeof(*out));
  int* rc = (int*)calloc(n ? n : 1, sizeof(*rc));
  uint8_t* state = (uint8_t*)calloc(n ? n : 1, 1);
  if (!out || !rc || !state) { free(out); free(rc); free(state); return -1; }
  for (d = 0; d < n; ++d) {
    const srt_drawcall_t* dc = &trace->drawcalls[d];
    const int on_device = (r->flags & SRT_DRAW3D_GPU_BINNING) || dc->num_primitives >= 10000u;
    if (d >= start_draw && d <= end_draw && !on_device) state[d] = BIN_QUEUED;
  }
  pthread_mutex_lock(&bp->mu);
  bp->draws = trace->drawcalls; bp->n = n; bp->next = 0;
  bp->width = r->width; bp->height = r->height; bp->format = (r->flags & SRT_DRAW3D_WIDE_TILES) ? SRT_TILEBUF_WIDE : SRT_TILEBUF_REFERENCE;
  bp->rank = r->rank; bp->num_ranks = r->num_ranks;
  bp->out = out; bp->rc = rc; bp->state = state;
  ++bp->epoch;
  pthread_cond_broadcast(&bp->cv_work);
  pthread_mutex_unlock(&bp->mu);
  return 0;
}

/* The bin of draw d (NULL when the draw is not binned on the host).  A draw no worker has started yet is binned here, on
 * the calling thread; one in progress is waited for.  The caller owns the result and frees it. */
static srt_binning_t* bin_job_get(srt_draw3d_t* r, uint32_t d, int* rc) {
  bin_pool_t* bp = r->pool;
  srt_binning_t* b = NULL;
  pthread_mutex_lock(&bp->mu);
  if (bp->state[d] == BIN_QUEUED) {
    bp->state[d] = BIN_TAKEN; /* the workers skip it from now on */
    pthread_mutex_unlock(&bp->mu);
    *rc = srt_draw3d_binning_rank(&bp->draws[d], bp->width, bp->height, RASTER_TILE_LOGSIZE, bp->format, bp->rank,
                                  bp->num_ranks, &bp->out[d]);
    pthread_mutex_lock(&bp->mu);
    bp->rc[d] = *rc;
    bp->state[d] = BIN_DONE;
  }
  if (bp->state[d] != BIN_SKIP) {
    while (bp->state[d] != BIN_DONE) pthread_cond_wait(&bp->cv_done, &bp->mu);
    b = &bp->out[d];
    *rc = bp->rc[d];
    bp->state[d] = BIN_SKIP; /* consumed */
  }
  pthread_mutex_unlock(&bp->mu);
  return b;
}

/* Ends the job: drains what is still queued, waits for the workers to leave it, frees unconsumed bins. */
static void bin_job_end(srt_draw3d_t* r) {
  bin_pool_t* bp = r->pool;
  uint32_t d;
  pthread_mutex_lock(&bp->mu);
  bp->next = bp->n; /* nothing more is started */
  while (bp->busy != 0) pthread_cond_wait(&bp->cv_done, &bp->mu);
  for (d = 0; d < bp->n; ++d) {
    while (bp->state[d] == BIN_TAKEN) pthread_cond_wait(&bp->cv_done, &bp->mu);
    if (bp->state[d] == BIN_DONE && bp->rc[d] == 0) srt_binning_free(&bp->out[d]);
  }
  free(bp->out); free(bp->rc); free(bp->state);
  bp->out = NULL; bp->rc = NULL; bp->state = NULL; bp->draws = NULL;
  bp->n = bp->next = 0;
  pthread_mutex_unlock(&bp->mu);
}

static uint32_t tile_format(const srt_draw3d_t* r) {
  return (r->flags & SRT_DRAW3D_WIDE_TILES) ? SRT_TILEBUF_WIDE : SRT_TILEBUF_REFERENCE;
}

/* bin + allocate + upload the buffers of one draw (main.cpp:186-212, :286-312) */
static int stage_draw(srt_draw3d_t* r, const srt_trace_t* trace, const srt_drawcall_t* dc, vx_buffer_h* tile_buffer,
                      vx_buffer_h* prim_buffer, vx_buffer_h* tex_buffer, const srt_texture_t** tex_out,
                      uint64_t addr[3], uint32_t* num_tiles, const srt_binning_t* prebinned, int prebinned_rc) {
  srt_binning_t b;
  double t0 = now_ms(), t1;
  int rc;
  *tex_out = NULL;
  if ((r->flags & SRT_DRAW3D_GPU_BINNING) || ((r->flags & SRT_DRAW3D_AUTO) && dc->num_primitives >= 10000u)) {
    /* the pre-pass on the device: upload the draw's vertices and indices instead of its binned form */
    vx_buffer_h vb = NULL, ib = NULL;
    srt_binning_info_t info;
    const uint64_t vbytes = (uint64_t)dc->num_vertices * sizeof(srt_vertex_t), ibytes = (uint64_t)dc->num_primitives * sizeof(srt_primitive_t);
    *num_tiles = 0;
    if (dc->num_primitives == 0 || dc->num_vertices == 0) return 0;
    if (*tile_buffer) vx_mem_free(*tile_buffer);
    if (*prim_buffer) vx_mem_free(*prim_buffer);
    *tile_buffer = *prim_buffer = NULL;
    rc = vx_upload_bytes(r->device, dc->vertices, vbytes, &vb) || vx_upload_bytes(r->device, dc->primitives, ibytes, &ib);
    t1 = now_ms();
    r->stats.upload_ms += t1 - t0;
    if (rc == 0)
      rc = vx_cuda_draw3d_binning(r->device, vb, dc->num_vertices, ib, dc->num_primitives, dc->viewport.near_, dc->viewport.far_,
                                  r->width, r->height, RASTER_TILE_LOGSIZE, tile_format(r), r->rank, r->num_ranks, tile_buffer,
                                  prim_buffer, &info);
    vx_mem_free(vb);
    vx_mem_free(ib);
    if (rc != 0) {
      printf("Error: binning failed (%d)%s\n", rc, rc == -2 ? ": 16-bit tile headers overflow, use wide tiles" : "");
      return rc;
    }
    r->stats.bin_ms += now_ms() - t1;
    t1 = now_ms();
    *num_tiles = info.num_tiles;
    if (info.num_tiles == 0) return 0;
    rc = vx_mem_address(*tile_buffer, &addr[0]) || vx_mem_address(*prim_buffer, &addr[1]);
    r->stats.h2d_bytes += vbytes + ibytes;
    r->stats.prims += info.num_prims; r->stats.tiles += info.num_tiles; r->stats.tile_prims += info.num_tile_prims;
    r->stats.draws_submitted += 1;
  } else {
    if (prebinned) { b = *prebinned; rc = prebinned_rc; } /* binned by the pool; this function frees it like its own */
    else rc = srt_draw3d_binning_rank(dc, r->width, r->height, RASTER_TILE_LOGSIZE, tile_format(r), r->rank, r->num_ranks, &b);
    if (rc != 0) {
      printf("Error: binning failed (%d)%s\n", rc, rc == -2 ? ": 16-bit tile headers overflow, use wide tiles" : "");
      return rc;
    }
    t1 = now_ms();
    r->stats.bin_ms += t1 - t0;
    *num_tiles = b.num_tiles;
    if (b.num_tiles == 0) { srt_binning_free(&b); return 0; } /* main.cpp:190-191 */
    if (*tile_buffer) vx_mem_free(*tile_buffer);
    if (*prim_buffer) vx_mem_free(*prim_buffer);
    *tile_buffer = *prim_buffer = NULL;
    rc = vx_mem_alloc(r->device, b.tilebuf_size, VX_MEM_READ, tile_buffer) || vx_mem_address(*tile_buffer, &addr[0]) ||
         vx_mem_alloc(r->device, b.primbuf_size, VX_MEM_READ, prim_buffer) || vx_mem_address(*prim_buffer, &addr[1]) ||
         vx_copy_to_dev(*tile_buffer, b.tilebuf, 0, b.tilebuf_size) || vx_copy_to_dev(*prim_buffer, b.primbuf, 0, b.primbuf_size);
    r->stats.h2d_bytes += b.tilebuf_size + b.primbuf_size;
    r->stats.prims += b.num_prims; r->stats.tiles += b.num_tiles; r->stats.tile_prims += b.num_tile_prims;
    r->stats.draws_submitted += 1;
    srt_binning_free(&b);
  }
  if (rc == 0 && dc->states.texture_enabled) {
    const srt_texture_t* tex = srt_trace_texture(trace, dc->texture_id);
    uint64_t level0;
    if (!tex) { printf("Error: texture %u is not in the trace\n", dc->texture_id); return -1; }
    level0 = (uint64_t)tex->width * tex->height * srt_format_bpp(tex->format);
    if (*tex_buffer) vx_mem_free(*tex_buffer);
    *tex_buffer = NULL;
    rc = vx_mem_alloc(r->device, level0, VX_MEM_READ, tex_buffer) || vx_mem_address(*tex_buffer, &addr[2]) ||
         vx_copy_to_dev(*tex_buffer, tex->pixels, 0, level0);
    r->stats.h2d_bytes += level0;
    *tex_out = tex;
  }
  r->stats.upload_ms += now_ms() - t1;
  return rc ? -1 : 0;
}

static int render_draws(srt_draw3d_t* r, const srt_trace_t* trace, uint32_t start_draw, uint32_t end_draw, int pooled) {
  uint32_t d;
  for (d = 0; d < trace->num_drawcalls; ++d) {
    const srt_drawcall_t* dc = &trace->drawcalls[d];
    const srt_texture_t* tex;
    const srt_binning_t* pre = NULL;
    uint64_t addr[3] = {0, 0, 0};
    uint32_t num_tiles;
    int pre_rc = 0;
    double t0;
    if (d < start_draw || d > end_draw) continue;
    if (pooled) {
      t0 = now_ms();
      pre = bin_job_get(r, d, &pre_rc);
      r->stats.bin_ms += now_ms() - t0; /* time this thread waited for (or spent on) bins */
    }
    CHECK(stage_draw(r, trace, dc, &r->tile_buffer, &r->prim_buffer, &r->tex_buffer, &tex, addr, &num_tiles, pre, pre_rc));
    if (num_tiles == 0) continue;
    t0 = now_ms();
    CHECK(program_draw(r, &r->kernel_arg, 1, &dc->states, tex, addr[0], addr[1], addr[2], num_tiles));
    /* the reference leaks the previous args buffer per draw (main.cpp:346); freed here */
    if (r->args_buffer) vx_mem_free(r->args_buffer);
    r->args_buffer = NULL;
    CHECK(vx_upload_bytes(r->device, &r->kernel_arg, sizeof(r->kernel_arg), &r->args_buffer));
    r->stats.h2d_bytes += sizeof(r->kernel_arg);
    CHECK(vx_start(r->device, r->krnl_buffer, r->args_buffer));
    if (!(r->flags & SRT_DRAW3D_AUTO)) CHECK(vx_ready_wait(r->device, VX_MAX_TIMEOUT)); /* main.cpp:357 */
    r->stats.run_ms += now_ms() - t0;
  }
  return 0;
}

/* SRT_DRAW3D_AUTO, every draw of the range binned on the host: the frame's tile buffers, primitive buffers and textures
 * (each texture once) are packed into one host block and travel as ONE upload into one device arena, the kernel
 * arguments of all draws follow, and only then are the draws started back to back.  The per-draw flow interleaves
 * ~3 small uploads with every kernel; under a concurrent 64 MB read-back each of those waits its turn on the host link
 * (measured: a frame rendered beside a read-back took 2.0 ms instead of 1.1), which is what kept the pipelined loop
 * from overlapping the two.  Same bins, same DCR values, same kernels: the frame is identical. */
typedef struct { uint64_t tile_off, prim_off, tex_off; uint32_t num_tiles; const srt_texture_t* tex; srt_binning_t bin; int have_bin; } pack_draw_t;

static uint64_t pack_align(uint64_t v) { return (v + 255) & ~255ull; }

static int render_frame_packed(srt_draw3d_t* r, const srt_trace_t* trace, uint32_t start_draw, uint32_t end_draw) {
  const uint32_t n = trace->num_drawcalls;
  pack_draw_t* pd = (pack_draw_t*)calloc(n ? n : 1, sizeof(*pd));
  uint64_t* tex_off = (uint64_t*)calloc(trace->num_textures ? trace->num_textures : 1, sizeof(uint64_t));
  draw3d_kernel_arg_t arg = r->kernel_arg;
  uint64_t total = 0, off;
  uint32_t d, t;
  int rc = 0;
  double t0 = now_ms(), t1;
  if (!pd || !tex_off) { free(pd); free(tex_off); return -1; }
  free_slots(r);
  r->slots = (draw_slot_t*)calloc(n ? n : 1, sizeof(draw_slot_t));
  if (!r->slots) { free(pd); free(tex_off); return -1; }
  /* 1. collect the bins in submission order and lay the arena out */
  for (t = 0; t < trace->num_textures; ++t) tex_off[t] = ~0ull;
  for (d = 0; d < n && rc == 0; ++d) {
    const srt_drawcall_t* dc = &trace->drawcalls[d];
    srt_binning_t* b;
    int brc = 0;
    if (d < start_draw || d > end_draw) continue;
    b = bin_job_get(r, d, &brc);
    if (!b) { rc = -1; break; } /* not host-binned: the caller checked that this cannot happen */
    if (brc != 0) {
      printf("Error: binning failed (%d)%s\n", brc, brc == -2 ? ": 16-bit tile headers overflow, use wide tiles" : "");
      rc = brc;
      break;
    }
    pd[d].bin = *b;
    pd[d].have_bin = 1;
    pd[d].num_tiles = b->num_tiles;
    if (b->num_tiles == 0) continue; /* main.cpp:190-191 */
    pd[d].tile_off = total; total = pack_align(total + b->tilebuf_size);
    pd[d].prim_off = total; total = pack_align(total + b->primbuf_size);
    r->stats.prims += b->num_prims; r->stats.tiles += b->num_tiles; r->stats.tile_prims += b->num_tile_prims;
    r->stats.draws_submitted += 1;
    if (dc->states.texture_enabled) {
      const srt_texture_t* tex = srt_trace_texture(trace, dc->texture_id);
      if (!tex) { printf("Error: texture %u is not in the trace\n", dc->texture_id); rc = -1; break; }
      t = (uint32_t)(tex - trace->textures);
      if (tex_off[t] == ~0ull) {
        tex_off[t] = total;
        total = pack_align(total + (uint64_t)tex->width * tex->height * srt_format_bpp(tex->format));
      }
      pd[d].tex = tex;
      pd[d].tex_off = tex_off[t];
    }
  }
  t1 = now_ms();
  r->stats.bin_ms += t1 - t0;
  /* 2. pack on the host, one arena on the device, one upload (in pieces the back-end stages without blocking) */
  if (rc == 0 && total > 0) {
    if (total > r->pack_capacity) {
      free(r->pack_host);
      r->pack_host = (uint8_t*)malloc(total);
      r->pack_capacity = r->pack_host ? total : 0;
      if (!r->pack_host) rc = -1;
    }
    if (rc == 0 && total > r->arena_capacity) {
      vx_mem_free(r->arena_buffer); /* ordered behind the draws that still read it */
      r->arena_buffer = NULL;
      r->arena_capacity = 0;
      rc = vx_mem_alloc(r->device, total + total / 4, VX_MEM_READ, &r->arena_buffer) || vx_mem_address(r->arena_buffer, &r->arena_addr);
      if (rc == 0) r->arena_capacity = total + total / 4;
    }
    if (rc == 0) {
      for (d = 0; d < n; ++d) {
        if (!pd[d].have_bin || pd[d].num_tiles == 0) continue;
        memcpy(r->pack_host + pd[d].tile_off, pd[d].bin.tilebuf, pd[d].bin.tilebuf_size);
        memcpy(r->pack_host + pd[d].prim_off, pd[d].bin.primbuf, pd[d].bin.primbuf_size);
      }
      for (t = 0; t < trace->num_textures; ++t)
        if (tex_off[t] != ~0ull) {
          const srt_texture_t* tex = &trace->textures[t];
          memcpy(r->pack_host + tex_off[t], tex->pixels, (uint64_t)tex->width * tex->height * srt_format_bpp(tex->format));
        }
      for (off = 0; off < total && rc == 0; off += 4ull << 20) {
        const uint64_t piece = total - off < (4ull << 20) ? total - off : (4ull << 20);
        rc = vx_copy_to_dev(r->arena_buffer, r->pack_host + off, off, piece);
      }
      r->stats.h2d_bytes += total;
    }
  }
  /* 3. the kernel arguments of every draw (mirrors only; submit() replays the DCRs), then all draws back to back */
  for (d = 0; d < n && rc == 0; ++d) {
    draw_slot_t* s;
    if (!pd[d].have_bin || pd[d].num_tiles == 0) continue;
    s = &r->slots[r->num_slots++];
    s->num_tiles = pd[d].num_tiles;
    rc = program_draw(r, &arg, 0, &trace->drawcalls[d].states, pd[d].tex, r->arena_addr + pd[d].tile_off,
                      r->arena_addr + pd[d].prim_off, r->arena_addr + pd[d].tex_off, s->num_tiles);
    if (rc != 0) break;
    s->arg = arg;
    rc = vx_upload_bytes(r->device, &s->arg, sizeof(s->arg), &s->args_buffer);
    r->stats.h2d_bytes += sizeof(s->arg);
  }
  for (d = 0; d < n; ++d)
    if (pd[d].have_bin) srt_binning_free(&pd[d].bin);
  free(pd);
  free(tex_off);
  r->stats.upload_ms += now_ms() - t1;
  if (rc != 0) return rc < 0 ? rc : -1;
  r->kernel_arg = arg; /* the mirrors persist across draws and frames (main.cpp:77) */
  t1 = now_ms();
  rc = srt_draw3d_submit(r);
  r->stats.run_ms += now_ms() - t1;
  return rc;
}

int srt_draw3d_render(srt_draw3d_t* r, const srt_trace_t* trace, uint32_t start_draw, uint32_t end_draw) {
  int rc, pooled;
  memset(&r->stats, 0, sizeof(r->stats));
  /* more than one draw to bin and a pool (SRT_DRAW3D_AUTO): the pre-pass of the frame runs beside the submission */
  pooled = r->pool && trace->num_drawcalls > 1 && end_draw > start_draw && bin_job_begin(r, trace, start_draw, end_draw) == 0;
  if (pooled && !(r->flags & SRT_DRAW3D_GPU_BINNING)) {
    /* every draw of the range small enough for the host pre-pass: the packed frame */
    uint32_t d, packed = 1;
    for (d = start_draw; d <= end_draw && d < trace->num_drawcalls; ++d)
      if (trace->drawcalls[d].num_primitives >= 10000u) packed = 0;
    if (packed) {
      rc = render_frame_packed(r, trace, start_draw, end_draw);
      bin_job_end(r);
      return rc;
    }
  }
  rc = render_draws(r, trace, start_draw, end_draw, pooled);
  if (pooled) bin_job_end(r);
  return rc;
}

int srt_draw3d_prepare(srt_draw3d_t* r, const srt_trace_t* trace) {
  uint32_t d;
  draw3d_kernel_arg_t arg = r->kernel_arg;
  memset(&r->stats, 0, sizeof(r->stats));
  free_slots(r);
  r->slots = (draw_slot_t*)calloc(trace->num_drawcalls ? trace->num_drawcalls : 1, sizeof(draw_slot_t));
  if (!r->slots) return -1;
  for (d = 0; d < trace->num_drawcalls; ++d) {
    const srt_drawcall_t* dc = &trace->drawcalls[d];
    const srt_texture_t* tex;
    draw_slot_t* s = &r->slots[r->num_slots];
    uint64_t addr[3] = {0, 0, 0};
    CHECK(stage_draw(r, trace, dc, &s->tile_buffer, &s->prim_buffer, &s->tex_buffer, &tex, addr, &s->num_tiles, NULL, 0));
    if (s->num_tiles == 0) continue;
    ++r->num_slots;
    /* mirrors only: the device DCRs are written at submission time */
    CHECK(program_draw(r, &arg, 0, &dc->states, tex, addr[0], addr[1], addr[2], s->num_tiles));
    s->arg = arg;
    CHECK(vx_upload_bytes(r->device, &s->arg, sizeof(s->arg), &s->args_buffer));
    r->stats.h2d_bytes += sizeof(s->arg);
  }
  return 0;
}

int srt_draw3d_submit(srt_draw3d_t* r) {
  uint32_t i, a;
  for (i = 0; i < r->num_slots; ++i) {
    const draw_slot_t* s = &r->slots[i];
    /* replay the draw's DCR state from its mirror so the "hardware unit" path sees it too */
    for (a = VX_DCR_RASTER_STATE_BEGIN; a < VX_DCR_RASTER_STATE_END; ++a)
      CHECK(vx_dcr_write(r->device, a, DRAW3D_RASTER_DCR(&s->arg, a)));
    for (a = VX_DCR_OM_STATE_BEGIN; a < VX_DCR_OM_STATE_END; ++a)
      CHECK(vx_dcr_write(r->device, a, DRAW3D_OM_DCR(&s->arg, a)));
    if (s->arg.tex_enabled) {
      CHECK(vx_dcr_write(r->device, VX_DCR_TEX_STAGE, 0));
      for (a = VX_DCR_TEX_ADDR; a <= VX_DCR_TEX_MIPOFF(0); ++a)
        CHECK(vx_dcr_write(r->device, a, DRAW3D_TEX_DCR(&s->arg, 0, a)));
    }
    CHECK(vx_start(r->device, r->krnl_buffer, s->args_buffer));
  }
  return 0;
}

int srt_draw3d_read(srt_draw3d_t* r, uint32_t* cbuf, uint32_t* zbuf) { /* main.cpp:380-382 */
  r->stats.d2h_bytes = 0;
  if (cbuf) { CHECK(vx_copy_from_dev(cbuf, r->color_buffer, 0, r->cbuf_size)); r->stats.d2h_bytes += r->cbuf_size; }
  if (zbuf) { CHECK(vx_copy_from_dev(zbuf, r->depth_buffer, 0, r->zbuf_size)); r->stats.d2h_bytes += r->zbuf_size; }
  return 0;
}

int srt_draw3d_read_async(srt_draw3d_t* r, uint32_t* cbuf, uint64_t* ticket) {
  CHECK(vx_cuda_copy_from_dev_async(cbuf, r->color_buffer, 0, r->cbuf_size, ticket));
  r->stats.d2h_bytes = r->cbuf_size;
  return 0;
}

int srt_draw3d_read_wait(srt_draw3d_t* r, uint64_t ticket) { return vx_cuda_copy_wait(r->device, ticket); }

const srt_draw3d_stats_t* srt_draw3d_stats(const srt_draw3d_t* r) { return &r->stats; }
vx_buffer_h srt_draw3d_color_buffer(const srt_draw3d_t* r) { return r->color_buffer; }
vx_buffer_h srt_draw3d_depth_buffer(const srt_draw3d_t* r) { return r->depth_buffer; }
