/*
 * skybox_draw3d.h -- the raster hot path (group R): the device-visible records,
 * DCR numbers and the kernel argument of the reference's `draw3d` app, plus this repo's C host API
 * for the pre-pass (trace loading, triangle setup + tile binning).
 *
 * Everything in the first half is the REFERENCE's layout, restated for C (the reference headers are
 * C++): a host program compiled against this header produces byte-identical buffers and a
 * byte-identical kernel argument to one compiled against the reference's
 *   sim/common/graphics.h:38-71            (rast_prim_t, rast_tile_header_t, rast_attribs_t)
 *   tests/regression/draw3d/common.h:18-34 (kernel_arg_t, 288 bytes)
 *   hw/rtl/VX_types.vh:309-458             (TEX / RASTER / OM constants and DCR numbers)
 * tests/test_draw3d_layout.py checks the sizes against the compiled reference (oracle/_ref).
 */
#ifndef SKYBOX_DRAW3D_H
#define SKYBOX_DRAW3D_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- constants: hw/rtl/VX_types.vh ------------------------------------------------------------ */
#define VX_TEX_STAGE_COUNT 2  /* :309 */
#define VX_TEX_LOD_MAX 15     /* :314-315 */
#define VX_TEX_FXD_FRAC 23    /* :319  u,v handed to the sampler are 9.23 fixed point */

#define VX_TEX_FILTER_POINT 0
#define VX_TEX_FILTER_BILINEAR 1
#define VX_TEX_WRAP_CLAMP 0
#define VX_TEX_WRAP_REPEAT 1
#define VX_TEX_WRAP_MIRROR 2
#define VX_TEX_FORMAT_A8R8G8B8 0
#define VX_TEX_FORMAT_R5G6B5 1
#define VX_TEX_FORMAT_A1R5G5B5 2
#define VX_TEX_FORMAT_A4R4G4B4 3
#define VX_TEX_FORMAT_A8L8 4
#define VX_TEX_FORMAT_L8 5
#define VX_TEX_FORMAT_A8 6

#define VX_DCR_TEX_STATE_BEGIN 0x006 /* == VX_DCR_BASE_STATE_END, :337 */
#define VX_DCR_TEX_STAGE (VX_DCR_TEX_STATE_BEGIN + 0)
#define VX_DCR_TEX_ADDR (VX_DCR_TEX_STATE_BEGIN + 1)
#define VX_DCR_TEX_LOGDIM (VX_DCR_TEX_STATE_BEGIN + 2)
#define VX_DCR_TEX_FORMAT (VX_DCR_TEX_STATE_BEGIN + 3)
#define VX_DCR_TEX_FILTER (VX_DCR_TEX_STATE_BEGIN + 4)
#define VX_DCR_TEX_WRAP (VX_DCR_TEX_STATE_BEGIN + 5)
#define VX_DCR_TEX_MIPOFF(lod) (VX_DCR_TEX_STATE_BEGIN + 6 + (lod))
/* Mip levels > 0: the DCR file stores all 16 MIPOFF words and the kernel_arg_t mirror carries them, but this
 * back-end SAMPLES LEVEL 0 ONLY -- the one call site, draw3d/kernel.cpp:160-164, passes lod = 0 to the sampler, and
 * the apps upload level 0 alone (apps/draw3d/draw3d_app.c).  A kernel that wants trilinear lookups
 * (sim/common/graphics.cpp:253-314 with lod > 0) is outside what this library renders. */
#define VX_DCR_TEX_STATE_END (VX_DCR_TEX_MIPOFF(VX_TEX_LOD_MAX) + 1)
#define VX_DCR_TEX_STATE_COUNT (VX_DCR_TEX_STATE_END - VX_DCR_TEX_STATE_BEGIN) /* 22 */

#define RASTER_TILE_LOGSIZE 5 /* hw/rtl/VX_config.vh:477-478 : 32x32-pixel tiles */
#define VX_RASTER_DIM_BITS 15 /* :352 */

#define VX_DCR_RASTER_STATE_BEGIN VX_DCR_TEX_STATE_END
#define VX_DCR_RASTER_TBUF_ADDR (VX_DCR_RASTER_STATE_BEGIN + 0)
#define VX_DCR_RASTER_TILE_COUNT (VX_DCR_RASTER_STATE_BEGIN + 1)
#define VX_DCR_RASTER_PBUF_ADDR (VX_DCR_RASTER_STATE_BEGIN + 2)
#define VX_DCR_RASTER_PBUF_STRIDE (VX_DCR_RASTER_STATE_BEGIN + 3)
#define VX_DCR_RASTER_SCISSOR_X (VX_DCR_RASTER_STATE_BEGIN + 4)
#define VX_DCR_RASTER_SCISSOR_Y (VX_DCR_RASTER_STATE_BEGIN + 5)
#define VX_DCR_RASTER_STATE_END (VX_DCR_RASTER_STATE_BEGIN + 6)
#define VX_DCR_RASTER_STATE_COUNT 6

#define VX_OM_DEPTH_BITS 24
#define VX_OM_DEPTH_MASK ((1u << VX_OM_DEPTH_BITS) - 1)
#define VX_OM_STENCIL_MASK 0xffu

#define VX_OM_DEPTH_FUNC_ALWAYS 0
#define VX_OM_DEPTH_FUNC_NEVER 1
#define VX_OM_DEPTH_FUNC_LESS 2
#define VX_OM_DEPTH_FUNC_LEQUAL 3
#define VX_OM_DEPTH_FUNC_EQUAL 4
#define VX_OM_DEPTH_FUNC_GEQUAL 5
#define VX_OM_DEPTH_FUNC_GREATER 6
#define VX_OM_DEPTH_FUNC_NOTEQUAL 7

#define VX_OM_STENCIL_OP_KEEP 0
#define VX_OM_STENCIL_OP_ZERO 1
#define VX_OM_STENCIL_OP_REPLACE 2
#define VX_OM_STENCIL_OP_INCR 3
#define VX_OM_STENCIL_OP_DECR 4
#define VX_OM_STENCIL_OP_INVERT 5
#define VX_OM_STENCIL_OP_INCR_WRAP 6
#define VX_OM_STENCIL_OP_DECR_WRAP 7

#define VX_OM_BLEND_MODE_ADD 0
#define VX_OM_BLEND_MODE_SUB 1
#define VX_OM_BLEND_MODE_REV_SUB 2
#define VX_OM_BLEND_MODE_MIN 3
#define VX_OM_BLEND_MODE_MAX 4
#define VX_OM_BLEND_MODE_LOGICOP 5

#define VX_OM_BLEND_FUNC_ZERO 0
#define VX_OM_BLEND_FUNC_ONE 1
#define VX_OM_BLEND_FUNC_SRC_RGB 2
#define VX_OM_BLEND_FUNC_ONE_MINUS_SRC_RGB 3
#define VX_OM_BLEND_FUNC_DST_RGB 4
#define VX_OM_BLEND_FUNC_ONE_MINUS_DST_RGB 5
#define VX_OM_BLEND_FUNC_SRC_A 6
#define VX_OM_BLEND_FUNC_ONE_MINUS_SRC_A 7
#define VX_OM_BLEND_FUNC_DST_A 8
#define VX_OM_BLEND_FUNC_ONE_MINUS_DST_A 9
#define VX_OM_BLEND_FUNC_CONST_RGB 10
#define VX_OM_BLEND_FUNC_ONE_MINUS_CONST_RGB 11
#define VX_OM_BLEND_FUNC_CONST_A 12
#define VX_OM_BLEND_FUNC_ONE_MINUS_CONST_A 13
#define VX_OM_BLEND_FUNC_ALPHA_SAT 14

#define VX_DCR_OM_STATE_BEGIN VX_DCR_RASTER_STATE_END
#define VX_DCR_OM_CBUF_ADDR (VX_DCR_OM_STATE_BEGIN + 0)
#define VX_DCR_OM_CBUF_PITCH (VX_DCR_OM_STATE_BEGIN + 1)
#define VX_DCR_OM_CBUF_WRITEMASK (VX_DCR_OM_STATE_BEGIN + 2)
#define VX_DCR_OM_ZBUF_ADDR (VX_DCR_OM_STATE_BEGIN + 3)
#define VX_DCR_OM_ZBUF_PITCH (VX_DCR_OM_STATE_BEGIN + 4)
#define VX_DCR_OM_DEPTH_FUNC (VX_DCR_OM_STATE_BEGIN + 5)
#define VX_DCR_OM_DEPTH_WRITEMASK (VX_DCR_OM_STATE_BEGIN + 6)
#define VX_DCR_OM_STENCIL_FUNC (VX_DCR_OM_STATE_BEGIN + 7)
#define VX_DCR_OM_STENCIL_ZPASS (VX_DCR_OM_STATE_BEGIN + 8)
#define VX_DCR_OM_STENCIL_ZFAIL (VX_DCR_OM_STATE_BEGIN + 9)
#define VX_DCR_OM_STENCIL_FAIL (VX_DCR_OM_STATE_BEGIN + 10)
#define VX_DCR_OM_STENCIL_REF (VX_DCR_OM_STATE_BEGIN + 11)
#define VX_DCR_OM_STENCIL_MASK (VX_DCR_OM_STATE_BEGIN + 12)
#define VX_DCR_OM_STENCIL_WRITEMASK (VX_DCR_OM_STATE_BEGIN + 13)
#define VX_DCR_OM_BLEND_MODE (VX_DCR_OM_STATE_BEGIN + 14)
#define VX_DCR_OM_BLEND_FUNC (VX_DCR_OM_STATE_BEGIN + 15)
#define VX_DCR_OM_BLEND_CONST (VX_DCR_OM_STATE_BEGIN + 16)
#define VX_DCR_OM_LOGIC_OP (VX_DCR_OM_STATE_BEGIN + 17)
#define VX_DCR_OM_STATE_END (VX_DCR_OM_STATE_BEGIN + 18)
#define VX_DCR_OM_STATE_COUNT 18

/* Counters of the last draw, read with vx_mpm_query (free user-CSR numbers next to the RT ones). */
#define VX_CSR_RASTER_FRAGMENTS 0x7C8  /* fragments shaded            (needs VX_DCR_RT_STATS = 1) */
#define VX_CSR_RASTER_QUAD_TESTS 0x7C9 /* quad x primitive tests      (needs VX_DCR_RT_STATS = 1) */
#define VX_CSR_RASTER_TILES 0x7CA      /* tiles (= CTAs) of the draw                                */

/* ---- device-visible records: sim/common/graphics.h:38-71 ---------------------------------------- */
typedef struct { int32_t x, y, z; } rast_vec3e_t;     /* fixed 16.16 (FloatE, graphics.h:26-31)        */
typedef struct { int32_t x, y, z; } rast_attrib_t;    /* fixed 8.24  (FloatA): d(v0-v2), d(v1-v2), v2  */
typedef struct {
  rast_attrib_t z, r, g, b, a, u, v;
} rast_attribs_t;                                     /* 84 B */
typedef struct {
  rast_vec3e_t edges[3];                              /* 36 B: a, b, c of the three edge equations     */
  rast_attribs_t attribs;
} rast_prim_t;                                        /* 120 B */
typedef struct {
  uint16_t tile_x, tile_y;                            /* in tiles                                      */
  uint16_t pids_offset;                               /* in u32s, counted from the END of this header  */
  uint16_t pids_count;
} rast_tile_header_t;                                 /* 8 B; all headers first, then the pid lists    */

/* ---- kernel argument: tests/regression/draw3d/common.h:18-34 (SW_ENABLE) -------------------------
 * The three leading members are the reference's RasterDCRS / TexDCRS / OMDCRS objects
 * (graphics.h:90-181): plain u32 arrays indexed by (dcr address - state begin).  TexDCRS drops the
 * STAGE register from its array and keeps the selected stage last (graphics.h:150-180). */
typedef struct {
  uint32_t raster_dcrs[VX_DCR_RASTER_STATE_COUNT];                          /*   0 ..  24 */
  uint32_t tex_dcrs[VX_TEX_STAGE_COUNT][VX_DCR_TEX_STATE_COUNT - 1];        /*  24 .. 192 */
  uint32_t tex_stage;                                                       /* 192 .. 196 */
  uint32_t om_dcrs[VX_DCR_OM_STATE_COUNT];                                  /* 196 .. 268 */
  uint32_t log_num_tasks;                                                   /* 268        */
  uint64_t prim_addr;                                                       /* 272        */
  uint8_t depth_enabled, color_enabled, tex_enabled, tex_modulate;          /* 280 (C++ bool) */
  uint8_t sw_rast, sw_tex, sw_om;
  uint8_t pad_[1];
} draw3d_kernel_arg_t;                                                      /* 288 B */

#define DRAW3D_RASTER_DCR(arg, addr) ((arg)->raster_dcrs[(addr) - VX_DCR_RASTER_STATE_BEGIN])
#define DRAW3D_OM_DCR(arg, addr) ((arg)->om_dcrs[(addr) - VX_DCR_OM_STATE_BEGIN])
#define DRAW3D_TEX_DCR(arg, stage, addr) ((arg)->tex_dcrs[stage][(addr) - VX_DCR_TEX_STATE_BEGIN - 1])

/* ---- draw-call trace: third_party/cocogfx/include/cgltrace.hpp:14-165 --------------------------- */
enum { SRT_FORMAT_A8 = 1, SRT_FORMAT_L8 = 2, SRT_FORMAT_A8L8 = 3, SRT_FORMAT_R5G6B5 = 4,
       SRT_FORMAT_A8R8G8B8 = 5, SRT_FORMAT_A1R5G5B5 = 6, SRT_FORMAT_A4R4G4B4 = 8 }; /* format.hpp:24-37 */
enum { SRT_FILTER_NONE = 0, SRT_FILTER_NEAREST = 1, SRT_FILTER_LINEAR = 2 };
enum { SRT_ADDRESS_WRAP = 0, SRT_ADDRESS_CLAMP = 1 };
enum { SRT_ENVMODE_ADD = 0, SRT_ENVMODE_BLEND, SRT_ENVMODE_REPLACE, SRT_ENVMODE_MODULATE, SRT_ENVMODE_DECAL };

typedef struct {
  uint32_t color_enabled, color_format, color_writemask;
  uint32_t depth_test, depth_writemask, depth_format, depth_func;
  uint32_t stencil_test, stencil_func, stencil_zpass, stencil_zfail, stencil_fail;
  uint32_t stencil_ref, stencil_mask, stencil_writemask;
  uint32_t texture_enabled, texture_envmode, texture_minfilter, texture_magfilter;
  uint32_t texture_addressU, texture_addressV;
  uint32_t blend_enabled, blend_src, blend_dst;
  float texture_envcolor[4];
} srt_states_t; /* cgltrace.hpp:117-146, one u32 per field */

typedef struct { float pos[4]; float color[4]; float texcoord[2]; } srt_vertex_t; /* 40 B, :104-108 */
typedef struct { uint32_t i0, i1, i2; } srt_primitive_t;
typedef struct { int32_t left, right, top, bottom; float near_, far_; } srt_viewport_t;

typedef struct {
  srt_states_t states;
  uint32_t texture_id;
  srt_viewport_t viewport;
  uint32_t num_vertices, num_primitives;
  const srt_vertex_t* vertices;      /* dense; primitives index into it */
  const srt_primitive_t* primitives;
} srt_drawcall_t;

typedef struct {
  uint32_t id, format, width, height, size;
  const uint8_t* pixels;
} srt_texture_t;

typedef struct {
  uint32_t num_drawcalls, num_textures;
  srt_drawcall_t* drawcalls;
  srt_texture_t* textures;
  uint8_t* blob; /* owns everything the pointers above reach */
  uint64_t blob_size;
} srt_trace_t;

/* Loads a "SKBXTRC1" trace (tests/golden/make_raster_fixtures.py; plain or gzip'ed).  0 on success. */
int srt_trace_load(const char* path, srt_trace_t* out);
/* Loads one of the reference's own .cgltrace files (Boost XML archive of cocogfx::CGLTrace,
 * third_party/cocogfx/src/cgltrace.cpp:24-165) without Boost.  Same result as the flat form. */
int srt_trace_load_cgltrace(const char* path, srt_trace_t* out);
/* Same from memory; the blob is copied. */
int srt_trace_from_blob(const void* blob, uint64_t size, srt_trace_t* out);
void srt_trace_free(srt_trace_t* trace);
const srt_texture_t* srt_trace_texture(const srt_trace_t* trace, uint32_t id);
const srt_drawcall_t* srt_drawcall_at(const srt_trace_t* trace, uint32_t index); /* NULL when out of range */

/* ---- host pre-pass: triangle setup + tile binning (role of graphics::Binning,
 * sim/common/gfxutil.cpp:103-276).  Same bytes as the reference for primbuf; tilebuf holds the same
 * tiles with the same pid lists in the same (tile_x, tile_y)-lexicographic order. ------------------ */
/* The reference's tile header holds pids_offset and pids_count in 16 bits (graphics.h:66-71), which
 * overflows once a draw bins more than ~64 K (tile, primitive) pairs -- any B200-sized frame.  The wide
 * layout is this back-end's extension for those: selected per draw with VX_DCR_RASTER_EXT_TBUF_FORMAT. */
#define SRT_TILEBUF_REFERENCE 0 /* rast_tile_header_t[num_tiles], then u32 pids                        */
#define SRT_TILEBUF_WIDE 1      /* srt_tile_header_wide_t[num_tiles], then u32 pids                    */
typedef struct {
  uint16_t tile_x, tile_y;
  uint32_t pids_count;
  uint32_t pids_offset; /* in u32s, from the first pid after the last header */
  uint32_t reserved;
} srt_tile_header_wide_t; /* 16 B */
#define VX_DCR_RASTER_EXT_TBUF_FORMAT 0x804 /* back-end DCR (outside the reference's ranges); 0 after open */
#define VX_DCR_RASTER_EXT_GENERIC 0x805     /* 1 = never use the state-specialised kernels (A/B testing)   */
#define VX_DCR_RASTER_EXT_PPT 0x806         /* pixels per thread: 0 = by grid size, 4 = quad, 1 = pixel    */
#define VX_DCR_RASTER_EXT_TILE_LOGSIZE 0x80D /* log2 tile size the app binned its tile buffers with, 0 = undeclared:
                                                vx_start refuses any other value than RASTER_TILE_LOGSIZE, because the
                                                kernel is built for that size and the buffer formats do not carry it */

typedef struct {
  uint8_t* tilebuf;   uint64_t tilebuf_size;
  uint8_t* primbuf;   uint64_t primbuf_size;
  uint32_t num_tiles; uint32_t num_prims;
  uint64_t num_tile_prims; /* sum of pids_count */
  uint32_t format;         /* SRT_TILEBUF_* */
} srt_binning_t;
/* Returns 0; -1 on bad arguments / out of memory; -2 when `format` is SRT_TILEBUF_REFERENCE and a
 * 16-bit field would overflow (the reference would truncate silently). */
int srt_draw3d_binning(const srt_drawcall_t* draw, uint32_t width, uint32_t height, uint32_t tile_logsize,
                       uint32_t format, srt_binning_t* out);
/* Same, keeping only the tile rows of one rank: tile_y % num_ranks == rank (multi-GPU sharding). */
int srt_draw3d_binning_rank(const srt_drawcall_t* draw, uint32_t width, uint32_t height, uint32_t tile_logsize,
                            uint32_t format, uint32_t rank, uint32_t num_ranks, srt_binning_t* out);
void srt_binning_free(srt_binning_t* b);

/* Texture upload image: level 0 followed by the box-filtered mip chain, and the per-level offsets the
 * app writes to VX_DCR_TEX_MIPOFF (cocogfx GenerateMipmaps, third_party/cocogfx/src/blitter.cpp:125-220;
 * draw3d samples level 0 only, tests/regression/draw3d/kernel.cpp:160-164). */
uint32_t srt_format_bpp(uint32_t format);
uint32_t srt_format_to_vx(uint32_t format); /* graphics::toVXFormat, gfxutil.cpp:280-296; 0xffffffff = invalid */

#ifdef __cplusplus
}
#endif

#endif /* SKYBOX_DRAW3D_H */
