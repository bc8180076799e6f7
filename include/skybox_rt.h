/*
 * skybox_rt.h -- contract of the `raytrace` regression app on the B200 back-end.
 *
 * The reference tree (/root/reference) has NO ray-tracing extension:
 * nothing in this file replaces a reference declaration one-to-one.  It is shaped after
 * the per-app `common.h` the reference's graphics apps share between host and device
 * (tests/regression/draw3d/common.h:18-34, raster/common.h:8-20, tex/common.h:14-24):
 * one POD kernel_arg_t with device addresses as uint64_t, uploaded with vx_upload_bytes
 * and handed to the device by vx_start.  The flat buffers it points to play the role
 * that primbuf/tilebuf (sim/common/gfxutil.cpp:253-271) play for draw3d: the output of
 * a host pre-pass ("the builder") that the device consumes as is.
 *
 * Everything here is plain C and is shared, byte for byte, by
 *   - the host app          (skybox_rt_b200/apps/raytrace/main.c)
 *   - the CUDA back-end     (skybox_rt_b200/csrc/)
 *   - the CPU oracle        (oracle/ref_rt.c; test infrastructure only)
 */
#ifndef SKYBOX_RT_H
#define SKYBOX_RT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------
 * Scene input: triangle soup.  prim id = index in this array.
 * ---------------------------------------------------------------------------------- */
typedef struct {
  float v0[3], v1[3], v2[3];
  uint32_t albedo; /* A8R8G8B8, a<<24|r<<16|g<<8|b (cocogfx color.hpp:24-33) */
} srt_triangle_t; /* 40 B */

/* ------------------------------------------------------------------------------------
 * Builder output #1: flat BVH2, 32 B per node, root = node 0.
 *   count == 0 : interior, children are nodes left_first and left_first+1
 *   count  > 0 : leaf, triangle records [left_first, left_first+count)
 * Boxes are the exact fp32 min/max of the contained vertices (no padding).
 * ---------------------------------------------------------------------------------- */
typedef struct {
  float bmin[3];
  uint32_t left_first;
  float bmax[3];
  uint32_t count;
} srt_bvh2_node_t; /* 32 B */

#define SRT_MAX_LEAF_TRIS 4

/* ------------------------------------------------------------------------------------
 * Builder output #2: triangle records in leaf order, 48 B = 3 x float4.
 *   e1 = v1 - v0, e2 = v2 - v0, each one fp32 subtraction per component, done ONCE by
 *   the builder so every consumer intersects the identical operands.
 * ---------------------------------------------------------------------------------- */
typedef struct {
  float v0[3];
  uint32_t prim_id; /* index into the input soup */
  float e1[3];
  uint32_t albedo;
  float e2[3];
  uint32_t pad;
} srt_tri_rec_t; /* 48 B */

/* Explicit rays (mode SRT_MODE_RAYS) and hit records. */
typedef struct {
  float o[3];
  float tmin;
  float d[3];
  float tmax;
} srt_ray_t; /* 32 B */

#define SRT_MISS 0xFFFFFFFFu
typedef struct {
  float t, u, v;
  uint32_t prim_id; /* SRT_MISS when nothing was hit (then t = tmax, u = v = 0) */
} srt_hit_t; /* 16 B */

/* ------------------------------------------------------------------------------------
 * kernel_arg_t
 * ---------------------------------------------------------------------------------- */
enum {
  SRT_MODE_PRIMARY = 0, /* closest hit + N.L, no shadow ray                     */
  SRT_MODE_SHADOW  = 1, /* + one shadow ray per lit hit                         */
  SRT_MODE_BOUNCE  = 2, /* + one diffuse bounce with its own shadow ray         */
  SRT_MODE_RAYS    = 3  /* explicit ray list -> hit records, no shading         */
};

typedef struct {
  uint32_t num_tasks;   /* like tex/common.h: row strips for the CPU kernel; ignored by CUDA */
  uint32_t dst_width;
  uint32_t dst_height;
  uint32_t cbuf_stride; /* 4                                                   */
  uint32_t cbuf_pitch;  /* dst_width * 4                                       */
  uint32_t mode;        /* SRT_MODE_*                                          */
  uint32_t spp;         /* samples per pixel (>= 1)                            */
  uint32_t seed;

  uint64_t cbuf_addr;   /* A8R8G8B8 framebuffer (this rank's strip buffer when num_ranks > 1) */
  uint64_t node_addr;   /* srt_bvh2_node_t[num_nodes]                          */
  uint64_t tri_addr;    /* srt_tri_rec_t[num_tris]                             */
  uint64_t ray_addr;    /* srt_ray_t[num_rays]   (SRT_MODE_RAYS)               */
  uint64_t hit_addr;    /* srt_hit_t[num_rays]   (SRT_MODE_RAYS)               */
  uint64_t num_rays;
  uint32_t num_nodes;
  uint32_t num_tris;

  float cam_pos[3];
  float cam_fwd[3];     /* unit                                               */
  float cam_right[3];   /* unit                                               */
  float cam_up[3];      /* unit                                               */
  float tan_half_fov;   /* vertical                                           */
  float light_pos[3];
  float ambient;
  float sky[3];         /* radiance returned by a bounce ray that misses      */
  float ray_eps;        /* origin offset along the normal for secondary rays  */

  /* multi-GPU row partition: row y belongs to rank (y / rows_per_block) % num_ranks and
   * is stored in that rank's strip buffer at local row
   *   (y / rows_per_block / num_ranks) * rows_per_block + y % rows_per_block.
   * num_ranks == 1 (or 0) => the whole frame, row y at cbuf_addr + y * cbuf_pitch. */
  uint32_t rank_id;
  uint32_t num_ranks;
  uint32_t rows_per_block;
  uint32_t reserved;
} kernel_arg_t;

/* rows owned by `rank` (size of its strip buffer in rows) */
static inline uint32_t srt_rank_rows(uint32_t height, uint32_t rows_per_block,
                                     uint32_t num_ranks, uint32_t rank) {
  uint32_t n = 0, y;
  if (num_ranks <= 1) return height;
  for (y = 0; y < height; y += rows_per_block) {
    if ((y / rows_per_block) % num_ranks == rank) {
      uint32_t r = height - y;
      n += r < rows_per_block ? r : rows_per_block;
    }
  }
  return n;
}

/* ------------------------------------------------------------------------------------
 * Kernel image.  The reference's hkernel holds a .vxbin: u64 min_vma, u64 max_vma, then a
 * RISC-V image (kernel/scripts/vxbin.py:54-75).  The CUDA back-end keeps the 16-byte
 * header and replaces the image by this descriptor naming a built-in pipeline.
 * ---------------------------------------------------------------------------------- */
#define SRT_KERNEL_MAGIC 0x5452584f42594b53ull /* "SKYBOXRT" little endian */
enum {
  SRT_PIPELINE_RAYTRACE = 1, /* kernel_arg_t above                                 */
  SRT_PIPELINE_DRAW3D   = 2  /* raster path (SURVEY.md section 8 group R)          */
};
typedef struct {
  uint64_t min_vma;
  uint64_t max_vma;
  uint64_t magic;
  uint32_t pipeline;
  uint32_t version;
} srt_kernel_image_t; /* 32 B */

/* ------------------------------------------------------------------------------------
 * Performance counters readable with vx_mpm_query(dev, addr, core_id, &v).
 * VX_CSR_MCYCLE / VX_CSR_MINSTRET keep the reference numbers (hw/rtl/VX_types.vh:73,77).
 * The RT counters use free user-CSR numbers.  All are totals for the last vx_start.
 * ---------------------------------------------------------------------------------- */
#define VX_CSR_MCYCLE          0xB00 /* SM clock cycles of the last run (elapsed ns * clock) */
#define VX_CSR_MINSTRET        0xB02 /* rays traced (the unit of retired work here)          */
#define VX_CSR_RT_RAYS_PRIMARY 0x7C0
#define VX_CSR_RT_RAYS_SHADOW  0x7C1
#define VX_CSR_RT_RAYS_BOUNCE  0x7C2
#define VX_CSR_RT_NODES        0x7C3 /* wide nodes fetched   (only when SRT_STATS is on)     */
#define VX_CSR_RT_TRIS         0x7C4 /* triangles tested     (only when SRT_STATS is on)     */
#define VX_CSR_RT_ELAPSED_NS   0x7C5 /* device time of the last run, CUDA events             */
#define VX_CSR_RT_NODE_BYTES   0x7C6 /* bytes per node fetch of the kernel variant that ran  */
#define VX_CSR_RT_LAUNCHES     0x7C7 /* kernels launched by the last vx_start                */
#define VX_CSR_RT_WARP_NODE_STEPS 0x7C8 /* warp-level node steps (STATS): NODES / this = lanes per node step */
#define VX_CSR_RT_WARP_TRI_STEPS  0x7C9 /* warp-level triangle steps (STATS)                                  */

/* ------------------------------------------------------------------------------------
 * Performance monitor, reference numbering (hw/rtl/VX_types.vh:27, :35-40, :66-247).
 * VX_DCR_BASE_MPM_CLASS selects what the class CSRs 0xB03.. mean; the reference's stub writes
 * it from the VORTEX_PROFILING environment variable before every vx_start
 * (runtime/stub/vortex.cpp:142-149) and so does this library.  While the class is non-zero
 * the kernels run their counting instantiations and VX_CSR_MCYCLE / VX_CSR_MINSTRET / the
 * class CSRs ACCUMULATE over every run since vx_dev_open (free-running, like the hardware);
 * with class 0 they describe the last vx_start.  core_id = SM index: MINSTRET is the work
 * (rays, fragments) started on that SM, MCYCLE is the same for every SM, and the TEX / RASTER
 * / OM / MEM class counters are device totals reported by every core -- the reference sums
 * them over cores and divides by the core count (runtime/stub/utils.cpp:726-729, :747-750).
 * Latency, stall and cache-miss CSRs of a class read 0: nothing inside a CUDA kernel measures
 * them (ncu does; profiles/).
 * ---------------------------------------------------------------------------------- */
#define VX_DCR_BASE_MPM_CLASS   0x005
#define VX_DCR_MPM_CLASS_NONE   0
#define VX_DCR_MPM_CLASS_CORE   1
#define VX_DCR_MPM_CLASS_MEM    2
#define VX_DCR_MPM_CLASS_TEX    3
#define VX_DCR_MPM_CLASS_RASTER 4
#define VX_DCR_MPM_CLASS_OM     5
#define VX_CSR_MPM_USER         0xB03 /* first class-dependent CSR */
#define VX_CSR_MPM_USER_END     0xB20
#define VX_CSR_MPM_LOADS        0xB0D /* CORE: records read  (nodes + triangles | tile, primitive, texel, framebuffer) */
#define VX_CSR_MPM_STORES       0xB0E /* CORE: words written (pixels | framebuffer words)                             */
#define VX_CSR_MPM_MEM_READS    0xB18 /* MEM: same as LOADS  */
#define VX_CSR_MPM_MEM_WRITES   0xB19 /* MEM: same as STORES */
#define VX_CSR_MPM_MEM_LT       0xB1A
#define VX_CSR_MPM_TEX_READS    0xB03 /* TEX: texels fetched */
#define VX_CSR_MPM_TCACHE_READS 0xB06
#define VX_CSR_MPM_RASTER_READS 0xB03 /* RASTER: tile headers + primitive ids + primitive records read */
#define VX_CSR_MPM_RCACHE_READS 0xB06
#define VX_CSR_MPM_OM_READS     0xB03 /* OM: framebuffer words (colour, depth/stencil) read    */
#define VX_CSR_MPM_OM_WRITES    0xB04 /* OM: framebuffer words written                          */
#define VX_CSR_MPM_OCACHE_READS 0xB07
#define VX_CSR_MPM_OCACHE_WRITES 0xB08

/* ------------------------------------------------------------------------------------
 * Host builder (libskybox_rt_host: plain C, no CUDA) -- the pre-pass that plays the role
 * of graphics::Binning for this path.  Deterministic: same soup => same bytes.
 * ---------------------------------------------------------------------------------- */
typedef struct {
  srt_bvh2_node_t* nodes;
  srt_tri_rec_t* tris;
  uint32_t num_nodes;
  uint32_t num_tris;
} srt_bvh2_t;

/* Binned-SAH BVH2 (16 bins, leaves <= SRT_MAX_LEAF_TRIS).  Returns 0 on success. */
int srt_bvh2_build(const srt_triangle_t* soup, uint32_t num_tris, srt_bvh2_t* out);
void srt_bvh2_free(srt_bvh2_t* bvh);

/* Procedural scenes of BASELINE.md section 3.  Caller frees *soup with free(). */
int srt_scene_cornell(srt_triangle_t** soup, uint32_t* num_tris);
int srt_scene_terrain(uint32_t quads_x, uint32_t quads_y, uint32_t seed,
                      srt_triangle_t** soup, uint32_t* num_tris);
/* Camera / light presets for the two scene families, written into `arg`. */
void srt_camera_cornell(kernel_arg_t* arg);
void srt_camera_terrain(kernel_arg_t* arg);
/* Incoherent ray stress (config 5): origin uniform in [lo,hi], direction uniform on the
 * sphere by rejection; ray i depends only on (seed, first + i). */
void srt_rays_random(const float lo[3], const float hi[3], uint32_t seed,
                     uint64_t first, uint64_t count, srt_ray_t* rays);

#ifdef __cplusplus
}
#endif

#endif /* SKYBOX_RT_H */
