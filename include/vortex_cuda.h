/*
 * vortex_cuda.h -- CUDA-specific additions to the vx_* C ABI (no counterpart in the reference).
 *
 * The reference host API (runtime/include/vortex.h) has no notion of a device-side queue that a
 * second library could share.  On B200 the framebuffer gather between GPUs (NCCL) and device-side
 * timing (CUDA events) must be ordered with the kernels vx_start enqueues, so the back-end exposes
 * its launch stream.  Plain C: the stream is passed as void* (a cudaStream_t).
 */
#ifndef SKYBOX_RT_B200_VORTEX_CUDA_H
#define SKYBOX_RT_B200_VORTEX_CUDA_H

#include "vortex.h"

#ifdef __cplusplus
extern "C" {
#endif

/* The CUDA stream vx_start launches on and vx_copy_* are ordered with (owned by the device). */
int vx_cuda_stream(vx_device_h hdevice, void** stream);

/* Page-lock a host range the app reuses for vx_copy_to_dev / vx_copy_from_dev (cudaHostRegister), so
 * those copies run as direct DMA at link rate instead of through the driver's staging buffer.  Optional:
 * the copies work on any host pointer.  The range must stay mapped until vx_cuda_host_unregister. */
int vx_cuda_host_register(vx_device_h hdevice, void* ptr, uint64_t size);
int vx_cuda_host_unregister(vx_device_h hdevice, void* ptr);

/* Asynchronous read-back for render loops.  vx_copy_from_dev blocks until the bytes are on the host, so a loop of
 * vx_start / vx_copy_from_dev leaves the device idle during every transfer.  This variant orders the copy after
 * everything queued on the launch stream so far, runs it on a second stream and returns at once with a ticket:
 * frames started afterwards overlap the transfer.  host_ptr should be page-locked (vx_cuda_host_register) -- a
 * pageable destination makes the copy synchronous.  The app must not let later work overwrite the source range, nor
 * read host_ptr, before vx_cuda_copy_wait(ticket) returns (double-buffer the render target).  Up to 16 copies
 * may be in flight; vx_dev_close waits for all of them. */
int vx_cuda_copy_from_dev_async(void* host_ptr, vx_buffer_h hbuffer, uint64_t src_offset, uint64_t size, uint64_t* ticket);
int vx_cuda_copy_wait(vx_device_h hdevice, uint64_t ticket);

/* Totals of the performance monitor (VX_DCR_BASE_MPM_CLASS != 0, include/skybox_rt.h) accumulated over every
 * profiled run since vx_dev_open: out[i] = slot i below for i < count.  vx_dump_perf prints from these. */
enum {
  VX_PERF_CYCLES = 0,   /* SM cycles (elapsed device time x SM clock)            */
  VX_PERF_NS,           /* elapsed device time, ns                                */
  VX_PERF_RUNS,         /* profiled vx_start calls                                */
  VX_PERF_RT_PRIMARY, VX_PERF_RT_SHADOW, VX_PERF_RT_BOUNCE, /* rays by class      */
  VX_PERF_RT_NODES,     /* BVH node records fetched                               */
  VX_PERF_RT_TRIS,      /* triangles tested                                       */
  VX_PERF_RT_NODE_BYTES,/* bytes of those node fetches                            */
  VX_PERF_RT_STORES,    /* pixels / hit records written                           */
  VX_PERF_FRAGMENTS,    /* fragments shaded                                       */
  VX_PERF_TESTS,        /* quad x primitive coverage tests                        */
  VX_PERF_TILES,        /* tiles rendered                                         */
  VX_PERF_TEX_READS,    /* texels fetched                                         */
  VX_PERF_OM_READS,     /* framebuffer words read                                 */
  VX_PERF_OM_WRITES,    /* framebuffer words written                              */
  VX_PERF_RASTER_READS, /* tile headers + primitive ids + primitive records read  */
  VX_PERF_COUNT
};
int vx_cuda_perf_totals(vx_device_h hdevice, uint64_t* out, uint32_t count);

/* Triangle setup + tile binning on the device: the role of graphics::Binning (sim/common/gfxutil.cpp:103-276),
 * which the reference runs on the host, for frames where the host pre-pass and the upload of its output would
 * dominate.  `vertices` holds srt_vertex_t[num_vertices] (40 B: pos xyzw, color rgba, uv -- CGLTrace::vertex_t)
 * and `primitives` srt_primitive_t[num_primitives] (3 x u32), both already on the device.  On success
 * *tilebuf / *primbuf are new buffers (free with vx_mem_free; NULL when empty) holding exactly the bytes
 * srt_draw3d_binning_rank() produces for the same input (include/skybox_draw3d.h: SRT_TILEBUF_* formats,
 * rank / num_ranks tile-row dealing).  Returns 0; -1 bad arguments / out of memory / an index past the vertex
 * buffer; -2 when SRT_TILEBUF_REFERENCE's 16-bit fields would overflow.  Enqueued on the launch stream; waits
 * once for the output sizes.
 * tile_logsize: buffers handed to vx_start must be binned with RASTER_TILE_LOGSIZE (5, 32x32 pixels): the render
 * kernel is built for that size like the reference's raster unit (hw/rtl/VX_config.vh:477-478) and neither tile
 * buffer format carries the size.  The tile buffer returned here remembers its size and vx_start refuses it (-1, with
 * a message) when it is not 5; for tile buffers binned on the host and uploaded, the app can declare the size in
 * VX_DCR_RASTER_EXT_TILE_LOGSIZE (include/skybox_draw3d.h) for the same check.  Other sizes (1..8) exist for tooling
 * and for the byte-for-byte comparison with graphics::Binning at its other sizes. */
typedef struct {
  uint32_t num_tiles, num_prims;
  uint64_t num_tile_prims; /* sum of pids_count */
  uint64_t tilebuf_size, primbuf_size;
  uint32_t format;
} srt_binning_info_t;
int vx_cuda_draw3d_binning(vx_device_h hdevice, vx_buffer_h vertices, uint32_t num_vertices, vx_buffer_h primitives,
                           uint32_t num_primitives, float near_, float far_, uint32_t width, uint32_t height,
                           uint32_t tile_logsize, uint32_t format, uint32_t rank, uint32_t num_ranks,
                           vx_buffer_h* tilebuf, vx_buffer_h* primbuf, srt_binning_info_t* info);

/* BVH2 build on the device for the ray-tracing path: `soup` holds srt_triangle_t[num_triangles] (include/skybox_rt.h)
 * already on the device; on success *nodes / *tris are new buffers (free with vx_mem_free) in exactly the two
 * formats the host builder srt_build_bvh2() emits -- srt_bvh2_node_t[info->num_nodes], root = node 0, leaves of
 * <= SRT_MAX_LEAF_TRIS triangles, and srt_tri_rec_t[num_triangles] in leaf order -- so they go straight into
 * kernel_arg_t.node_addr / tri_addr.  A linear BVH (Morton order + radix tree), not the host's binned SAH: the
 * tree differs, what vx_start renders from it does not (closest hit is tree-independent by construction).
 * Some node slots are unreachable (zero).  Returns 0, or -1 on bad arguments / out of memory.  Asynchronous on the
 * launch stream. */
typedef struct {
  uint32_t num_nodes, num_tris;
} srt_lbvh_info_t;
int vx_cuda_rt_build_lbvh(vx_device_h hdevice, vx_buffer_h soup, uint32_t num_triangles, vx_buffer_h* nodes,
                          vx_buffer_h* tris, srt_lbvh_info_t* info);

/* The same build for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The caller passes the input AND both outputs: `nodes` must hold (num_triangles == 1 ? 1 :
 * 2 x num_triangles - 1) x sizeof(srt_bvh2_node_t) bytes and `tris` num_triangles x sizeof(srt_tri_rec_t) bytes,
 * all three distinct buffers of this device; on success they hold exactly the bytes vx_cuda_rt_build_lbvh produces
 * for the same soup and info has the same counts.  The call allocates and frees nothing: the builder's scratch is
 * one fixed array inside the library (5.5 MiB of device memory that every GPU which loads libvortex.so keeps until
 * the process exits), sized for VX_CUDA_LBVH_NOALLOC_MAX_TRIS triangles.  Returns 0; -1 on bad arguments or
 * buffers that are too small; -3, with a message on stderr, when num_triangles exceeds the maximum (use
 * vx_cuda_rt_build_lbvh then).  Asynchronous on the launch stream.  All calls on one GPU share the array, which
 * keeps the previous call's contents (the builder clears what it needs at zero itself): calls that can overlap on
 * the GPU -- two vx_device handles opened on the same GPU, whose launch streams are independent -- race and must
 * be serialised by the application (vx_ready_wait on the first before the second starts). */
#define VX_CUDA_LBVH_NOALLOC_MAX_TRIS 65536u
int vx_cuda_rt_build_lbvh_noalloc(vx_device_h hdevice, vx_buffer_h soup, uint32_t num_triangles, vx_buffer_h nodes,
                                  vx_buffer_h tris, srt_lbvh_info_t* info);

/* The wide-BVH image (128-byte SoA Node4 records, csrc/rt_repack.h) the last ray-tracing vx_start traversed, for
 * tests and tooling: *size = its size in bytes; copied to dst when dst != NULL and capacity suffices. */
int vx_cuda_rt_node4_image(vx_device_h hdevice, void* dst, uint64_t capacity, uint64_t* size);

/* Fill `size` bytes of a buffer at `offset` with a 32-bit pattern on the device (both multiples of 4), ordered on the
 * launch stream.  The reference's apps clear their render targets by uploading a host array of the clear value
 * (draw3d/main.cpp:478-490: 2 x W x H x 4 bytes over the host link per frame); an app that knows it runs on this
 * back-end clears here instead. */
int vx_cuda_fill32(vx_buffer_h hbuffer, uint64_t offset, uint64_t size, uint32_t value);

/* Read-bandwidth probe for roofline denominators measured on the box a benchmark runs on: #SM x 8 CTAs stream
 * `bytes` (a multiple of 16) of a private device buffer `repeats` times with 16-byte read-only loads, after one
 * untimed pass; *ms = device time of the timed passes (CUDA events on the launch stream).  bytes well inside the
 * L2 (<= 64 MiB on B200) measures L2 -> SM bandwidth, bytes far beyond it (>= 1 GiB) HBM read bandwidth. */
int vx_cuda_probe_read_bandwidth(vx_device_h hdevice, uint64_t bytes, uint32_t repeats, float* ms);

/* ---- multi-GPU: one process (or thread) per GPU, one vx_device each, scene replicated, pixels partitioned ----
 * No counterpart in the reference (single device).  A communicator binds this device to `num_ranks` peers over
 * NCCL (NVLink / NVSwitch inside a box).  `id` is VX_CUDA_COMM_ID_BYTES bytes made ONCE by
 * vx_cuda_comm_unique_id on any one rank and handed to the others by the application (a file, a pipe, an
 * environment variable, MPI ... -- the apps use a file, apps/raytrace/main.c).  vx_cuda_comm_init is collective:
 * every rank must call it.  libnccl.so.2 is loaded at the first of these calls; everything else in this library
 * works without it. */
#define VX_CUDA_COMM_ID_BYTES 128
typedef void* vx_comm_h;
int vx_cuda_comm_unique_id(void* id);
int vx_cuda_comm_init(vx_device_h hdevice, const void* id, uint32_t rank, uint32_t num_ranks, vx_comm_h* hcomm);
int vx_cuda_comm_destroy(vx_comm_h hcomm);

/* Assemble a frame whose rows were rendered rank by rank (kernel_arg_t.rank_id / num_ranks / rows_per_block:
 * row y lives on rank (y / rows_per_block) % num_ranks at local row (y / rows_per_block / num_ranks) *
 * rows_per_block + y % rows_per_block of that rank's strip buffer).  Collective, enqueued on the launch stream
 * behind the frame's kernels, returns at once: every rank passes its `strip` (>= srt_rank_rows() x width x 4
 * bytes); the root also passes `frame` (>= height x width x 4 bytes, rows at pitch width x 4; NULL elsewhere),
 * which holds the whole frame once the stream reaches this point (vx_ready_wait, or any later vx_copy_from_dev).
 * One NCCL group of N-1 send/receive pairs plus one de-interleave kernel on the root. */
int vx_cuda_frame_gather(vx_device_h hdevice, vx_comm_h hcomm, vx_buffer_h strip, vx_buffer_h frame, uint32_t width,
                         uint32_t height, uint32_t rows_per_block, uint32_t root);

/* The same exchange for ranks that render into FULL-SIZE frame buffers with their row blocks at the final positions
 * (the raster path deals 32-row tile rows round-robin): a sending rank packs its blocks into one strip and sends it
 * as one message, the root scatters the received strips to their rows of its own `frame`; its own rows never move.
 * Every rank passes its full-size buffer. */
int vx_cuda_frame_gather_inplace(vx_device_h hdevice, vx_comm_h hcomm, vx_buffer_h frame, uint32_t width, uint32_t height,
                                 uint32_t rows_per_block, uint32_t root);

/* Concatenating gather for record lists split over the ranks (hit records of a ray list): rank r's `bytes` bytes of
 * `src` land at offsets[r] of the root's `dst`; offsets / sizes (num_ranks entries each, sizes[r] == that rank's
 * `bytes`) are read on the root only.  Same ordering rules as vx_cuda_frame_gather. */
int vx_cuda_gather_bytes(vx_device_h hdevice, vx_comm_h hcomm, vx_buffer_h src, uint64_t bytes, vx_buffer_h dst,
                         const uint64_t* offsets, const uint64_t* sizes, uint32_t root);

/* CUDA device ordinal the handle was opened on. */
int vx_cuda_device(vx_device_h hdevice, int* ordinal);

#ifdef __cplusplus
}
#endif

#endif
