// vx_internal.h -- private state of the CUDA back-end behind include/vortex.h.
//
// Mirrors the shape of the reference drivers: a `vx_device` object per opened device that owns
// memory, a DCR file and the launch machinery (reference runtime/simx/vortex.cpp:32-246), and a
// small `vx_buffer` record per allocation (reference runtime/common/callbacks.inc:14-18).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <mutex>

#include "skybox_draw3d.h"
#include "skybox_rt.h"
#include "vortex.h"
#include "vortex_cuda.h"

struct vx_device;

struct vx_buffer {
  vx_device* dev;
  void* dptr;             // cudaMalloc'ed storage
  uint64_t size;
  uint64_t reported_addr; // == (uint64_t)dptr except for vx_mem_reserve'd kernel images
  int flags;
  uint8_t* shadow;        // host copy of small host-written buffers (kernel args, kernel images)
  uint64_t generation;    // bumped by every vx_copy_to_dev (invalidates derived device data)
  uint64_t uid;           // device-wide allocation number, never reused (the pool does reuse addresses)
  uint32_t tile_logsize;  // tile buffers made by vx_cuda_draw3d_binning: the tile size they were binned with, else 0
  vx_buffer* prev;
  vx_buffer* next;
};

constexpr uint64_t kShadowMaxBytes = 4096;
constexpr uint64_t kStageSlotBytes = 4ull << 20;   // largest upload that is staged
constexpr uint64_t kStageRingBytes = 32ull << 20;  // page-locked staging ring, allocated by the byte (256-byte granules)
constexpr int kStageEntries = 256;                 // staged transfers that may be in flight

// Device-side counters written by the kernels (one u64 each).
enum {
  CTR_RAYS_SHADOW = 0,
  CTR_RAYS_BOUNCE = 1,
  CTR_NODES = 2,
  CTR_TRIS = 3,
  CTR_CURSOR = 4,  // job cursor of the persistent kernels
  CTR_WARP_NODE_STEPS = 5,  // warp-level node steps / triangle steps (STATS instantiations of the Node4 kernels)
  CTR_WARP_TRI_STEPS = 6,
  CTR_PER_SM = 8,  // rays started on each SM (%smid), STATS instantiations only
  CTR_COUNT = 8 + 256
};

// Accumulators of the performance monitor: slots VX_PERF_* of include/vortex_cuda.h, free-running over every
// profiled run since vx_dev_open, like the reference's MPM CSRs (hw/rtl/VX_types.vh:66-247).
constexpr int kMaxSms = 256;

namespace srt {
struct RtState;      // rt_pipeline.cu
struct RasterState;  // raster_pipeline.cu
}

struct vx_device {
  int cuda_dev;
  cudaDeviceProp prop;
  cudaStream_t stream;
  cudaEvent_t ev_start, ev_stop;
  // asynchronous read-back (vx_cuda_copy_from_dev_async): a second stream + a ring of completion events
  cudaStream_t copy_stream;     // created by the first asynchronous copy
  cudaEvent_t copy_fork;        // launch stream -> copy stream ordering
  cudaEvent_t copy_done[16];
  uint64_t copy_next_ticket;
  // staged uploads (vx_copy_to_dev of up to kStageSlotBytes): page-locked ring the caller's bytes are copied into, so
  // that the transfer is queued on the launch stream and the host does not wait for the kernels in front of it.  The
  // ring is handed out by the byte: a frame of small uploads (a draw's tile / primitive buffers are tens of KB) takes
  // a few MB, so the host can run several frames ahead of the device before it has to wait for a transfer
  uint8_t* stage_ring;          // kStageRingBytes, cudaHostAlloc'ed at the first staged upload
  bool stage_mapped;            // the ring is visible to kernels at its host address (copy kernel instead of DMA)
  cudaEvent_t stage_done[kStageEntries];                        // FIFO of the transfers in flight: completion event ...
  struct { uint64_t begin, end; } stage_span[kStageEntries];   // ... and the bytes of the ring each one reads
  uint32_t stage_head, stage_count;                             // oldest entry, entries in flight
  uint64_t stage_tail;                                          // end of the newest entry's bytes
  bool run_recorded;            // ev_start/ev_stop bracket the last vx_start
  std::map<uint32_t, uint32_t> dcrs;
  vx_buffer* buffers;           // intrusive list of live allocations
  uint64_t next_uid;            // uid of the next allocation
  uint64_t bytes_used;
  srt::RtState* rt;
  // results of the last run, filled lazily by vx_mpm_query / vx_dump_perf
  uint64_t last_rays_primary;
  uint64_t last_counters[CTR_COUNT];
  bool last_counters_valid;
  int last_pipeline;
  uint32_t last_launches;       // kernels enqueued by the last vx_start
  uint32_t last_node_bytes;     // bytes per node fetch of the kernel variant that ran
  uint64_t last_rt_stores;      // pixels (or hit records) written by the last run
  // raster pipeline (draw3d)
  srt::RasterState* raster;
  uint32_t last_raster_tiles;       // tiles (= CTAs) of the last draw
  uint64_t last_raster_fragments;   // fragments shaded    (only when VX_DCR_RT_STATS is on)
  uint64_t last_raster_tests;       // quad x primitive coverage tests (only when VX_DCR_RT_STATS is on)
  uint64_t last_raster_cls[4];      // texels fetched, framebuffer words read, written, raster records read
  // performance monitor
  bool harvested;                   // the last run's counters are in last_* (and in acc when it ran profiled)
  int run_mpm_class;                // VX_DCR_BASE_MPM_CLASS latched by the last vx_start
  uint64_t last_instrs_sm[kMaxSms]; // work units (rays / fragments) per SM of the last run, counting runs only
  uint64_t acc[VX_PERF_COUNT];
  uint64_t acc_instrs_sm[kMaxSms];
};

// DCRs owned by this back-end (outside every range the reference decodes,
// hw/rtl/VX_types.vh:23-29, :332-339, :353-359, :439-458).
#define VX_DCR_RT_BASE    0x800
#define VX_DCR_RT_VARIANT 0x800 /* 0 = default: the measured best per kind of work (frames: 4, ray lists: 7);
                                   1 = baseline BVH2 megakernel, 2 = Node4 persistent megakernel while-while,
                                   3 = ray pool (ray lists), 4 / 5 / 6 = wavefront pipeline with while-while /
                                   vote / vote + deferred-triangle trace kernels, 7 = Node4 persistent megakernel,
                                   vote scheduler (the round-1 default)                                       */
#define VX_DCR_RT_STATS   0x801 /* 1 = count node fetches / triangle tests (slower)               */
#define VX_DCR_RT_KEEP_ACTIVE 0x802 /* tuning: refill threshold of the persistent kernels (1..32), 0 = default */
#define VX_DCR_RT_PREFETCH 0x803 /* 1 = warm L2 with the scene before tracing when it fits (measured: no gain, off by default) */
#define VX_DCR_RT_SPLIT_SAMPLES 0x807 /* multi-sample frames as (pixel, sample) jobs + ordered resolve: 0 = when spp > 1,
                                        1 = never (a pixel's samples run sequentially in one lane), 2 = always */
#define VX_DCR_RT_CHAINS  0x809 /* wavefront pipeline: independent chains a frame is split into so that the tail of one
                                   chain's trace kernel overlaps the other chain's next kernel (0 = default = 1;
                                   2..4 forced)                                      */
#define VX_DCR_RT_PASS_JOBS 0x80A /* wavefront pipeline: jobs (pixels x samples) one pass may hold, 0 = default 2^27; a frame
                                    with more runs as several passes over sample ranges (tests force this path)      */
#define VX_DCR_RT_FUSE_BOUNCE 0x80B /* wavefront pipeline: 0 = bounce hits shaded and their shadow rays traced inside the shadow +
                                     bounce trace launch (default), 1 = as separate launches K4 / K5 (A/B)             */
#define VX_DCR_RT_LBVH_ROTATE 0x80C /* vx_cuda_rt_build_lbvh: SAH tree-rotation passes over the radix tree, 0 = default (2),
                                     k > 0 = k - 1 passes (1 = none)                                               */
#define VX_DCR_RT_REPACK  0x808 /* BVH2 -> Node4: 0 = on the device in one cooperative launch (rt_repack_device.cu), 1 = on the
                                   host (rt_repack.cpp), 2 = on the device, one set of launches per tree level (round 1); A/B */
#define VX_DCR_RT_END     0x810

inline uint32_t vx_dev_dcr(const vx_device* dev, uint32_t addr) {
  auto it = dev->dcrs.find(addr);
  return it == dev->dcrs.end() ? 0u : it->second;
}
// the counting instantiations run when asked for directly or when the performance monitor is on
inline bool vx_dev_counting(const vx_device* dev) {
  return vx_dev_dcr(dev, VX_DCR_RT_STATS) != 0 || vx_dev_dcr(dev, VX_DCR_BASE_MPM_CLASS) != 0;
}

namespace srt {
// rt_pipeline.cu
int rt_launch(vx_device* dev, const kernel_arg_t& arg);
void rt_destroy(vx_device* dev);
int rt_read_counters(vx_device* dev);
void rt_buffer_freed(vx_device* dev, const vx_buffer* b);  // drops derived data (Node4 image) built from b
// raster_pipeline.cu
int raster_launch(vx_device* dev, const draw3d_kernel_arg_t& arg);
int raster_read_counters(vx_device* dev);
int raster_take_fault(vx_device* dev);  // -1 (+ a message) when a tile kernel met data outside its buffers since the last call
void raster_destroy(vx_device* dev);
}  // namespace srt

#define VX_CUDA_OK(expr)                                                                          \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess) {                                                                      \
      fprintf(stderr, "[VXDRV] CUDA error %s at %s:%d: %s\n", cudaGetErrorName(_e), __FILE__,     \
              __LINE__, cudaGetErrorString(_e));                                                  \
      return -1;                                                                                  \
    }                                                                                             \
  } while (0)
