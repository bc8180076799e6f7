// rt_pipeline.cu -- the `raytrace` pipeline that vx_start runs for SRT_PIPELINE_RAYTRACE.
//
// Stands where the reference's device program stands: the code reached through vx_start ->
// processor_.run() (reference runtime/simx/vortex.cpp:172-193) that reads kernel_arg_t and runs
// kernel_body over all tasks (kernel/src/vx_spawn.c:157-322).  Everything is enqueued on the
// device stream; nothing here blocks the host.
#include <stdio.h>
#include <string.h>

#include "vx_internal.h"
// clang-format off
#include "rt_baseline.cuh"
// clang-format on

namespace srt {

struct RtState {
  unsigned long long* d_counters;  // CTR_COUNT u64, zeroed at every launch
};

static int rt_state(vx_device* dev, RtState** out) {
  if (!dev->rt) {
    RtState* s = new RtState();
    memset(s, 0, sizeof(*s));
    VX_CUDA_OK(cudaMalloc(&s->d_counters, sizeof(unsigned long long) * CTR_COUNT));
    dev->rt = s;
  }
  *out = dev->rt;
  return 0;
}

void rt_destroy(vx_device* dev) {
  if (!dev->rt) return;
  cudaFree(dev->rt->d_counters);
  delete dev->rt;
  dev->rt = nullptr;
}

static int validate(const kernel_arg_t& a) {
  if (a.node_addr == 0 || a.tri_addr == 0 || a.num_nodes == 0 || a.num_tris == 0) return -1;
  if (a.mode > SRT_MODE_RAYS) return -1;
  if (a.mode == SRT_MODE_RAYS) {
    if (a.num_rays && (a.ray_addr == 0 || a.hit_addr == 0)) return -1;
    return 0;
  }
  if (a.dst_width == 0 || a.dst_height == 0 || a.cbuf_addr == 0 || a.spp == 0) return -1;
  if (a.cbuf_stride != 4 || a.cbuf_pitch < a.dst_width * 4) return -1;
  if (a.num_ranks > 1 && (a.rank_id >= a.num_ranks || a.rows_per_block == 0)) return -1;
  return 0;
}

int rt_launch(vx_device* dev, const kernel_arg_t& a) {
  RtState* st;
  if (validate(a) != 0) {
    fprintf(stderr, "[VXDRV] raytrace: invalid kernel_arg_t\n");
    return -1;
  }
  if (rt_state(dev, &st) != 0) return -1;
  cudaStream_t s = dev->stream;
  const bool stats = dev->dcrs.count(VX_DCR_RT_STATS) && dev->dcrs[VX_DCR_RT_STATS] != 0;
  dev->last_launches = 0;
  dev->last_node_bytes = 32;  // baseline: one srt_bvh2_node_t per fetch
  VX_CUDA_OK(cudaMemsetAsync(st->d_counters, 0, sizeof(unsigned long long) * CTR_COUNT, s));

  if (a.mode == SRT_MODE_RAYS) {
    dev->last_rays_primary = a.num_rays;
    if (a.num_rays == 0) return 0;
    const uint64_t blocks = (a.num_rays + 255) / 256;
    if (stats) rt_baseline_rays_kernel<true><<<(unsigned)blocks, 256, 0, s>>>(a, st->d_counters);
    else rt_baseline_rays_kernel<false><<<(unsigned)blocks, 256, 0, s>>>(a, st->d_counters);
    VX_CUDA_OK(cudaGetLastError());
    dev->last_launches = 1;
    return 0;
  }

  const uint32_t rows = srt_rank_rows(a.dst_height, a.rows_per_block, a.num_ranks, a.rank_id);
  dev->last_rays_primary = (uint64_t)rows * a.dst_width * a.spp;
  if (rows == 0) return 0;
  const uint32_t tiles = ((a.dst_width + 7) / 8) * ((rows + 3) / 4);  // one warp per 8x4 tile
  const uint32_t blocks = (tiles + 7) / 8;                            // 8 warps per block
  if (stats) rt_baseline_kernel<true><<<blocks, 256, 0, s>>>(a, rows, st->d_counters);
  else rt_baseline_kernel<false><<<blocks, 256, 0, s>>>(a, rows, st->d_counters);
  VX_CUDA_OK(cudaGetLastError());
  dev->last_launches = 1;
  return 0;
}

// read back the device counters of the last run (called after the stream is idle)
int rt_read_counters(vx_device* dev) {
  if (!dev->rt) return -1;
  unsigned long long h[CTR_COUNT];
  VX_CUDA_OK(cudaMemcpy(h, dev->rt->d_counters, sizeof(h), cudaMemcpyDeviceToHost));
  for (int i = 0; i < CTR_COUNT; ++i) dev->last_counters[i] = h[i];
  dev->last_counters_valid = true;
  return 0;
}

}  // namespace srt
