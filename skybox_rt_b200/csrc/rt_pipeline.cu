// rt_pipeline.cu -- the `raytrace` pipeline that vx_start runs for SRT_PIPELINE_RAYTRACE.
//
// Stands where the reference's device program stands: the code reached through vx_start ->
// processor_.run() (reference runtime/simx/vortex.cpp:172-193) that reads kernel_arg_t and runs
// kernel_body over all tasks (kernel/src/vx_spawn.c:157-322).  Everything is enqueued on the
// device stream; nothing here blocks the host.
#include <stdio.h>
#include <string.h>

#include <vector>

#include "rt_repack.h"
#include "vx_internal.h"
// clang-format off
#include "rt_baseline.cuh"
#include "rt_wide.cuh"
#include "rt_pool.cuh"
#include "rt_wave.cuh"
// clang-format on

namespace srt {

// The wide-BVH image of one uploaded scene.  Built on the first vx_start that sees a given
// (node buffer, triangle buffer) pair and reused until the app overwrites the node buffer
// (vx_copy_to_dev bumps its generation) -- the role the tile/primitive buffers of draw3d play per
// draw call (tests/regression/draw3d/main.cpp:190-211), except that the B200 layout is derived by
// the back-end instead of being uploaded by the app.
struct WideCache {
  uint64_t node_addr, tri_addr;
  uint64_t node_uid, tri_uid, node_gen, tri_gen;  // which allocations, and which contents of them
  uint32_t num_nodes, num_tris;
  float4* d_nodes4;
  uint32_t num_nodes4, depth;
  bool valid;
};

constexpr int kMaxChains = 4;

struct RtState {
  unsigned long long* d_counters;  // CTR_COUNT u64, zeroed at every launch
  WideCache wide;
  int wide_blocks_per_sm[12];      // occupancy of the persistent kernels, queried once
  int wide_split_blocks_per_sm[12];// same for the (pixel, sample)-job render kernels
  int pool_blocks_per_sm[4];
  uint2* d_spill;                  // stack spill area of the pool kernels (rarely touched)
  size_t spill_warps;
  float4* d_samples;               // per-sample radiance of split multi-sample frames, [spp][pixels]
  size_t samples_capacity;         // in float4
  // wavefront pipeline (rt_wave.cuh): queues and path state for `wave_jobs` jobs, one allocation
  char* d_wave;
  uint64_t wave_jobs, wave_sum_jobs;
  WaveBufs wave;
  WaveCtl* d_wave_ctl;             // kMaxChains control blocks
  cudaStream_t chain_stream[3];    // the extra chains of a frame run here (the first one on the device stream)
  cudaEvent_t chain_fork, chain_join[3];
  int wave_blocks_per_sm[24];      // occupancy per (stage, scheduler, stats) instantiation
};

static int rt_state(vx_device* dev, RtState** out) {
  if (!dev->rt) {
    RtState* s = new RtState();
    memset(s, 0, sizeof(*s));
    VX_CUDA_OK(cudaMalloc(&s->d_counters, sizeof(unsigned long long) * CTR_COUNT));
    VX_CUDA_OK(cudaMalloc(&s->d_wave_ctl, sizeof(WaveCtl) * kMaxChains));
    VX_CUDA_OK(cudaEventCreateWithFlags(&s->chain_fork, cudaEventDisableTiming));
    for (int i = 0; i < kMaxChains - 1; ++i) {
      VX_CUDA_OK(cudaStreamCreateWithFlags(&s->chain_stream[i], cudaStreamNonBlocking));
      VX_CUDA_OK(cudaEventCreateWithFlags(&s->chain_join[i], cudaEventDisableTiming));
    }
    dev->rt = s;
  }
  *out = dev->rt;
  return 0;
}

static vx_buffer* find_buffer(vx_device* dev, uint64_t addr) {
  for (vx_buffer* b = dev->buffers; b; b = b->next)
    if (addr >= b->reported_addr && addr < b->reported_addr + b->size) return b;
  return nullptr;
}

// Make sure the Node4 image of the scene in `a` is resident; (re)build it when the scene changed.
static int wide_prepare(vx_device* dev, RtState* st, const kernel_arg_t& a) {
  vx_buffer* nb = find_buffer(dev, a.node_addr);
  if (!nb || a.node_addr + (uint64_t)a.num_nodes * sizeof(srt_bvh2_node_t) > nb->reported_addr + nb->size) {
    fprintf(stderr, "[VXDRV] raytrace: node_addr is not inside a device buffer of %u nodes\n", a.num_nodes);
    return -1;
  }
  vx_buffer* tb = find_buffer(dev, a.tri_addr);
  if (!tb || a.tri_addr + (uint64_t)a.num_tris * sizeof(srt_tri_rec_t) > tb->reported_addr + tb->size) {
    fprintf(stderr, "[VXDRV] raytrace: tri_addr is not inside a device buffer of %u triangles\n", a.num_tris);
    return -1;
  }
  WideCache& c = st->wide;
  if (c.valid && c.node_addr == a.node_addr && c.tri_addr == a.tri_addr && c.num_nodes == a.num_nodes &&
      c.num_tris == a.num_tris && c.node_uid == nb->uid && c.tri_uid == tb->uid && c.node_gen == nb->generation &&
      c.tri_gen == tb->generation)
    return 0;
  const srt_bvh2_node_t* d_n2 = reinterpret_cast<const srt_bvh2_node_t*>((const char*)nb->dptr + (a.node_addr - nb->reported_addr));
  const srt_tri_rec_t* d_tr = reinterpret_cast<const srt_tri_rec_t*>((const char*)tb->dptr + (a.tri_addr - tb->reported_addr));
  float4* d_n4 = nullptr;
  uint32_t depth = 0, count4 = 0;
  if (vx_dev_dcr(dev, VX_DCR_RT_REPACK) == 1) {
    // A/B path: read the builder's BVH2 back and repack it on the host (rt_repack.cpp) -- the same bytes
    std::vector<srt_bvh2_node_t> n2(a.num_nodes);
    VX_CUDA_OK(cudaMemcpyAsync(n2.data(), d_n2, n2.size() * sizeof(srt_bvh2_node_t), cudaMemcpyDeviceToHost, dev->stream));
    std::vector<srt_tri_rec_t> tr(a.num_tris);
    VX_CUDA_OK(cudaMemcpyAsync(tr.data(), d_tr, tr.size() * sizeof(srt_tri_rec_t), cudaMemcpyDeviceToHost, dev->stream));
    VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
    std::vector<Node4> n4;
    if (repack_bvh4(n2.data(), a.num_nodes, tr.data(), a.num_tris, n4, &depth) != 0) {
      fprintf(stderr, "[VXDRV] raytrace: malformed BVH2 (child / triangle range out of bounds, or a leaf with more "
                      "than %d triangles)\n", SRT_MAX_LEAF_TRIS);
      return -1;
    }
    count4 = (uint32_t)n4.size();
    VX_CUDA_OK(cudaMalloc(&d_n4, n4.size() * sizeof(Node4)));
    VX_CUDA_OK(cudaMemcpyAsync(d_n4, n4.data(), n4.size() * sizeof(Node4), cudaMemcpyHostToDevice, dev->stream));
    VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
  } else if (repack_bvh4_device(dev->stream, d_n2, a.num_nodes, d_tr, a.num_tris, &d_n4, &count4, &depth,
                                vx_dev_dcr(dev, VX_DCR_RT_REPACK) == 2) != 0) {
    // the repack where the data is: once per scene, level by level (rt_repack_device.cu)
    fprintf(stderr, "[VXDRV] raytrace: malformed BVH2 (child / triangle range out of bounds, or a leaf with more "
                    "than %d triangles)\n", SRT_MAX_LEAF_TRIS);
    return -1;
  }
  if (3u * depth + 1u > (uint32_t)(kSmStack + kLocalStack)) {
    fprintf(stderr, "[VXDRV] raytrace: BVH too deep for the traversal stack (%u wide levels)\n", depth);
    cudaFree(d_n4);
    return -1;
  }
  if (c.d_nodes4) {
    VX_CUDA_OK(cudaStreamSynchronize(dev->stream));  // a queued frame may still read the previous image
    cudaFree(c.d_nodes4);
  }
  c.d_nodes4 = d_n4;
  c.valid = false;
  c.node_addr = a.node_addr; c.tri_addr = a.tri_addr;
  c.node_uid = nb->uid; c.tri_uid = tb->uid; c.node_gen = nb->generation; c.tri_gen = tb->generation;
  c.num_nodes = a.num_nodes; c.num_tris = a.num_tris;
  c.num_nodes4 = count4; c.depth = depth;
  c.valid = true;
  return 0;
}

template <class K>
static int persistent_grid(vx_device* dev, int* cached, K kernel, size_t smem, uint64_t jobs) {
  if (*cached == 0) {
    int n = 0;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kWideThreads, smem) != cudaSuccess || n <= 0)
      return -1;
    *cached = n;
  }
  const uint64_t full = (uint64_t)dev->prop.multiProcessorCount * (uint64_t)*cached;
  const uint64_t need = (jobs + kWideThreads - 1) / kWideThreads;
  return (int)(need < full ? need : full);
}

// grid of a pool kernel (kPoolThreads per CTA, one 64-ray pool per warp) + its spill area
template <class K>
static int pool_grid(vx_device* dev, RtState* st, int* cached, K kernel, size_t smem, uint64_t jobs, int cap) {
  if (*cached == 0) {
    int n = 0;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kPoolThreads, smem) != cudaSuccess || n <= 0)
      return -1;
    *cached = n;
  }
  const int per_sm = cap > 0 && cap < *cached ? cap : *cached;
  const uint64_t full = (uint64_t)dev->prop.multiProcessorCount * (uint64_t)per_sm;
  const uint64_t need = (jobs + kPoolSlots * (kPoolThreads / 32) - 1) / (kPoolSlots * (kPoolThreads / 32));
  const int grid = (int)(need < full ? need : full);
  const size_t warps = (size_t)grid * (kPoolThreads / 32);
  if (warps > st->spill_warps) {
    if (st->d_spill) cudaFree(st->d_spill);
    st->d_spill = nullptr;
    st->spill_warps = 0;
    const size_t w = (size_t)dev->prop.multiProcessorCount * (size_t)*cached * (kPoolThreads / 32);
    if (cudaMalloc(&st->d_spill, w * kPoolSlots * kPoolSpill * sizeof(uint2)) != cudaSuccess) return -1;
    st->spill_warps = w;
  }
  return grid;
}

// vx_mem_free of a scene buffer: the stream-ordered pool hands the same address to the next allocation of that
// size, so the image derived from the freed buffer must not outlive it.
void rt_buffer_freed(vx_device* dev, const vx_buffer* b) {
  if (!dev->rt) return;
  WideCache& c = dev->rt->wide;
  if (c.valid && (c.node_uid == b->uid || c.tri_uid == b->uid)) c.valid = false;
}

// ---- wavefront pipeline (variants 4 / 5, rt_wave.cuh) ------------------------------------------------------------
static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

// queues and path state for `jobs` jobs of one pass; sum_jobs != 0: plus a per-pixel running sum between passes
static int wave_reserve(RtState* st, uint64_t jobs, uint64_t sum_jobs) {
  const bool need_sum = sum_jobs != 0;
  if (st->d_wave && st->wave_jobs >= jobs && (!need_sum || (st->wave.sum && st->wave_sum_jobs >= sum_jobs))) return 0;
  if (st->d_wave) cudaFree(st->d_wave);  // cudaFree waits for queued work that still reads it
  st->d_wave = nullptr;
  st->wave_jobs = 0;
  const size_t n = (size_t)jobs;
  const size_t sz_ctl = 0, sz_hit = align256(n * sizeof(uint2)), sz_path = align256(n * sizeof(float4)),
               sz_q = align256(n * 2 * sizeof(float4)), sz_sum = need_sum ? align256((size_t)sum_jobs * sizeof(float4)) : 0;
  const size_t total = sz_ctl + 2 * sz_hit + sz_path + 2 * sz_q + sz_sum;
  if (cudaMalloc(&st->d_wave, total) != cudaSuccess) {
    cudaGetLastError();
    fprintf(stderr, "[VXDRV] raytrace: out of device memory for the ray queues (%zu MB)\n", total >> 20);
    return -1;
  }
  char* p = st->d_wave;
  st->wave.ctl = st->d_wave_ctl;
  st->wave.job_base = 0;
  st->wave.hit1 = reinterpret_cast<uint2*>(p); p += sz_hit;
  st->wave.hit2 = reinterpret_cast<uint2*>(p); p += sz_hit;
  st->wave.path = reinterpret_cast<float4*>(p); p += sz_path;
  st->wave.q_shadow = reinterpret_cast<float4*>(p); p += sz_q;
  st->wave.q_bounce = reinterpret_cast<float4*>(p); p += sz_q;
  st->wave.sum = need_sum ? reinterpret_cast<float4*>(p) : nullptr;
  st->wave_jobs = jobs;
  st->wave_sum_jobs = sum_jobs;
  return 0;
}

template <int STAGE, int SCHED, bool STATS>
static int wave_trace(vx_device* dev, RtState* st, const kernel_arg_t& a, const WideScene& sc, const WaveBufs& wb,
                      cudaStream_t stream, uint32_t rows, uint32_t jobs, uint32_t plane_jobs, uint32_t s_base) {
  const size_t smem = sizeof(uint2) * kSmStack * kWideThreads;
  int* occ = st->wave_blocks_per_sm + (STAGE * 6 + SCHED * 2 + (STATS ? 1 : 0));
  auto kernel = rt_wave_trace_kernel<STAGE, SCHED, STATS>;
  if (*occ == 0) {
    int n = 0;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kWideThreads, smem) != cudaSuccess || n <= 0)
      return -1;
    *occ = n;
  }
  // camera stage: no more CTAs than there are jobs; queue stages: the queue size is only known on the device
  uint64_t grid = (uint64_t)dev->prop.multiProcessorCount * (uint64_t)*occ;
  if (STAGE == 0) {
    const uint64_t need = ((uint64_t)jobs + kWideThreads - 1) / kWideThreads;
    if (need < grid) grid = need;
  }
  kernel<<<(unsigned)grid, kWideThreads, smem, stream>>>(a, sc, wb, rows, jobs, plane_jobs, s_base, st->d_counters);
  return 0;
}

// One chain of one pass: the five launches of rt_wave.cuh in front of the combine, for the jobs
// [wb.job_base, wb.job_base + jobs) of the pass (job = pass-local sample * plane_jobs + pixel job).
template <int SCHED, bool STATS>
static int wave_chain(vx_device* dev, RtState* st, const kernel_arg_t& a, const WideScene& sc, const WaveBufs& wb,
                      cudaStream_t s, uint32_t rows, uint32_t jobs, uint32_t plane_jobs, uint32_t s_base,
                      uint32_t* launches) {
  const float4* tris = sc.tris;
  const unsigned blocks = (jobs + 255u) / 256u;
  const unsigned sm_blocks = (unsigned)dev->prop.multiProcessorCount * 8u;
  if (wave_trace<0, SCHED, STATS>(dev, st, a, sc, wb, s, rows, jobs, plane_jobs, s_base) != 0) return -1;
  rt_wave_shade1_kernel<<<blocks, 256, 0, s>>>(a, tris, wb, rows, jobs, plane_jobs, s_base);
  *launches += 2;
  const bool fused = a.mode >= SRT_MODE_BOUNCE && vx_dev_dcr(dev, VX_DCR_RT_FUSE_BOUNCE) == 0;
  if (fused) {
    // K3': shadow + bounce queues, bounce hits shaded in-lane and their shadow rays traced by the same lane
    if (wave_trace<3, SCHED, STATS>(dev, st, a, sc, wb, s, rows, jobs, plane_jobs, s_base) != 0) return -1;
    *launches += 1;
  } else {
    if (a.mode >= SRT_MODE_SHADOW) {
      if (wave_trace<1, SCHED, STATS>(dev, st, a, sc, wb, s, rows, jobs, plane_jobs, s_base) != 0) return -1;
      *launches += 1;
    }
    if (a.mode >= SRT_MODE_BOUNCE) {
      rt_wave_shade2_kernel<<<blocks < sm_blocks ? blocks : sm_blocks, 256, 0, s>>>(a, tris, wb);
      if (wave_trace<2, SCHED, STATS>(dev, st, a, sc, wb, s, rows, jobs, plane_jobs, s_base) != 0) return -1;
      *launches += 2;
    }
  }
  return 0;
}

// A frame = passes over sample ranges (one pass whenever pixels x samples fit the queue budget), each pass =
// `chains` independent chains over disjoint job ranges (whole 256-job groups, so that the shading CTAs of different
// chains never share a tile) on their own streams, joined in front of the combine.  Every trace kernel is persistent
// and ends with a tail in which a few long rays keep a few warps busy (ncu: K5 of C3 ran at 30 % warps active);
// with two chains the other chain's next kernel moves into the SM slots a tail frees, so the tails overlap useful
// work instead of idling the GPU.  The chains share hit1 / path (indexed by pass-wide job id) and split the queues;
// each has its own control block.
constexpr uint64_t kWavePassJobs = 1ull << 27;  // 160 B of queues and path state per job: 21 GB at most

template <int SCHED, bool STATS>
static int wave_frame(vx_device* dev, RtState* st, const kernel_arg_t& a, const WideScene& sc, uint32_t rows,
                      uint32_t plane_jobs, uint32_t chains_wanted, uint32_t* launches) {
  cudaStream_t s0 = dev->stream;
  const uint64_t budget = vx_dev_dcr(dev, VX_DCR_RT_PASS_JOBS) ? vx_dev_dcr(dev, VX_DCR_RT_PASS_JOBS) : kWavePassJobs;
  uint32_t per_pass = (uint32_t)(budget / plane_jobs);  // samples per pass
  if (per_pass < 1u) per_pass = 1u;
  if (per_pass > a.spp) per_pass = a.spp;
  if (wave_reserve(st, (uint64_t)plane_jobs * per_pass, per_pass < a.spp ? plane_jobs : 0) != 0) return -1;
  VX_CUDA_OK(cudaMemsetAsync(st->d_wave_ctl, 0, sizeof(WaveCtl) * kMaxChains, s0));
  for (uint32_t s_first = 0; s_first < a.spp; s_first += per_pass) {
    const uint32_t s_count = a.spp - s_first < per_pass ? a.spp - s_first : per_pass;
    const uint32_t jobs = plane_jobs * s_count;
    const uint32_t groups = (jobs + 255u) / 256u;
    // one chain by default: with the bounce stage folded into K3 the launches that had the long tails are gone, and
    // two chains measured equal on one GPU (C3, C4) and 2 % slower with the frame gather behind them (N = 2)
    uint32_t chains = chains_wanted ? chains_wanted : 1u;
    if (chains > (uint32_t)kMaxChains) chains = kMaxChains;
    if (chains > groups) chains = groups;
    if (chains > 1u) {
      VX_CUDA_OK(cudaEventRecord(st->chain_fork, s0));
      for (uint32_t c = 1; c < chains; ++c) VX_CUDA_OK(cudaStreamWaitEvent(st->chain_stream[c - 1], st->chain_fork, 0));
    }
    for (uint32_t c = 0; c < chains; ++c) {
      const uint32_t g0 = (uint32_t)((uint64_t)groups * c / chains), g1 = (uint32_t)((uint64_t)groups * (c + 1) / chains);
      const uint32_t first = g0 * 256u, last = g1 * 256u < jobs ? g1 * 256u : jobs;
      WaveBufs wb = st->wave;
      wb.job_base = first;
      wb.ctl = st->d_wave_ctl + c;
      wb.q_shadow = st->wave.q_shadow + 2 * (size_t)first;  // a chain queues at most one shadow / bounce ray per job
      wb.q_bounce = st->wave.q_bounce + 2 * (size_t)first;
      wb.hit2 = st->wave.hit2 + first;
      if (wave_chain<SCHED, STATS>(dev, st, a, sc, wb, c == 0 ? s0 : st->chain_stream[c - 1], rows, last - first, plane_jobs,
                                   s_first, launches) != 0)
        return -1;
    }
    for (uint32_t c = 1; c < chains; ++c) {
      VX_CUDA_OK(cudaEventRecord(st->chain_join[c - 1], st->chain_stream[c - 1]));
      VX_CUDA_OK(cudaStreamWaitEvent(s0, st->chain_join[c - 1], 0));
    }
    rt_wave_combine_kernel<<<(plane_jobs + 255u) / 256u, 256, 0, s0>>>(a, st->wave, rows, plane_jobs, s_first, s_count,
                                                                      st->d_wave_ctl, chains, st->d_counters);
    *launches += 1;
  }
  return 0;
}

void rt_destroy(vx_device* dev) {
  if (!dev->rt) return;
  cudaFree(dev->rt->d_counters);
  cudaFree(dev->rt->d_wave_ctl);
  cudaEventDestroy(dev->rt->chain_fork);
  for (int i = 0; i < kMaxChains - 1; ++i) {
    cudaStreamDestroy(dev->rt->chain_stream[i]);
    cudaEventDestroy(dev->rt->chain_join[i]);
  }
  if (dev->rt->wide.d_nodes4) cudaFree(dev->rt->wide.d_nodes4);
  if (dev->rt->d_spill) cudaFree(dev->rt->d_spill);
  if (dev->rt->d_samples) cudaFree(dev->rt->d_samples);
  if (dev->rt->d_wave) cudaFree(dev->rt->d_wave);
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

// [addr, addr + bytes) must lie inside one live allocation: a short or stale address is refused here (-1, like
// the other argument checks) instead of faulting inside a kernel, which would poison the CUDA context.
static bool inside_buffer(vx_device* dev, uint64_t addr, uint64_t bytes) {
  const vx_buffer* b = find_buffer(dev, addr);
  return b && addr + bytes <= b->reported_addr + b->size;
}

int rt_launch(vx_device* dev, const kernel_arg_t& a) {
  RtState* st;
  if (validate(a) != 0) {
    fprintf(stderr, "[VXDRV] raytrace: invalid kernel_arg_t\n");
    return -1;
  }
  if (a.mode == SRT_MODE_RAYS) {
    if (a.num_rays && (!inside_buffer(dev, a.ray_addr, a.num_rays * sizeof(srt_ray_t)) ||
                       !inside_buffer(dev, a.hit_addr, a.num_rays * sizeof(srt_hit_t)))) {
      fprintf(stderr, "[VXDRV] raytrace: ray_addr / hit_addr do not hold %llu records\n", (unsigned long long)a.num_rays);
      return -1;
    }
  } else {
    const uint32_t rows_ = srt_rank_rows(a.dst_height, a.rows_per_block, a.num_ranks, a.rank_id);
    if (rows_ && !inside_buffer(dev, a.cbuf_addr, (uint64_t)(rows_ - 1) * a.cbuf_pitch + (uint64_t)a.dst_width * 4)) {
      fprintf(stderr, "[VXDRV] raytrace: cbuf_addr does not hold %u rows of pitch %u\n", rows_, a.cbuf_pitch);
      return -1;
    }
  }
  if (rt_state(dev, &st) != 0) return -1;
  cudaStream_t s = dev->stream;
  const bool stats = vx_dev_counting(dev);
  uint32_t variant = dev->dcrs.count(VX_DCR_RT_VARIANT) ? dev->dcrs[VX_DCR_RT_VARIANT] : 0u;
  // 0 = the measured best per kind of work: frames run the wavefront pipeline with while-while
  // trace kernels (C3: 1.60 ms against 1.92 for the megakernel), explicit ray lists the vote-scheduled persistent kernel
  if (variant == 0u) variant = a.mode == SRT_MODE_RAYS ? 7u : 4u;
  if (variant > 7u) {
    fprintf(stderr, "[VXDRV] raytrace: unknown kernel variant %u\n", variant);
    return -1;
  }
  const bool wide = variant != 1u;
  const bool pooled = variant == 3u;
  const bool wave = variant >= 4u && variant <= 6u;  // wavefront pipeline; trace kernels scheduled 4: while-while, 5: vote, 6: vote + deferred triangle
  const int sched = (variant == 2u || variant == 4u) ? SCHED_WHILE : (variant == 6u ? SCHED_DEFER : SCHED_VOTE);  // 5, 7: vote
  dev->last_launches = 0;
  dev->last_node_bytes = wide ? (uint32_t)sizeof(Node4) : 32u;
  if (wide && wide_prepare(dev, st, a) != 0) return -1;
  uint32_t tune = dev->dcrs.count(VX_DCR_RT_KEEP_ACTIVE) ? dev->dcrs[VX_DCR_RT_KEEP_ACTIVE] : 0u;
  const int pool_cap = (int)((tune >> 16) & 0xffu);  // tuning: CTAs per SM of the pool kernels (0 = all that fit)
  if (pooled) {
    uint32_t tf = tune & 0xffu, tt = (tune >> 8) & 0xffu;
    if (tf == 0u || tf > 64u) tf = kPoolFetchAt;
    if (tt == 0u || tt > 64u) tt = kPoolTriAt;
    tune = tf | (tt << 8);
  } else if (sched == SCHED_WHILE) {
    if (tune == 0u || tune > 32u) tune = kKeepActive;
  } else {
    uint32_t ta = tune & 0xffu, tt = (tune >> 8) & 0xffu;
    if (ta == 0u || ta > 32u) ta = kAdvanceAt;
    if (tt == 0u || tt > 32u) tt = kTriAt;
    tune = ta | (tt << 8);
  }
  const WideScene sc = {st->wide.d_nodes4, reinterpret_cast<const float4*>(a.tri_addr), (int)tune};
  const size_t smem = sizeof(uint2) * kSmStack * kWideThreads;
  int* occ = st->wide_blocks_per_sm + (sched * 4 + (stats ? 2 : 0));
  VX_CUDA_OK(cudaMemsetAsync(st->d_counters, 0, sizeof(unsigned long long) * CTR_COUNT, s));

  uint32_t extra_launches = 0;
  {
    const uint32_t pf = dev->dcrs.count(VX_DCR_RT_PREFETCH) ? dev->dcrs[VX_DCR_RT_PREFETCH] : 0u;
    const size_t nb = (size_t)st->wide.num_nodes4 * sizeof(Node4), tbytes = (size_t)a.num_tris * sizeof(srt_tri_rec_t);
    if (wide && pf == 1u && nb + tbytes <= (size_t)dev->prop.l2CacheSize * 3 / 4) {
      rt_prefetch_l2_kernel<<<dev->prop.multiProcessorCount * 4, 256, 0, s>>>(
          reinterpret_cast<const char*>(st->wide.d_nodes4), nb / 128, reinterpret_cast<const char*>(a.tri_addr),
          (tbytes + 127) / 128);
      extra_launches = 1;
    }
  }

// launch one of the four instantiations of a persistent kernel
#define SRT_LAUNCH_WIDE(KERNEL, OCC, JOBS, ...)                                                    \
  do {                                                                                             \
    int g_;                                                                                        \
    if (sched == SCHED_DEFER && stats) {                                                           \
      if ((g_ = persistent_grid(dev, OCC, KERNEL<SCHED_DEFER, true>, smem, JOBS)) <= 0) return -1; \
      KERNEL<SCHED_DEFER, true><<<g_, kWideThreads, smem, s>>>(__VA_ARGS__);                       \
    } else if (sched == SCHED_DEFER) {                                                             \
      if ((g_ = persistent_grid(dev, OCC, KERNEL<SCHED_DEFER, false>, smem, JOBS)) <= 0) return -1;\
      KERNEL<SCHED_DEFER, false><<<g_, kWideThreads, smem, s>>>(__VA_ARGS__);                      \
    } else if (sched == SCHED_VOTE && stats) {                                                            \
      if ((g_ = persistent_grid(dev, OCC, KERNEL<SCHED_VOTE, true>, smem, JOBS)) <= 0) return -1;  \
      KERNEL<SCHED_VOTE, true><<<g_, kWideThreads, smem, s>>>(__VA_ARGS__);                        \
    } else if (sched == SCHED_VOTE) {                                                              \
      if ((g_ = persistent_grid(dev, OCC, KERNEL<SCHED_VOTE, false>, smem, JOBS)) <= 0) return -1; \
      KERNEL<SCHED_VOTE, false><<<g_, kWideThreads, smem, s>>>(__VA_ARGS__);                       \
    } else if (stats) {                                                                            \
      if ((g_ = persistent_grid(dev, OCC, KERNEL<SCHED_WHILE, true>, smem, JOBS)) <= 0) return -1; \
      KERNEL<SCHED_WHILE, true><<<g_, kWideThreads, smem, s>>>(__VA_ARGS__);                       \
    } else {                                                                                       \
      if ((g_ = persistent_grid(dev, OCC, KERNEL<SCHED_WHILE, false>, smem, JOBS)) <= 0) return -1;\
      KERNEL<SCHED_WHILE, false><<<g_, kWideThreads, smem, s>>>(__VA_ARGS__);                      \
    }                                                                                              \
  } while (0)

  if (a.mode == SRT_MODE_RAYS) {
    dev->last_rays_primary = a.num_rays;
    dev->last_rt_stores = a.num_rays;
    // the job cursor is 32 bits wide: long lists go out in slices of 2^30 rays
    const uint64_t slice = 1ull << 30;
    for (uint64_t first = 0; first < a.num_rays; first += slice) {
      const uint64_t n = a.num_rays - first < slice ? a.num_rays - first : slice;
      if (pooled) {
        const float4* rays = reinterpret_cast<const float4*>(a.ray_addr) + 2 * first;
        float4* hits = reinterpret_cast<float4*>(a.hit_addr) + first;
        if (first) VX_CUDA_OK(cudaMemsetAsync(st->d_counters + CTR_CURSOR, 0, sizeof(unsigned long long), s));
        const size_t psmem = sizeof(WarpPool) * (kPoolThreads / 32);
        int g;
        if (stats) {
          if ((g = pool_grid(dev, st, &st->pool_blocks_per_sm[0], rt_pool_rays_kernel<true>, psmem, n, pool_cap)) <= 0) return -1;
          rt_pool_rays_kernel<true><<<g, kPoolThreads, psmem, s>>>(rays, hits, sc, st->d_spill, (uint32_t)n, st->d_counters);
        } else {
          if ((g = pool_grid(dev, st, &st->pool_blocks_per_sm[1], rt_pool_rays_kernel<false>, psmem, n, pool_cap)) <= 0) return -1;
          rt_pool_rays_kernel<false><<<g, kPoolThreads, psmem, s>>>(rays, hits, sc, st->d_spill, (uint32_t)n, st->d_counters);
        }
      } else if (wide) {
        const float4* rays = reinterpret_cast<const float4*>(a.ray_addr) + 2 * first;
        float4* hits = reinterpret_cast<float4*>(a.hit_addr) + first;
        if (first) VX_CUDA_OK(cudaMemsetAsync(st->d_counters + CTR_CURSOR, 0, sizeof(unsigned long long), s));
        SRT_LAUNCH_WIDE(rt_wide_rays_kernel, occ + 1, n, rays, hits, sc, (uint32_t)n, st->d_counters);
      } else {
        kernel_arg_t b = a;
        b.ray_addr = a.ray_addr + first * sizeof(srt_ray_t);
        b.hit_addr = a.hit_addr + first * sizeof(srt_hit_t);
        b.num_rays = n;
        const uint64_t blocks = (n + 255) / 256;
        if (stats) rt_baseline_rays_kernel<true><<<(unsigned)blocks, 256, 0, s>>>(b, st->d_counters);
        else rt_baseline_rays_kernel<false><<<(unsigned)blocks, 256, 0, s>>>(b, st->d_counters);
      }
      VX_CUDA_OK(cudaGetLastError());
      dev->last_launches += 1;
    }
    dev->last_launches += extra_launches;
    return 0;
  }

  const uint32_t rows = srt_rank_rows(a.dst_height, a.rows_per_block, a.num_ranks, a.rank_id);
  dev->last_rays_primary = (uint64_t)rows * a.dst_width * a.spp;
  dev->last_rt_stores = (uint64_t)rows * a.dst_width;
  if (rows == 0) return 0;
  const uint32_t tiles = ((a.dst_width + 7) / 8) * ((rows + 3) / 4);  // one 8x4 pixel tile = 32 jobs
  if (wave) {
    const uint64_t jobs = (uint64_t)tiles * 32u;
    if (jobs > 0x7FFFFF00ull) return -1;
    uint32_t n = 0;
    int rc;
    const uint32_t chains = vx_dev_dcr(dev, VX_DCR_RT_CHAINS);  // 0 = by pass size
    if (sched == SCHED_WHILE) rc = stats ? wave_frame<SCHED_WHILE, true>(dev, st, a, sc, rows, (uint32_t)jobs, chains, &n)
                                         : wave_frame<SCHED_WHILE, false>(dev, st, a, sc, rows, (uint32_t)jobs, chains, &n);
    else if (sched == SCHED_DEFER) rc = stats ? wave_frame<SCHED_DEFER, true>(dev, st, a, sc, rows, (uint32_t)jobs, chains, &n)
                                              : wave_frame<SCHED_DEFER, false>(dev, st, a, sc, rows, (uint32_t)jobs, chains, &n);
    else rc = stats ? wave_frame<SCHED_VOTE, true>(dev, st, a, sc, rows, (uint32_t)jobs, chains, &n)
                    : wave_frame<SCHED_VOTE, false>(dev, st, a, sc, rows, (uint32_t)jobs, chains, &n);
    if (rc != 0) return -1;
    VX_CUDA_OK(cudaGetLastError());
    dev->last_launches = n + extra_launches;
    return 0;
  }
  if (wide) {
    uint64_t jobs = (uint64_t)tiles * 32u;
    if (jobs > 0xFFFFFF00ull) return -1;
    // Multi-sample frames: one job per (pixel, sample) instead of per pixel, resolved in sample order afterwards
    // (rt_wide.cuh).  VX_DCR_RT_SPLIT_SAMPLES: 0 = when it pays (spp > 1), 1 = never, 2 = always.
    const uint32_t split_dcr = vx_dev_dcr(dev, VX_DCR_RT_SPLIT_SAMPLES);
    float4* samples = nullptr;
    const uint64_t need = (uint64_t)rows * a.dst_width * a.spp;
    if (split_dcr != 1 && (a.spp > 1 || split_dcr == 2) && jobs * a.spp <= 0xFFFFFF00ull && need <= (4ull << 30) / sizeof(float4)) {
      if (st->samples_capacity < need) {
        if (st->d_samples) cudaFree(st->d_samples);
        st->d_samples = nullptr;
        st->samples_capacity = 0;
        if (cudaMalloc(&st->d_samples, need * sizeof(float4)) == cudaSuccess) st->samples_capacity = need;
        else cudaGetLastError();  // no room: per-pixel jobs, same result
      }
      if (st->samples_capacity >= need) {
        samples = st->d_samples;
        jobs *= a.spp;
      }
    }
    if (samples) SRT_LAUNCH_WIDE(rt_wide_render_split_kernel, st->wide_split_blocks_per_sm + (occ - st->wide_blocks_per_sm), jobs, a, sc, rows, (uint32_t)jobs, st->d_counters, samples);
    else SRT_LAUNCH_WIDE(rt_wide_render_kernel, occ, jobs, a, sc, rows, (uint32_t)jobs, st->d_counters, samples);
    if (samples) {
      rt_resolve_kernel<<<dim3((a.dst_width + 63) / 64, (rows + 3) / 4), 256, 0, s>>>(a, rows, samples);
      ++extra_launches;
    }
  } else {
    const uint32_t blocks = (tiles + 7) / 8;  // 8 warps per block
    if (stats) rt_baseline_kernel<true><<<blocks, 256, 0, s>>>(a, rows, st->d_counters);
    else rt_baseline_kernel<false><<<blocks, 256, 0, s>>>(a, rows, st->d_counters);
  }
#undef SRT_LAUNCH_WIDE
  VX_CUDA_OK(cudaGetLastError());
  dev->last_launches = 1 + extra_launches;
  return 0;
}

// read back the device counters of the last run (called after the stream is idle)
int rt_read_counters(vx_device* dev) {
  if (!dev->rt) return -1;
  unsigned long long h[CTR_COUNT];
  VX_CUDA_OK(cudaMemcpy(h, dev->rt->d_counters, sizeof(h), cudaMemcpyDeviceToHost));
  for (int i = 0; i < CTR_COUNT; ++i) dev->last_counters[i] = h[i];
  for (int i = 0; i < kMaxSms; ++i) dev->last_instrs_sm[i] = h[CTR_PER_SM + i];
  dev->last_counters_valid = true;
  return 0;
}

}  // namespace srt

// Test / tooling hook (include/vortex_cuda.h): the Node4 image the last ray-tracing vx_start traversed.
extern "C" int vx_cuda_rt_node4_image(vx_device_h hdevice, void* dst, uint64_t capacity, uint64_t* size) {
  if (nullptr == hdevice || nullptr == size) return -1;
  vx_device* dev = static_cast<vx_device*>(hdevice);
  if (!dev->rt || !dev->rt->wide.valid) return -1;
  const srt::WideCache& c = dev->rt->wide;
  *size = (uint64_t)c.num_nodes4 * sizeof(srt::Node4);
  if (nullptr == dst) return 0;
  if (capacity < *size) return -1;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  VX_CUDA_OK(cudaMemcpyAsync(dst, c.d_nodes4, *size, cudaMemcpyDeviceToHost, dev->stream));
  VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
  return 0;
}
