// scan.cuh -- exclusive prefix sum of a u32 array on the device, shared by the pre-passes
// (raster_binning.cu, rt_repack_device.cu).  Three launches: per-CTA reduce, one-CTA spine, per-CTA apply.
// Every element v is scanned as the pair (v, v != 0): out[i].x = sum of in[0..i), out[i].y = number of non-zero
// elements in in[0..i); totals[0] = the grand totals.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "vx_internal.h"

namespace srt {
namespace {

// ---- exclusive scan of (v, v != 0) over a u32 array: reduce / spine / apply ---------------------------
constexpr int kScanThreads = 256, kScanPerThread = 8, kScanTile = kScanThreads * kScanPerThread;

__device__ __forceinline__ uint2 add2(uint2 a, uint2 b) { return make_uint2(a.x + b.x, a.y + b.y); }

// CTA-wide exclusive scan of one uint2 per thread; returns the exclusive prefix, *total = CTA sum
__device__ __forceinline__ uint2 cta_exclusive(uint2 v, uint2* total) {
  __shared__ uint2 s_warp[kScanThreads / 32];
  __shared__ uint2 s_total;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint2 inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t ax = __shfl_up_sync(0xffffffffu, inc.x, o), ay = __shfl_up_sync(0xffffffffu, inc.y, o);
    if (lane >= (uint32_t)o) { inc.x += ax; inc.y += ay; }
  }
  __syncthreads();  // s_warp may still be read from the previous call
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint2 w = lane < kScanThreads / 32 ? s_warp[lane] : make_uint2(0, 0);
    uint2 winc = w;
#pragma unroll
    for (int o = 1; o < kScanThreads / 32; o <<= 1) {
      const uint32_t ax = __shfl_up_sync(0xffffffffu, winc.x, o), ay = __shfl_up_sync(0xffffffffu, winc.y, o);
      if (lane >= (uint32_t)o) { winc.x += ax; winc.y += ay; }
    }
    if (lane < kScanThreads / 32) s_warp[lane] = make_uint2(winc.x - w.x, winc.y - w.y);
    if (lane == kScanThreads / 32 - 1) s_total = winc;
  }
  __syncthreads();
  const uint2 base = s_warp[warp];
  *total = s_total;
  return make_uint2(base.x + inc.x - v.x, base.y + inc.y - v.y);
}

__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const uint32_t* __restrict__ in, uint32_t n,
                                                                    uint2* __restrict__ block_sums) {
  const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * kScanPerThread;
  uint2 s = make_uint2(0, 0);
#pragma unroll
  for (int k = 0; k < kScanPerThread; ++k) {
    const uint32_t v = base + k < n ? in[base + k] : 0u;
    s.x += v; s.y += v != 0u;
  }
  uint2 total;
  cta_exclusive(s, &total);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// one CTA: block_sums -> exclusive prefixes in place, grand total to totals[0]
__global__ void __launch_bounds__(kScanThreads) scan_spine_kernel(uint2* __restrict__ block_sums, uint32_t num_blocks,
                                                                   uint2* __restrict__ totals) {
  uint2 carry = make_uint2(0, 0);
  for (uint32_t base = 0; base < num_blocks; base += kScanThreads) {
    const uint32_t i = base + threadIdx.x;
    const uint2 v = i < num_blocks ? block_sums[i] : make_uint2(0, 0);
    uint2 total;
    const uint2 ex = cta_exclusive(v, &total);
    if (i < num_blocks) block_sums[i] = add2(carry, ex);
    carry = add2(carry, total);
  }
  if (threadIdx.x == 0) totals[0] = carry;
}

__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint32_t* __restrict__ in, uint32_t n,
                                                                   const uint2* __restrict__ block_sums,
                                                                   uint2* __restrict__ out) {
  const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * kScanPerThread;
  uint32_t v[kScanPerThread];
  uint2 s = make_uint2(0, 0);
#pragma unroll
  for (int k = 0; k < kScanPerThread; ++k) {
    v[k] = base + k < n ? in[base + k] : 0u;
    s.x += v[k]; s.y += v[k] != 0u;
  }
  uint2 total;
  uint2 run = add2(block_sums[blockIdx.x], cta_exclusive(s, &total));
#pragma unroll
  for (int k = 0; k < kScanPerThread; ++k) {
    if (base + k < n) out[base + k] = run;
    run.x += v[k]; run.y += v[k] != 0u;
  }
}

int scan_u32(cudaStream_t s, const uint32_t* in, uint32_t n, uint2* block_sums, uint2* out, uint2* totals) {
  const uint32_t blocks = (n + kScanTile - 1) / kScanTile;
  scan_reduce_kernel<<<blocks, kScanThreads, 0, s>>>(in, n, block_sums);
  scan_spine_kernel<<<1, kScanThreads, 0, s>>>(block_sums, blocks, totals);
  scan_apply_kernel<<<blocks, kScanThreads, 0, s>>>(in, n, block_sums, out);
  VX_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace
}  // namespace srt
