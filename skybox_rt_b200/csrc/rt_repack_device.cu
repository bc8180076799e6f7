// rt_repack_device.cu -- the builder's BVH2 -> the 128-byte SoA Node4 image the traversal kernels fetch
// (north_star: "the reference's BVH builder output is repacked into a 16B-aligned SoA wide-BVH node ... layout"),
// on the device.  Same result, byte for byte, as the host routine repack_bvh4 (rt_repack.cpp), which stays as the
// A/B path (VX_DCR_RT_REPACK = 1) and as what tests/test_gpu_lbvh.py compares this against.
//
// The host routine is a breadth-first walk: wide node w opens the children of BVH2 node src[w] by largest surface
// area until four entries are held (collapse4), every entry that stays an inner node gets the next free wide-node
// index.  Breadth-first order makes that index "end of this level + number of inner entries of the wide nodes before
// me in this level + my own earlier slots" -- a prefix sum.  So, per level: count (thread / wide node) -> scan ->
// emit (thread / wide node: planes rounded outward in double precision exactly as the host does, refs, and src[] of
// the next level).
// Round 1 ran that as five launches and one 8-byte read-back PER LEVEL (a 1 M-triangle SAH tree has 12-15 levels:
// 548 us of 30+ launches, launch-bound on the narrow top levels).  Now it is ONE cooperative launch
// (repack_all_kernel): a persistent grid walks the levels with two grid-wide barriers each -- count + CTA-local
// scan | barrier | CTA bases from the block sums, emit | barrier -- and the host reads back one 16-byte record at
// the end.  The level-per-launch path stays as VX_DCR_RT_REPACK = 2 for A/B; all three produce the same bytes.
#include <cooperative_groups.h>
#include <math.h>
#include <stdio.h>

#include "rt_repack.h"
#include "scan.cuh"
#include "vx_internal.h"

namespace srt {
namespace {

struct Ent {
  uint32_t id;
  bool tri;
  float area;
};

__device__ __forceinline__ float half_area(const srt_bvh2_node_t& n) {
  const float dx = n.bmax[0] - n.bmin[0], dy = n.bmax[1] - n.bmin[1], dz = n.bmax[2] - n.bmin[2];
  return dx * dy + dy * dz + dz * dx;
}

__device__ __forceinline__ srt_bvh2_node_t load_node(const srt_bvh2_node_t* n2, uint32_t i) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(n2 + i)), b = __ldg(reinterpret_cast<const float4*>(n2 + i) + 1);
  srt_bvh2_node_t n;
  n.bmin[0] = a.x; n.bmin[1] = a.y; n.bmin[2] = a.z; n.left_first = __float_as_uint(a.w);
  n.bmax[0] = b.x; n.bmax[1] = b.y; n.bmax[2] = b.z; n.count = __float_as_uint(b.w);
  return n;
}

// rt_repack.cpp:collapse4, statement for statement
__device__ int collapse4(const srt_bvh2_node_t* n2, uint32_t num_nodes, uint32_t num_tris, uint32_t root, Ent out[4],
                         int& count) {
  count = 0;
  auto add_node = [&](uint32_t n) -> int {
    if (n >= num_nodes || n == 0) return -1;
    const srt_bvh2_node_t c = load_node(n2, n);
    if (c.count > SRT_MAX_LEAF_TRIS || (c.count > 0 && (uint64_t)c.left_first + c.count > num_tris)) return -1;
    out[count++] = Ent{n, false, half_area(c)};
    return 0;
  };
  const srt_bvh2_node_t r = load_node(n2, root);
  if (r.count > 0) {
    if (r.count > SRT_MAX_LEAF_TRIS || (uint64_t)r.left_first + r.count > num_tris) return -1;
    for (uint32_t t = 0; t < r.count; ++t) out[count++] = Ent{r.left_first + t, true, 0.0f};
    return 0;
  }
  if (add_node(r.left_first) || add_node(r.left_first + 1)) return -1;
  for (;;) {
    int best = -1;
    for (int i = 0; i < count; ++i) {
      if (out[i].tri) continue;
      const srt_bvh2_node_t c = load_node(n2, out[i].id);
      const int grow = c.count > 0 ? (int)c.count - 1 : 1;
      if (count + grow > 4) continue;
      if (best < 0 || out[i].area > out[best].area) best = i;
    }
    if (best < 0) break;
    const srt_bvh2_node_t c = load_node(n2, out[best].id);
    if (c.count > 0) {
      out[best] = Ent{c.left_first, true, 0.0f};
      for (uint32_t t = 1; t < c.count; ++t) out[count++] = Ent{c.left_first + t, true, 0.0f};
    } else {
      const uint32_t l = c.left_first;
      if (l + 1 >= num_nodes || l == 0) return -1;
      const srt_bvh2_node_t cl = load_node(n2, l);
      if (cl.count > SRT_MAX_LEAF_TRIS || (cl.count > 0 && (uint64_t)cl.left_first + cl.count > num_tris)) return -1;
      out[best] = Ent{l, false, half_area(cl)};
      if (add_node(l + 1)) return -1;
    }
  }
  return 0;
}

// wide nodes [begin, begin + n) of one level: how many of each one's entries stay inner nodes
__global__ void __launch_bounds__(128) repack_count_kernel(const srt_bvh2_node_t* __restrict__ n2, uint32_t num_nodes,
                                                           uint32_t num_tris, const uint32_t* __restrict__ src,
                                                           uint32_t begin, uint32_t n, uint32_t* __restrict__ inner,
                                                           uint32_t* __restrict__ error) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Ent ch[4];
  int count;
  if (collapse4(n2, num_nodes, num_tris, src[begin + i], ch, count) != 0) {
    atomicOr(error, 1u);
    inner[i] = 0;
    return;
  }
  uint32_t k = 0;
  for (int s = 0; s < count; ++s) k += ch[s].tri ? 0u : 1u;
  inner[i] = k;
}

__device__ __forceinline__ double dmin(double a, double b) { return b < a ? b : a; }  // std::min
__device__ __forceinline__ double dmax(double a, double b) { return a < b ? b : a; }  // std::max

// wide node begin + i: its Node4 record, and src[] of the wide nodes its inner entries become (first one = `next`)
__device__ __forceinline__ void emit_node(const srt_bvh2_node_t* __restrict__ n2, uint32_t num_nodes, uint32_t num_tris,
                                          const srt_tri_rec_t* __restrict__ tris, uint32_t* __restrict__ src,
                                          uint32_t begin, uint32_t i, uint32_t next, uint32_t capacity, double delta,
                                          float4* __restrict__ out) {
  Ent ch[4];
  int count;
  if (collapse4(n2, num_nodes, num_tris, src[begin + i], ch, count) != 0) return;  // flagged by the count pass
  float plane[6][4];
  uint32_t ref[4];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
#pragma unroll
    for (int k = 0; k < 6; ++k) plane[k][s] = (k & 1) ? -__int_as_float(0x7f800000) : __int_as_float(0x7f800000);
    ref[s] = kEmptyRef;
  }
  for (int s = 0; s < count; ++s) {
    double lo[3], hi[3];
    if (ch[s].tri) {
      const float4* tp = reinterpret_cast<const float4*>(tris + ch[s].id);
      const float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
      const float v0[3] = {r0.x, r0.y, r0.z}, e1[3] = {r1.x, r1.y, r1.z}, e2[3] = {r2.x, r2.y, r2.z};
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const double a = (double)v0[k], b = a + (double)e1[k], c = a + (double)e2[k];
        lo[k] = dmin(a, dmin(b, c));
        hi[k] = dmax(a, dmax(b, c));
      }
      ref[s] = kTriBit | ch[s].id;
    } else {
      const srt_bvh2_node_t c = load_node(n2, ch[s].id);
#pragma unroll
      for (int k = 0; k < 3; ++k) { lo[k] = (double)c.bmin[k]; hi[k] = (double)c.bmax[k]; }
      ref[s] = next;
      if (next < capacity) src[next] = ch[s].id;  // beyond capacity = a cycle; the host stops after this level
      ++next;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      // rt_repack.cpp round_down / round_up: the nearest float on the outside of the inflated bound
      plane[2 * k][s] = __double2float_rd(lo[k] - delta);
      plane[2 * k + 1][s] = __double2float_ru(hi[k] + delta);
    }
  }
  float4* o = out + 8 * (size_t)(begin + i);  // Node4 = plane[6][4], ref[4], pad[4]
#pragma unroll
  for (int k = 0; k < 6; ++k) o[k] = make_float4(plane[k][0], plane[k][1], plane[k][2], plane[k][3]);
  o[6] = make_float4(__uint_as_float(ref[0]), __uint_as_float(ref[1]), __uint_as_float(ref[2]), __uint_as_float(ref[3]));
  o[7] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

__global__ void __launch_bounds__(128) repack_emit_kernel(const srt_bvh2_node_t* __restrict__ n2, uint32_t num_nodes,
                                                          uint32_t num_tris, const srt_tri_rec_t* __restrict__ tris,
                                                          uint32_t* __restrict__ src, uint32_t begin, uint32_t n,
                                                          const uint2* __restrict__ offset, uint32_t next_begin,
                                                          uint32_t capacity, double delta, float4* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  emit_node(n2, num_nodes, num_tris, tris, src, begin, i, next_begin + offset[i].x, capacity, delta, out);
}

// ---- the whole repack in one cooperative launch ---------------------------------------------------------------
struct RepackCtl {
  uint32_t end, levels, error, pad;
};

__device__ __forceinline__ uint32_t count_inner(const srt_bvh2_node_t* n2, uint32_t num_nodes, uint32_t num_tris,
                                                uint32_t root, uint32_t* error) {
  Ent ch[4];
  int count;
  if (collapse4(n2, num_nodes, num_tris, root, ch, count) != 0) {
    atomicOr(error, 1u);
    return 0u;
  }
  uint32_t k = 0;
  for (int s = 0; s < count; ++s) k += ch[s].tri ? 0u : 1u;
  return k;
}

// Persistent grid (#SM x resident CTAs).  Every level's wide nodes [begin, end) are dealt to the CTAs in contiguous
// chunks, so that "inner entries of the wide nodes before me" = the sums of the CTAs before mine + a CTA-local scan.
__global__ void __launch_bounds__(kScanThreads)
repack_all_kernel(const srt_bvh2_node_t* __restrict__ n2, uint32_t num_nodes, uint32_t num_tris,
                  const srt_tri_rec_t* __restrict__ tris, uint32_t* __restrict__ src, uint32_t* __restrict__ local_off,
                  uint32_t* __restrict__ block_sum, uint32_t capacity, double delta, float4* __restrict__ out,
                  RepackCtl* __restrict__ ctl) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  const uint32_t nb = gridDim.x, b = blockIdx.x;
  uint32_t begin = 0, end = 1, levels = 0;
  while (begin < end) {
    const uint32_t n = end - begin;
    ++levels;
    const uint32_t chunk = (n + nb - 1) / nb;
    const uint32_t c0 = b * chunk < n ? b * chunk : n, c1 = c0 + chunk < n ? c0 + chunk : n;
    // count + CTA-local exclusive scan (CTA-uniform trip count)
    uint32_t carry = 0;
    for (uint32_t t0 = c0; t0 < c1; t0 += kScanThreads) {
      const uint32_t i = t0 + threadIdx.x;
      const uint32_t k = i < c1 ? count_inner(n2, num_nodes, num_tris, src[begin + i], &ctl->error) : 0u;
      uint2 total;
      const uint2 ex = cta_exclusive(make_uint2(k, 0u), &total);
      if (i < c1) local_off[begin + i] = carry + ex.x;
      carry += total.x;
    }
    if (threadIdx.x == 0) block_sum[b] = carry;
    grid.sync();
    // my CTA's base and the level's total from the block sums
    uint2 part = make_uint2(0u, 0u);
    for (uint32_t j = threadIdx.x; j < nb; j += kScanThreads) {
      const uint32_t v = __ldcg(block_sum + j);
      part.y += v;
      if (j < b) part.x += v;
    }
    uint2 sums;
    cta_exclusive(part, &sums);
    const uint32_t level_total = sums.y;
    const bool bad = __ldcg(&ctl->error) != 0u || (uint64_t)end + level_total > capacity;  // malformed, or a cycle
    if (!bad)
      for (uint32_t i = c0 + threadIdx.x; i < c1; i += kScanThreads)
        emit_node(n2, num_nodes, num_tris, tris, src, begin, i, end + sums.x + local_off[begin + i], capacity, delta, out);
    if (bad) {
      if (b == 0 && threadIdx.x == 0) ctl->error = 1u;
      break;  // every thread of the grid takes this branch together
    }
    grid.sync();  // src[] of the next level is complete, block_sum may be overwritten
    begin = end;
    end += level_total;
  }
  if (b == 0 && threadIdx.x == 0) {
    ctl->end = end;
    ctl->levels = levels;
  }
}

}  // namespace

// d_n2 / d_tris: the app's buffers.  On success *d_out (cudaMalloc'ed, owned by the caller) holds *num_nodes4
// Node4 records and *depth the number of wide levels.  -1: malformed BVH2 (as repack_bvh4) or out of memory.
int repack_bvh4_device(cudaStream_t s, const srt_bvh2_node_t* d_n2, uint32_t num_nodes, const srt_tri_rec_t* d_tris,
                       uint32_t num_tris, float4** d_out, uint32_t* num_nodes4, uint32_t* depth, bool per_level) {
  *d_out = nullptr;
  if (!d_n2 || !d_tris || num_nodes == 0 || num_tris == 0 || num_tris > 0x7FFFFFFFu) return -1;
  srt_bvh2_node_t root;
  VX_CUDA_OK(cudaMemcpyAsync(&root, d_n2, sizeof(root), cudaMemcpyDeviceToHost, s));
  VX_CUDA_OK(cudaStreamSynchronize(s));
  if (root.count > SRT_MAX_LEAF_TRIS || (root.count > 0 && root.count > num_tris)) return -1;
  const double delta = (double)scene_delta_host(root);
  // a wide node is made from a distinct BVH2 node, so there are at most num_nodes of them
  const uint32_t capacity = num_nodes;
  float4* out = nullptr;
  uint32_t *src = nullptr, *inner = nullptr, *error = nullptr;
  uint2 *offset = nullptr, *block_sums = nullptr, *totals = nullptr;
  auto release = [&](bool keep_out) {
    if (!keep_out && out) cudaFree(out);
    cudaFreeAsync(src, s); cudaFreeAsync(inner, s); cudaFreeAsync(error, s);
    cudaFreeAsync(offset, s); cudaFreeAsync(block_sums, s); cudaFreeAsync(totals, s);
  };
  if (cudaMalloc(&out, (size_t)capacity * sizeof(Node4)) != cudaSuccess ||
      cudaMallocAsync(&src, (size_t)capacity * 4, s) != cudaSuccess || cudaMallocAsync(&inner, (size_t)capacity * 4, s) != cudaSuccess ||
      cudaMallocAsync(&offset, (size_t)capacity * 8, s) != cudaSuccess ||
      cudaMallocAsync(&block_sums, ((size_t)capacity / kScanTile + 1) * 8, s) != cudaSuccess ||
      cudaMallocAsync(&totals, 16, s) != cudaSuccess || cudaMallocAsync(&error, 4, s) != cudaSuccess) {
    cudaGetLastError();
    fprintf(stderr, "[VXDRV] raytrace: out of device memory for the Node4 image\n");
    release(false);
    return -1;
  }
  int rc = -1;
  uint32_t begin = 0, end = 1, levels = 0;
  bool launched = false;  // the cooperative kernel ran (its verdict stands); false = it could not be launched
  if (!per_level) {
    // one cooperative launch; `inner` serves as the CTA-local offsets, `offset` as the block sums, `totals` as the
    // control record (16 bytes)
    do {
      int dev_id = 0, per_sm = 0, sms = 0;
      if (cudaGetDevice(&dev_id) != cudaSuccess ||
          cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev_id) != cudaSuccess ||
          cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, repack_all_kernel, kScanThreads, 0) != cudaSuccess ||
          per_sm <= 0)
        break;
      uint64_t grid = (uint64_t)sms * (uint64_t)per_sm;  // all CTAs resident: a cooperative launch requires it
      if (grid > capacity) grid = capacity;              // (also keeps the block sums inside the `offset` allocation)
      if (cudaMemsetAsync(src, 0, 4, s) != cudaSuccess || cudaMemsetAsync(totals, 0, 16, s) != cudaSuccess) break;
      RepackCtl* ctl = reinterpret_cast<RepackCtl*>(totals);
      uint32_t* block_sum = reinterpret_cast<uint32_t*>(offset);  // capacity * 8 bytes >= grid * 4 (grid <= capacity)
      const srt_bvh2_node_t* a_n2 = d_n2;
      const srt_tri_rec_t* a_tris = d_tris;
      uint32_t a_nodes = num_nodes, a_ntris = num_tris, a_cap = capacity;
      double a_delta = delta;
      void* args[] = {&a_n2, &a_nodes, &a_ntris, &a_tris, &src, &inner, &block_sum, &a_cap, &a_delta, &out, &ctl};
      if (cudaLaunchCooperativeKernel((const void*)repack_all_kernel, dim3((unsigned)grid), dim3(kScanThreads), args, 0, s) !=
          cudaSuccess)
        break;
      launched = true;
      RepackCtl h;
      if (cudaMemcpyAsync(&h, ctl, sizeof(h), cudaMemcpyDeviceToHost, s) != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess)
        break;
      if (h.error) break;
      end = h.end;
      levels = h.levels;
      rc = 0;
    } while (0);
    if (!launched) {
      // no cooperative launch on this device / context (MPS share, a partitioned GPU): the per-level schedule
      // produces the same bytes
      cudaGetLastError();
      per_level = true;
    }
  }
  if (per_level && rc != 0 && !launched)
  do {
    if (cudaMemsetAsync(src, 0, 4, s) != cudaSuccess || cudaMemsetAsync(error, 0, 4, s) != cudaSuccess) break;
    bool ok = true;
    while (begin < end) {
      const uint32_t n = end - begin;
      ++levels;
      repack_count_kernel<<<(n + 127) / 128, 128, 0, s>>>(d_n2, num_nodes, num_tris, src, begin, n, inner, error);
      if (scan_u32(s, inner, n, block_sums, offset, totals) != 0) { ok = false; break; }
      repack_emit_kernel<<<(n + 127) / 128, 128, 0, s>>>(d_n2, num_nodes, num_tris, d_tris, src, begin, n, offset, end,
                                                         capacity, delta, out);
      struct { uint2 t; } h;
      uint32_t herr = 0;
      if (cudaMemcpyAsync(&h, totals, sizeof(h), cudaMemcpyDeviceToHost, s) != cudaSuccess ||
          cudaMemcpyAsync(&herr, error, 4, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
          cudaStreamSynchronize(s) != cudaSuccess) { ok = false; break; }
      if (herr || (uint64_t)end + h.t.x > capacity) { ok = false; break; }  // malformed, or more wide nodes than BVH2 nodes: a cycle
      begin = end;
      end += h.t.x;
    }
    if (!ok) break;
    rc = 0;
  } while (0);
  if (rc != 0) cudaGetLastError();
  release(rc == 0);
  if (rc != 0) return -1;
  *d_out = out;
  *num_nodes4 = end;
  *depth = levels;
  return 0;
}

}  // namespace srt
