// raster_binning.cu -- triangle setup + tile binning on the GPU.
//
// Same bytes as the host pre-pass (csrc/host/draw3d_binning.c), which restates the reference's
// graphics::Binning (sim/common/gfxutil.cpp:103-276): primbuf = one 120-byte rast_prim_t per kept
// triangle in submission order, tilebuf = the non-empty tiles in (tile_x, tile_y) order, each with the
// ids of the primitives whose bounding box touches it, ascending.  The reference builds a std::map of
// vectors on the host; here the same result comes from passes that have no order dependence except two
// prefix sums and a per-tile sort:
//
//   setup      thread / triangle   edge equations, attribute planes, tile rectangle, keep flag; records leave
//                                  the CTA through shared memory as one coalesced block
//   scan       keep flags -> primitive ids (order preserving compaction)
//   count      lane / triangle     atomicAdd on a dense (tile_x, tile_y) grid; a triangle over more than 8
//                                  tiles is walked by its whole warp
//   scan       grid counts -> start of every tile's id list, non-empty flags -> tile index
//   (one 28-byte read-back: the sizes of the two outputs are data)
//   compact    thread / word       kept records -> primbuf[pid]
//   headers    thread / grid cell
//   fill       lane / triangle     pid into its tiles' lists (atomic cursor: arrival order)
//   sort       CTA / tile          ascending pid = submission order (bitonic, shared or global memory)
//
// Float expressions keep the host file's operation order; the library is built with -fmad=false and IEEE
// division, and float -> int goes through x86_cvtt so out-of-range values convert as on the host CPU.
#include <stdio.h>
#include <string.h>

#include "raster_tile.cuh"
#include "scan.cuh"
#include "vx_internal.h"

namespace srt {
namespace {

struct BinParams {
  const srt_vertex_t* vertices;
  const srt_primitive_t* primitives;
  uint32_t num_vertices, num_primitives;
  uint32_t width, height, L, tiles_x, tiles_y, rank, num_ranks, format;
  float min_x, scale_x, min_y, scale_y, min_z, scale_z;
};

struct TileRect { uint16_t tx0, tx1, ty0, ty1; };  // [tx0, tx1) x [ty0, ty1)
struct V4 { float x, y, z, w; };

__device__ __forceinline__ int32_t to_fixed(float v, float one) { return x86_cvtt(v * one); }
__device__ __forceinline__ float max_f(float a, float b) { return a < b ? b : a; }
__device__ __forceinline__ float min_f(float a, float b) { return b < a ? b : a; }

__device__ __forceinline__ V4 load_pos8(const srt_vertex_t* v) {  // 40-byte records: 8-byte aligned only
  const float2 a = __ldg(reinterpret_cast<const float2*>(v->pos)), b = __ldg(reinterpret_cast<const float2*>(v->pos) + 1);
  return V4{a.x, a.y, b.x, b.y};
}
__device__ __forceinline__ V4 to_hdc(V4 c, const BinParams& P) {  // math.hpp:1234-1255
  return V4{c.x * P.scale_x + c.w * P.min_x, c.y * P.scale_y + c.w * P.min_y, c.z * P.scale_z + c.w * P.min_z, c.w};
}
__device__ __forceinline__ V4 to_screen(V4 c, const BinParams& P) {  // math.hpp:1222-1230, :1258-1294
  const float rhw = (c.w != 0.0f) ? (1.0f / c.w) : 0.0f;
  return V4{(c.x * rhw) * P.scale_x + P.min_x, (c.y * rhw) * P.scale_y + P.min_y, (c.z * rhw) * P.scale_z + P.min_z, rhw};
}
__device__ __forceinline__ void plane(int32_t* d, float a0, float a1, float a2) {  // ATTRIBUTE_DELTA, gfxutil.cpp:189-192
  d[0] = to_fixed(a0 - a2, 16777216.0f);
  d[1] = to_fixed(a1 - a2, 16777216.0f);
  d[2] = to_fixed(a2, 16777216.0f);
}

// ---- setup: one thread per submitted triangle -------------------------------------------------------
// Each thread builds its record in shared memory (row = thread, 30 words padded to 31: conflict-free for the
// per-thread stores and for the linear read-out); the CTA then writes its 128 records to tmp_prims as one
// contiguous, coalesced block.
constexpr int kSetupThreads = 128, kRecPitch = 31;

__device__ __forceinline__ bool setup_one(const BinParams& P, uint32_t n, int32_t* rec, TileRect* rect, uint32_t* error);

__global__ void __launch_bounds__(kSetupThreads) bin_setup_kernel(const BinParams P, int32_t* __restrict__ tmp_prims,
                                                                  TileRect* __restrict__ rects, uint32_t* __restrict__ keep,
                                                                  uint32_t* __restrict__ error) {
  __shared__ int32_t s_rec[kSetupThreads * kRecPitch];
  const uint32_t first = blockIdx.x * kSetupThreads, n = first + threadIdx.x;
  if (n < P.num_primitives) {
    TileRect r = {0, 0, 0, 0};
    const bool k = setup_one(P, n, s_rec + threadIdx.x * kRecPitch, &r, error);
    keep[n] = k ? 1u : 0u;
    rects[n] = r;
  }
  __syncthreads();
  const uint32_t in_cta = min((uint32_t)kSetupThreads, P.num_primitives - first);
  int32_t* dst = tmp_prims + (size_t)first * 30;
  for (uint32_t w = threadIdx.x; w < in_cta * 30; w += kSetupThreads) {
    const uint32_t t = w / 30u;
    dst[w] = s_rec[t * kRecPitch + (w - t * 30u)];
  }
}

__device__ __forceinline__ bool setup_one(const BinParams& P, uint32_t n, int32_t* rec, TileRect* rect, uint32_t* error) {
  const srt_primitive_t tri = P.primitives[n];
  if (tri.i0 >= P.num_vertices || tri.i1 >= P.num_vertices || tri.i2 >= P.num_vertices) {
    atomicOr(error, 1u);  // the host loader refuses such a trace; never read out of bounds
    return false;
  }
  const srt_vertex_t *v0 = P.vertices + tri.i0, *v1 = P.vertices + tri.i1, *v2 = P.vertices + tri.i2;
  const V4 c0 = load_pos8(v0), c1 = load_pos8(v1), c2 = load_pos8(v2);
  const V4 h0 = to_hdc(c0, P), h1 = to_hdc(c1, P), h2 = to_hdc(c2, P);
  float e[3][3];
  // edge equations in homogeneous 2D (gfxutil.cpp:35-75)
  e[0][0] = (h1.y * h2.w) - (h2.y * h1.w);
  e[1][0] = (h2.y * h0.w) - (h0.y * h2.w);
  e[2][0] = (h0.y * h1.w) - (h1.y * h0.w);
  e[0][1] = (h2.x * h1.w) - (h1.x * h2.w);
  e[1][1] = (h0.x * h2.w) - (h2.x * h0.w);
  e[2][1] = (h1.x * h0.w) - (h0.x * h1.w);
  e[0][2] = (h1.x * h2.y) - (h2.x * h1.y);
  e[1][2] = (h2.x * h0.y) - (h0.x * h2.y);
  e[2][2] = (h0.x * h1.y) - (h1.x * h0.y);
  const float det = e[0][2] * h0.w + e[1][2] * h1.w + e[2][2] * h2.w;
  if (det == 0) return false;  // degenerate, gfxutil.cpp:139-143
  if (det < 0) {
#pragma unroll
    for (int i = 0; i < 3; ++i) { e[i][0] *= -1.0f; e[i][1] *= -1.0f; e[i][2] *= -1.0f; }
  }
  const V4 s0 = to_screen(c0, P), s1 = to_screen(c1, P), s2 = to_screen(c2, P);
  // bounding box clamped to the target (gfxutil.cpp:152-178); compared as uint32_t like rast_bbox_t
  const int l = x86_cvtt(floorf(min_f(s0.x, min_f(s1.x, s2.x)))), r = x86_cvtt(ceilf(max_f(s0.x, max_f(s1.x, s2.x))));
  const int t = x86_cvtt(floorf(min_f(s0.y, min_f(s1.y, s2.y)))), b = x86_cvtt(ceilf(max_f(s0.y, max_f(s1.y, s2.y))));
  const uint32_t left = (uint32_t)(l > 0 ? l : 0), right = (uint32_t)(r < (int)P.width ? r : (int)P.width);
  const uint32_t top = (uint32_t)(t > 0 ? t : 0), bottom = (uint32_t)(b < (int)P.height ? b : (int)P.height);
  if (right <= left || bottom <= top) return false;
#pragma unroll
  for (int i = 0; i < 3; ++i) e[i][2] += e[i][0] * 0.5f + e[i][1] * 0.5f;  // pixel centres, :195-198
  // normalise by the largest |a|, |b| and convert to 16.16 (gfxutil.cpp:79-96)
  float m = fabsf(e[0][0]);
  m = max_f(m, fabsf(e[1][0])); m = max_f(m, fabsf(e[2][0]));
  m = max_f(m, fabsf(e[0][1])); m = max_f(m, fabsf(e[1][1])); m = max_f(m, fabsf(e[2][1]));
  const float scale = 1.0f / m;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    rec[3 * i + 0] = to_fixed(e[i][0] * scale, 65536.0f);
    rec[3 * i + 1] = to_fixed(e[i][1] * scale, 65536.0f);
    rec[3 * i + 2] = to_fixed(e[i][2] * scale, 65536.0f);
  }
  plane(rec + 9, s0.z, s1.z, s2.z);
#pragma unroll
  for (int k = 0; k < 4; ++k) plane(rec + 12 + 3 * k, v0->color[k], v1->color[k], v2->color[k]);
  plane(rec + 24, v0->texcoord[0], v1->texcoord[0], v2->texcoord[0]);
  plane(rec + 27, v0->texcoord[1], v1->texcoord[1], v2->texcoord[1]);
  // tiles touched (gfxutil.cpp:221-235), clamped to the grid exactly as the host pre-pass does
  const uint32_t tile = 1u << P.L;
  uint32_t tx0 = left >> P.L, tx1 = (right + tile - 1) >> P.L, ty0 = top >> P.L, ty1 = (bottom + tile - 1) >> P.L;
  if ((int32_t)right < 0) tx1 = tx0;
  if ((int32_t)bottom < 0) ty1 = ty0;
  if (tx1 > P.tiles_x) tx1 = P.tiles_x;
  if (ty1 > P.tiles_y) ty1 = P.tiles_y;
  if (tx0 > tx1) tx0 = tx1;
  if (ty0 > ty1) ty0 = ty1;
  *rect = TileRect{(uint16_t)tx0, (uint16_t)tx1, (uint16_t)ty0, (uint16_t)ty1};
  return true;
}

// ---- count / fill: one lane per submitted triangle; a triangle touching many tiles is shared by its warp ----
constexpr uint32_t kLaneCells = 8;  // rectangles up to this many tiles are walked by the owning lane alone

// Calls visit(tx, ty, pid) once for every tile of every kept triangle of this warp's 32 triangles.
template <class Visit>
__device__ __forceinline__ void for_each_tile(const BinParams& P, const TileRect* __restrict__ rects,
                                              const uint32_t* __restrict__ keep, const uint2* __restrict__ pid_of, Visit visit) {
  const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
  const bool valid = n < P.num_primitives && keep[n] != 0;
  uint32_t xs = 0, ys = 0, pid = 0, cells = 0;
  if (valid) {
    const uint2 r = *reinterpret_cast<const uint2*>(rects + n);  // (tx0 | tx1 << 16, ty0 | ty1 << 16)
    xs = r.x; ys = r.y;
    pid = pid_of[n].x;
    cells = ((xs >> 16) - (xs & 0xffff)) * ((ys >> 16) - (ys & 0xffff));
  }
  if (valid && cells <= kLaneCells) {
    const uint32_t h = (ys >> 16) - (ys & 0xffff);
    for (uint32_t c = 0; c < cells; ++c) visit((xs & 0xffff) + c / h, (ys & 0xffff) + c % h, pid);
  }
  uint32_t big = __ballot_sync(0xffffffffu, valid && cells > kLaneCells);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t bx = __shfl_sync(0xffffffffu, xs, src), by = __shfl_sync(0xffffffffu, ys, src);
    const uint32_t bpid = __shfl_sync(0xffffffffu, pid, src), bcells = __shfl_sync(0xffffffffu, cells, src);
    const uint32_t h = (by >> 16) - (by & 0xffff);
    for (uint32_t c = lane; c < bcells; c += 32) visit((bx & 0xffff) + c / h, (by & 0xffff) + c % h, bpid);
  }
}

__global__ void __launch_bounds__(256) bin_count_kernel(const BinParams P, const TileRect* __restrict__ rects,
                                                        const uint32_t* __restrict__ keep, const uint2* __restrict__ pid_of,
                                                        uint32_t* __restrict__ count, unsigned long long* __restrict__ pairs) {
  uint32_t mine = 0;
  for_each_tile(P, rects, keep, pid_of, [&](uint32_t tx, uint32_t ty, uint32_t) {
    if (ty % P.num_ranks == P.rank) {
      atomicAdd(count + (size_t)tx * P.tiles_y + ty, 1u);
      ++mine;
    }
  });
  mine = __reduce_add_sync(0xffffffffu, mine);
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(pairs, (unsigned long long)mine);  // 64-bit total: the u32 scan may wrap
}

// ---- compact: kept records -> primbuf[pid]; 8-byte words, four per thread in flight (coalesced both ways) ----
constexpr uint32_t kCompactPerThread = 4;

__global__ void __launch_bounds__(256) bin_compact_kernel(const uint2* __restrict__ tmp_prims, const uint32_t* __restrict__ keep,
                                                          const uint2* __restrict__ pid_of, uint32_t num_words,
                                                          uint2* __restrict__ primbuf) {
  const uint32_t base = blockIdx.x * (256u * kCompactPerThread) + threadIdx.x;  // 15 words per 120-byte record
  uint2 v[kCompactPerThread];
  uint32_t dst[kCompactPerThread];
#pragma unroll
  for (uint32_t k = 0; k < kCompactPerThread; ++k) {
    const uint32_t w = base + k * 256u;
    dst[k] = 0xffffffffu;
    if (w < num_words) {
      const uint32_t n = w / 15u;
      if (keep[n]) {
        dst[k] = pid_of[n].x * 15u + (w - n * 15u);  // < 2^27 * 15
        v[k] = __ldcs(tmp_prims + w);
      }
    }
  }
#pragma unroll
  for (uint32_t k = 0; k < kCompactPerThread; ++k)
    if (dst[k] != 0xffffffffu) primbuf[dst[k]] = v[k];
}

// ---- headers: one thread per grid cell, x-major like std::map<pair<tile_x, tile_y>> (gfxutil.cpp:243-262)
__global__ void __launch_bounds__(256) bin_headers_kernel(const BinParams P, const uint32_t* __restrict__ count,
                                                          const uint2* __restrict__ cursor, uint32_t num_tiles,
                                                          uint8_t* __restrict__ tilebuf, uint32_t* __restrict__ error) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= P.tiles_x * P.tiles_y) return;
  const uint32_t cnt = count[g];
  if (!cnt) return;
  const uint32_t tx = g / P.tiles_y, ty = g % P.tiles_y, t = cursor[g].y, start = cursor[g].x;
  if (P.format == SRT_TILEBUF_WIDE) {
    uint4 h;
    h.x = tx | (ty << 16); h.y = cnt; h.z = start; h.w = 0;
    reinterpret_cast<uint4*>(tilebuf)[t] = h;
  } else {
    // offset counted in u32s from the end of this header (gfxutil.cpp:252)
    const uint64_t off = (uint64_t)(num_tiles - 1 - t) * (sizeof(rast_tile_header_t) / 4) + start;
    if (off > 0xffff || cnt > 0xffff) { atomicOr(error, 2u); return; }
    uint2 h;
    h.x = tx | (ty << 16); h.y = (uint32_t)off | (cnt << 16);
    reinterpret_cast<uint2*>(tilebuf)[t] = h;
  }
}

// ---- fill: pid into its tiles' lists, arrival order inside a tile (sorted afterwards) ------------------------
__global__ void __launch_bounds__(256) bin_fill_kernel(const BinParams P, const TileRect* __restrict__ rects,
                                                       const uint32_t* __restrict__ keep, const uint2* __restrict__ pid_of,
                                                       const uint2* __restrict__ cursor, uint32_t* __restrict__ fill,
                                                       uint32_t* __restrict__ pids) {
  for_each_tile(P, rects, keep, pid_of, [&](uint32_t tx, uint32_t ty, uint32_t pid) {
    if (ty % P.num_ranks != P.rank) return;
    const size_t g = (size_t)tx * P.tiles_y + ty;
    pids[cursor[g].x + atomicAdd(fill + g, 1u)] = pid;
  });
}

// ---- sort: one CTA per grid cell; ascending pid = the reference's submission order ------------------------
// Bitonic network in its all-ascending form (first step of every stage mirrors the upper half), so the
// virtual +inf padding above n never has to move.
constexpr uint32_t kSortShared = 4096;

template <class Ptr>
__device__ __forceinline__ void bitonic_ascending(Ptr a, uint32_t n) {
  uint32_t pow2 = 1;
  while (pow2 < n) pow2 <<= 1;
  for (uint32_t size = 2; size <= pow2; size <<= 1) {
    const uint32_t half = size >> 1;
    for (uint32_t t = threadIdx.x; t < pow2 / 2; t += blockDim.x) {
      // half and stride are powers of two: masks and shifts, no division
      const uint32_t off = t & (half - 1), base = (t & ~(half - 1)) << 1, i = base + off, j = base + size - 1 - off;
      if (j < n) {
        const uint32_t x = a[i], y = a[j];
        if (x > y) { a[i] = y; a[j] = x; }
      }
    }
    __syncthreads();
    for (uint32_t stride = half >> 1; stride > 0; stride >>= 1) {
      for (uint32_t t = threadIdx.x; t < pow2 / 2; t += blockDim.x) {
        const uint32_t i = ((t & ~(stride - 1)) << 1) + (t & (stride - 1)), j = i + stride;
        if (j < n) {
          const uint32_t x = a[i], y = a[j];
          if (x > y) { a[i] = y; a[j] = x; }
        }
      }
      __syncthreads();
    }
  }
}

__global__ void __launch_bounds__(256) bin_sort_kernel(const uint32_t* __restrict__ count, const uint2* __restrict__ cursor,
                                                       uint32_t* __restrict__ pids) {
  __shared__ uint32_t s_list[kSortShared];
  const uint32_t g = blockIdx.x, n = count[g];
  if (n < 2) return;
  uint32_t* list = pids + cursor[g].x;
  {  // arrival order is often already submission order: one pass to find out
    int unsorted = 0;
    for (uint32_t i = threadIdx.x + 1; i < n; i += blockDim.x) unsorted |= list[i - 1] > list[i];
    if (!__syncthreads_or(unsorted)) return;
  }
  if (n <= kSortShared) {
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) s_list[i] = list[i];
    __syncthreads();
    bitonic_ascending(s_list, n);
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) list[i] = s_list[i];
  } else {
    bitonic_ascending((volatile uint32_t*)list, n);  // in place; __syncthreads orders the CTA's global accesses
  }
}

// One stream-ordered allocation per call, carved into 256-byte aligned pieces; the pieces that must start at zero
// come first and are cleared with one memset.
struct Scratch {
  cudaStream_t s;
  uint8_t* base = nullptr;
  size_t used = 0, planned = 0;
  explicit Scratch(cudaStream_t st) : s(st) {}
  ~Scratch() { if (base) cudaFreeAsync(base, s); }
  static size_t pad(size_t bytes) { return (bytes + 255) & ~(size_t)255; }
  template <class T> void plan(size_t count) { planned += pad((count ? count : 1) * sizeof(T)); }
  bool commit() { return cudaMallocAsync((void**)&base, planned, s) == cudaSuccess; }
  template <class T> T* take(size_t count) {
    T* q = reinterpret_cast<T*>(base + used);
    used += pad((count ? count : 1) * sizeof(T));
    return q;
  }
};

}  // namespace
}  // namespace srt

using namespace srt;

extern "C" int vx_cuda_draw3d_binning(vx_device_h hdevice, vx_buffer_h hvertices, uint32_t num_vertices,
                                      vx_buffer_h hprimitives, uint32_t num_primitives, float near_, float far_,
                                      uint32_t width, uint32_t height, uint32_t tile_logsize, uint32_t format,
                                      uint32_t rank, uint32_t num_ranks, vx_buffer_h* htilebuf, vx_buffer_h* hprimbuf,
                                      srt_binning_info_t* info) {
  if (nullptr == hdevice || nullptr == htilebuf || nullptr == hprimbuf || nullptr == info) return -1;
  *htilebuf = nullptr;
  *hprimbuf = nullptr;
  memset(info, 0, sizeof(*info));
  info->format = format;
  if (width == 0 || height == 0 || tile_logsize < 1 || tile_logsize > 8 || width > 32768 || height > 32768 ||
      (format != SRT_TILEBUF_REFERENCE && format != SRT_TILEBUF_WIDE) || num_ranks == 0 || rank >= num_ranks)
    return -1;
  if (num_primitives == 0) return 0;
  if (nullptr == hvertices || nullptr == hprimitives) return -1;
  vx_device* dev = static_cast<vx_device*>(hdevice);
  vx_buffer* vb = static_cast<vx_buffer*>(hvertices);
  vx_buffer* pb = static_cast<vx_buffer*>(hprimitives);
  if (vb->dev != dev || pb->dev != dev || (uint64_t)num_vertices * sizeof(srt_vertex_t) > vb->size ||
      (uint64_t)num_primitives * sizeof(srt_primitive_t) > pb->size || num_primitives > (1u << 27))
    return -1;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  cudaStream_t s = dev->stream;

  BinParams P;
  P.vertices = static_cast<const srt_vertex_t*>(vb->dptr);
  P.primitives = static_cast<const srt_primitive_t*>(pb->dptr);
  P.num_vertices = num_vertices; P.num_primitives = num_primitives;
  P.width = width; P.height = height; P.L = tile_logsize;
  P.tiles_x = (width + (1u << tile_logsize) - 1) >> tile_logsize;
  P.tiles_y = (height + (1u << tile_logsize) - 1) >> tile_logsize;
  P.rank = rank; P.num_ranks = num_ranks; P.format = format;
  // constants of ClipToHDC / NDCToScreen (math.hpp:1243-1250, :1267-1274), computed on the host as the pre-pass does
  P.min_x = (float)(int32_t)width / 2; P.scale_x = (float)(int32_t)width / 2;
  P.min_y = (float)(int32_t)height / 2; P.scale_y = (float)(int32_t)height / 2;
  P.min_z = (near_ + far_) / 2; P.scale_z = (far_ - near_) / 2;
  const uint32_t np = num_primitives, ngrid = P.tiles_x * P.tiles_y;
  const uint32_t scan_blocks = ((np > ngrid ? np : ngrid) + kScanTile - 1) / kScanTile;

  Scratch tmp(s);
  tmp.plan<uint32_t>(ngrid); tmp.plan<uint32_t>(ngrid); tmp.plan<uint2>(2); tmp.plan<uint32_t>(1);
  tmp.plan<unsigned long long>(1);
  const size_t zeroed = tmp.planned;
  tmp.plan<int32_t>((size_t)np * 30); tmp.plan<TileRect>(np); tmp.plan<uint32_t>(np); tmp.plan<uint2>(np);
  tmp.plan<uint2>(ngrid); tmp.plan<uint2>(scan_blocks);
  if (!tmp.commit()) {
    cudaGetLastError();
    fprintf(stderr, "[VXDRV] draw3d binning: out of device memory\n");
    return -1;
  }
  uint32_t* count = tmp.take<uint32_t>(ngrid);
  uint32_t* fill = tmp.take<uint32_t>(ngrid);
  uint2* totals = tmp.take<uint2>(2);   // [0] kept primitives, [1] (pairs, non-empty tiles)
  uint32_t* error = tmp.take<uint32_t>(1);
  unsigned long long* pairs = tmp.take<unsigned long long>(1);
  VX_CUDA_OK(cudaMemsetAsync(tmp.base, 0, zeroed, s));
  int32_t* tmp_prims = tmp.take<int32_t>((size_t)np * 30);
  TileRect* rects = tmp.take<TileRect>(np);
  uint32_t* keep = tmp.take<uint32_t>(np);
  uint2* pid_of = tmp.take<uint2>(np);
  uint2* cursor = tmp.take<uint2>(ngrid);
  uint2* block_sums = tmp.take<uint2>(scan_blocks);
  bin_setup_kernel<<<(np + kSetupThreads - 1) / kSetupThreads, kSetupThreads, 0, s>>>(P, tmp_prims, rects, keep, error);
  if (scan_u32(s, keep, np, block_sums, pid_of, totals + 0) != 0) return -1;
  const uint32_t prim_blocks = (np + 255) / 256;
  bin_count_kernel<<<prim_blocks, 256, 0, s>>>(P, rects, keep, pid_of, count, pairs);
  if (scan_u32(s, count, ngrid, block_sums, cursor, totals + 1) != 0) return -1;
  VX_CUDA_OK(cudaGetLastError());
  // the sizes of the two outputs are data: one small read-back
  struct { uint2 t[2]; } h;
  uint32_t herr = 0;
  unsigned long long hpairs = 0;
  VX_CUDA_OK(cudaMemcpyAsync(&hpairs, pairs, sizeof(hpairs), cudaMemcpyDeviceToHost, s));
  VX_CUDA_OK(cudaMemcpyAsync(&h, totals, sizeof(h), cudaMemcpyDeviceToHost, s));
  VX_CUDA_OK(cudaMemcpyAsync(&herr, error, sizeof(herr), cudaMemcpyDeviceToHost, s));
  VX_CUDA_OK(cudaStreamSynchronize(s));
  if (herr) {
    fprintf(stderr, "[VXDRV] draw3d binning: a primitive indexes past the vertex buffer\n");
    return -1;
  }
  if (hpairs > 0xffffffffull) return -1;  // pid offsets are 32-bit (draw3d_binning.c makes the same check)
  const uint32_t num_prims = h.t[0].x, total = h.t[1].x, num_tiles = h.t[1].y;
  const size_t hdr = format == SRT_TILEBUF_WIDE ? sizeof(srt_tile_header_wide_t) : sizeof(rast_tile_header_t);
  info->num_prims = num_prims;
  info->num_tiles = num_tiles;
  info->num_tile_prims = total;
  info->primbuf_size = (uint64_t)num_prims * sizeof(rast_prim_t);
  info->tilebuf_size = (uint64_t)num_tiles * hdr + (uint64_t)total * 4;
  vx_buffer_h tb = nullptr, prb = nullptr;
  if (num_prims) {
    if (vx_mem_alloc(hdevice, info->primbuf_size, VX_MEM_READ, &prb) != 0) return -1;
    const uint32_t words = np * 15u;  // np <= 2^27
    bin_compact_kernel<<<(words + 256 * kCompactPerThread - 1) / (256 * kCompactPerThread), 256, 0, s>>>(reinterpret_cast<const uint2*>(tmp_prims), keep, pid_of, words,
                                                           static_cast<uint2*>(static_cast<vx_buffer*>(prb)->dptr));
  }
  if (num_tiles) {
    if (vx_mem_alloc(hdevice, info->tilebuf_size, VX_MEM_READ, &tb) != 0) {
      vx_mem_free(prb);
      return -1;
    }
    uint8_t* tilebuf = static_cast<uint8_t*>(static_cast<vx_buffer*>(tb)->dptr);
    uint32_t* pids = reinterpret_cast<uint32_t*>(tilebuf + (size_t)num_tiles * hdr);
    bin_headers_kernel<<<(ngrid + 255) / 256, 256, 0, s>>>(P, count, cursor, num_tiles, tilebuf, error);
    bin_fill_kernel<<<prim_blocks, 256, 0, s>>>(P, rects, keep, pid_of, cursor, fill, pids);
    bin_sort_kernel<<<ngrid, 256, 0, s>>>(count, cursor, pids);
    VX_CUDA_OK(cudaGetLastError());
    if (format == SRT_TILEBUF_REFERENCE) {  // a 16-bit header field may not hold the value: -2, like the host pre-pass
      VX_CUDA_OK(cudaMemcpyAsync(&herr, error, sizeof(herr), cudaMemcpyDeviceToHost, s));
      VX_CUDA_OK(cudaStreamSynchronize(s));
      if (herr & 2u) {
        vx_mem_free(tb);
        vx_mem_free(prb);
        return -2;
      }
    }
  }
  if (tb) static_cast<vx_buffer*>(tb)->tile_logsize = tile_logsize;  // raster_launch refuses sizes its kernel is not built for
  *htilebuf = tb;
  *hprimbuf = prb;
  return 0;
}
