// rt_lbvh.cu -- BVH2 build on the device: triangle soup in, the same two flat
// arrays the host builder emits (include/skybox_rt.h: srt_bvh2_node_t[], srt_tri_rec_t[]) out.
//
// The host builder (csrc/host/rt_bvh2.c) is a binned-SAH build: better trees, 0.76 s for 1 M triangles on one
// core.  This one is a linear BVH: 30-bit Morton code of each triangle's box centre, 64-bit keys
// (code << 32 | triangle index, so keys are unique and the tree is a function of the input alone), a bitonic
// sort, the Karras 2012 radix-tree construction (one thread per internal node), a bottom-up refit, and an emit
// pass that turns every subtree of <= SRT_MAX_LEAF_TRIS triangles into one leaf.  Boxes are exact fp32 min / max
// of the contained vertices, as the format requires, so they do not depend on the order threads arrive in.
// What the path renders cannot depend on the tree at all: closest hit = min t with ties to the lowest primitive
// id, on the identical (v0, e1, e2) operands -- tests/test_gpu_lbvh.py holds frames and hit records from both
// builders equal bit for bit.
//
// Node layout of the output: root = node 0; the children of radix-tree node k live at nodes 1 + 2k and 2 + 2k
// (adjacent, as `left_first` demands).  Slots under a subtree that became a leaf are unreachable and zero.
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "vx_internal.h"

namespace srt {
namespace {

struct Box { float lo[3], hi[3]; };

// order-preserving float <-> uint mapping for atomicMin / atomicMax on floats
__device__ __forceinline__ uint32_t f2o(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float o2f(uint32_t o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}
__device__ __forceinline__ float min3(float a, float b, float c) { return fminf(a, fminf(b, c)); }
__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(a, fmaxf(b, c)); }

__device__ __forceinline__ Box tri_box(const srt_triangle_t& t) {
  Box b;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    b.lo[k] = min3(t.v0[k], t.v1[k], t.v2[k]);
    b.hi[k] = max3(t.v0[k], t.v1[k], t.v2[k]);
  }
  return b;
}

__device__ __forceinline__ srt_triangle_t load_tri(const srt_triangle_t* soup, uint32_t i) {
  // 40-byte records: five 8-byte loads
  const uint2* p = reinterpret_cast<const uint2*>(soup + i);
  uint2 w[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) w[k] = __ldg(p + k);
  srt_triangle_t t;
  memcpy(&t, w, sizeof(t));
  return t;
}

// ---- scene bounds of the box centres' boxes (= of all vertices) ------------------------------------------
__global__ void __launch_bounds__(256) lbvh_bounds_kernel(const srt_triangle_t* __restrict__ soup, uint32_t n,
                                                          uint32_t* __restrict__ scene /* lo[3], hi[3] ordered */) {
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const Box b = tri_box(load_tri(soup, i));
#pragma unroll
    for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], b.lo[k]); hi[k] = fmaxf(hi[k], b.hi[k]); }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    uint32_t l = f2o(lo[k]), h = f2o(hi[k]);
    l = __reduce_min_sync(0xffffffffu, l);
    h = __reduce_max_sync(0xffffffffu, h);
    if ((threadIdx.x & 31) == 0) { atomicMin(scene + k, l); atomicMax(scene + 3 + k, h); }
  }
}

__device__ __forceinline__ uint32_t spread10(uint32_t v) {  // 10 bits -> every third bit
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

__global__ void __launch_bounds__(256) lbvh_keys_kernel(const srt_triangle_t* __restrict__ soup, uint32_t n, uint32_t npow2,
                                                        const uint32_t* __restrict__ scene, unsigned long long* __restrict__ keys) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npow2) return;
  if (i >= n) { keys[i] = ~0ull; return; }  // padding sorts to the end and never moves
  const Box b = tri_box(load_tri(soup, i));
  // one scale for the three axes (the largest extent): Morton cells are cubes, so a flat scene spends its bits on
  // the two axes it extends along instead of slicing itself into thin layers
  float ext = 0.0f;
#pragma unroll
  for (int k = 0; k < 3; ++k) ext = fmaxf(ext, o2f(scene[3 + k]) - o2f(scene[k]));
  uint32_t q[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float lo = o2f(scene[k]), c = (b.lo[k] + b.hi[k]) * 0.5f;
    float g = ext > 0.0f ? ((c - lo) / ext) * 1024.0f : 0.0f;
    g = fminf(fmaxf(g, 0.0f), 1023.0f);  // NaN -> 0
    q[k] = (uint32_t)g;
  }
  const uint32_t code = (spread10(q[0]) << 2) | (spread10(q[1]) << 1) | spread10(q[2]);
  keys[i] = ((unsigned long long)code << 32) | i;
}

// ---- bitonic sort of npow2 u64 keys, all-ascending form ------------------------------------------------------
constexpr uint32_t kChunk = 2048;  // keys per CTA in the shared-memory stages (16 KB)

__device__ __forceinline__ void cmpx(unsigned long long& a, unsigned long long& b) {
  if (a > b) { const unsigned long long t = a; a = b; b = t; }
}

// strides `from` .. 1 of the half-cleaner on a chunk held in shared memory
__device__ __forceinline__ void local_halves(unsigned long long* s, uint32_t from) {
  for (uint32_t stride = from; stride > 0; stride >>= 1) {
    for (uint32_t t = threadIdx.x; t < kChunk / 2; t += blockDim.x) {
      const uint32_t i = ((t & ~(stride - 1)) << 1) + (t & (stride - 1));  // stride is a power of two
      cmpx(s[i], s[i + stride]);
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(512) sort_local_kernel(unsigned long long* __restrict__ keys) {
  __shared__ unsigned long long s[kChunk];
  unsigned long long* g = keys + (size_t)blockIdx.x * kChunk;
  for (uint32_t t = threadIdx.x; t < kChunk; t += blockDim.x) s[t] = g[t];
  __syncthreads();
  for (uint32_t size = 2; size <= kChunk; size <<= 1) {
    const uint32_t half = size >> 1;
    for (uint32_t t = threadIdx.x; t < kChunk / 2; t += blockDim.x) {
      const uint32_t off = t & (half - 1), base = (t & ~(half - 1)) << 1;
      cmpx(s[base + off], s[base + size - 1 - off]);
    }
    __syncthreads();
    local_halves(s, half >> 1);
  }
  for (uint32_t t = threadIdx.x; t < kChunk; t += blockDim.x) g[t] = s[t];
}

__global__ void __launch_bounds__(256) sort_flip_kernel(unsigned long long* __restrict__ keys, uint32_t size, uint32_t pairs) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= pairs) return;
  const uint32_t half = size >> 1, off = t & (half - 1);
  const size_t base = (size_t)(t & ~(half - 1)) << 1, i = base + off, j = base + size - 1 - off;
  unsigned long long a = keys[i], b = keys[j];
  if (a > b) { keys[i] = b; keys[j] = a; }
}

__global__ void __launch_bounds__(256) sort_half_kernel(unsigned long long* __restrict__ keys, uint32_t stride, uint32_t pairs) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= pairs) return;
  const size_t i = ((size_t)(t & ~(stride - 1)) << 1) + (t & (stride - 1));
  unsigned long long a = keys[i], b = keys[i + stride];
  if (a > b) { keys[i] = b; keys[i + stride] = a; }
}

// Two global steps per pass over the array (four keys per thread): the flip of stage `size` together with its
// first half-cleaner (stride size / 4) ...
__global__ void __launch_bounds__(256) sort_flip_half_kernel(unsigned long long* __restrict__ keys, uint32_t size, uint32_t quads) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= quads) return;
  const uint32_t quarter = size >> 2, off = t & (quarter - 1);
  const size_t base = (size_t)(t & ~(quarter - 1)) << 2;  // start of this thread's group of `size` keys
  const size_t e0 = base + off, e1 = e0 + quarter, m0 = base + size - 1 - off, m1 = m0 - quarter;
  unsigned long long a = keys[e0], b = keys[e1], c = keys[m1], d = keys[m0];
  cmpx(a, d); cmpx(b, c);  // flip: i against its mirror
  cmpx(a, b); cmpx(c, d);  // stride size / 4 inside each half
  keys[e0] = a; keys[e1] = b; keys[m1] = c; keys[m0] = d;
}

// ... and two consecutive half-cleaners (strides `stride` and `stride / 2`).
__global__ void __launch_bounds__(256) sort_half2_kernel(unsigned long long* __restrict__ keys, uint32_t stride, uint32_t quads) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= quads) return;
  const uint32_t h = stride >> 1, off = t & (h - 1);
  const size_t i0 = ((size_t)(t & ~(h - 1)) << 2) + off, i1 = i0 + h, i2 = i0 + stride, i3 = i2 + h;
  unsigned long long a = keys[i0], b = keys[i1], c = keys[i2], d = keys[i3];
  cmpx(a, c); cmpx(b, d);
  cmpx(a, b); cmpx(c, d);
  keys[i0] = a; keys[i1] = b; keys[i2] = c; keys[i3] = d;
}

__global__ void __launch_bounds__(512) sort_local_halves_kernel(unsigned long long* __restrict__ keys) {
  __shared__ unsigned long long s[kChunk];
  unsigned long long* g = keys + (size_t)blockIdx.x * kChunk;
  for (uint32_t t = threadIdx.x; t < kChunk; t += blockDim.x) s[t] = g[t];
  __syncthreads();
  local_halves(s, kChunk / 2);
  for (uint32_t t = threadIdx.x; t < kChunk; t += blockDim.x) g[t] = s[t];
}

int sort_keys(cudaStream_t st, unsigned long long* keys, uint32_t npow2) {
  if (npow2 < kChunk) return -1;  // the caller pads to at least one chunk
  const uint32_t chunks = npow2 / kChunk, pairs = npow2 / 2;
  sort_local_kernel<<<chunks, 512, 0, st>>>(keys);
  const uint32_t quads = npow2 / 4;
  for (uint32_t size = 2 * kChunk; size != 0 && size <= npow2; size <<= 1) {
    // steps of this stage that reach across chunks: the flip, then strides size/4 ... kChunk; two per launch
    uint32_t stride = size >> 2;
    if (stride >= kChunk) {
      sort_flip_half_kernel<<<(quads + 255) / 256, 256, 0, st>>>(keys, size, quads);
      stride >>= 1;
    } else {
      sort_flip_kernel<<<(pairs + 255) / 256, 256, 0, st>>>(keys, size, pairs);
    }
    while (stride >= kChunk) {
      if ((stride >> 1) >= kChunk) {
        sort_half2_kernel<<<(quads + 255) / 256, 256, 0, st>>>(keys, stride, quads);
        stride >>= 2;
      } else {
        sort_half_kernel<<<(pairs + 255) / 256, 256, 0, st>>>(keys, stride, pairs);
        stride >>= 1;
      }
    }
    sort_local_halves_kernel<<<chunks, 512, 0, st>>>(keys);
  }
  VX_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---- leaves: triangle records in sorted order + their boxes ---------------------------------------------------
__global__ void __launch_bounds__(256) lbvh_leaves_kernel(const srt_triangle_t* __restrict__ soup, uint32_t n,
                                                          const unsigned long long* __restrict__ keys,
                                                          float4* __restrict__ recs, Box* __restrict__ leaf_box) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t id = (uint32_t)keys[i];
  const srt_triangle_t t = load_tri(soup, id);
  // e1 = v1 - v0, e2 = v2 - v0: one fp32 subtraction per component (include/skybox_rt.h, rt_bvh2.c)
  recs[3 * (size_t)i + 0] = make_float4(t.v0[0], t.v0[1], t.v0[2], __uint_as_float(id));
  recs[3 * (size_t)i + 1] = make_float4(t.v1[0] - t.v0[0], t.v1[1] - t.v0[1], t.v1[2] - t.v0[2], __uint_as_float(t.albedo));
  recs[3 * (size_t)i + 2] = make_float4(t.v2[0] - t.v0[0], t.v2[1] - t.v0[1], t.v2[2] - t.v0[2], __uint_as_float(0u));
  leaf_box[i] = tri_box(t);
}

// ---- Karras 2012: one thread per internal node of the radix tree over the sorted unique keys ----------------
constexpr uint32_t kLeafFlag = 0x80000000u;

__device__ __forceinline__ int delta(const unsigned long long* __restrict__ keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  return __clzll((long long)(keys[i] ^ keys[j]));
}

__global__ void __launch_bounds__(256) lbvh_hierarchy_kernel(const unsigned long long* __restrict__ keys, uint32_t n,
                                                             uint2* __restrict__ children, uint2* __restrict__ range,
                                                             uint32_t* __restrict__ parent_internal,
                                                             uint32_t* __restrict__ parent_leaf, uint32_t* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, N = (int)n;
  if (i >= N - 1) return;
  const int d = delta(keys, N, i, i + 1) - delta(keys, N, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = delta(keys, N, i, i - d);
  int lmax = 2;
  while (delta(keys, N, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (delta(keys, N, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = delta(keys, N, i, j);
  int s = 0, t = l;
  do {
    t = (t + 1) >> 1;
    if (delta(keys, N, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  const int gamma = i + s * d + min(d, 0);
  const int first = min(i, j), last = max(i, j);
  const uint32_t left = first == gamma ? ((uint32_t)gamma | kLeafFlag) : (uint32_t)gamma;
  const uint32_t right = last == gamma + 1 ? ((uint32_t)(gamma + 1) | kLeafFlag) : (uint32_t)(gamma + 1);
  children[i] = make_uint2(left, right);
  range[i] = make_uint2((uint32_t)first, (uint32_t)last);
  count[i] = (uint32_t)(last - first + 1);
  if (left & kLeafFlag) parent_leaf[gamma] = (uint32_t)i; else parent_internal[gamma] = (uint32_t)i;
  if (right & kLeafFlag) parent_leaf[gamma + 1] = (uint32_t)i; else parent_internal[gamma + 1] = (uint32_t)i;
  if (i == 0) parent_internal[0] = 0xffffffffu;
}

// ---- refit: leaves walk up, the second child to arrive computes the parent's box -----------------------------
__global__ void __launch_bounds__(256) lbvh_refit_kernel(uint32_t n, const uint2* __restrict__ children,
                                                         const uint32_t* __restrict__ parent_internal,
                                                         const uint32_t* __restrict__ parent_leaf,
                                                         const Box* __restrict__ leaf_box, Box* __restrict__ node_box,
                                                         uint32_t* __restrict__ arrived) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t p = parent_leaf[i];
  while (p != 0xffffffffu) {
    __threadfence();  // this thread's box stores before the flag
    if (atomicAdd(arrived + p, 1u) == 0u) return;
    __threadfence();  // the sibling's box stores after the flag
    const uint2 c = children[p];
    const volatile Box* a = (c.x & kLeafFlag) ? leaf_box + (c.x & ~kLeafFlag) : node_box + c.x;
    const volatile Box* b = (c.y & kLeafFlag) ? leaf_box + (c.y & ~kLeafFlag) : node_box + c.y;
    Box u;
#pragma unroll
    for (int k = 0; k < 3; ++k) { u.lo[k] = fminf(a->lo[k], b->lo[k]); u.hi[k] = fmaxf(a->hi[k], b->hi[k]); }
    node_box[p] = u;
    p = parent_internal[p];
  }
}

// ---- tree rotations: a bottom-up pass that lowers the SAH cost of the radix tree -----------------------------------
// The radix tree splits where the Morton codes first differ -- spatial medians, ~8 % slower to trace than the host's
// binned-SAH tree (C3).  At every internal node p = (L, R), once both subtrees are final (same arrival counters as
// the refit), the four rotations that swap a child with a grandchild on the other side are priced by the one term
// of the SAH cost they change, the surface area of the restructured child: R <-> L1 makes L = (R, L2), R <-> L2 makes
// L = (L1, R), L <-> R1 makes R = (L, R2), L <-> R2 makes R = (R1, L).  The cheapest one is applied if it beats the
// current tree (Kensler 2008 / Kopta et al. 2012).  Subtrees of <= SRT_MAX_LEAF_TRIS triangles become leaves over
// a CONTIGUOUS run of sorted triangles in the emit, so a node is only restructured while it holds more than that
// before and after; smaller subtrees move as a whole.  Deterministic: a node's decision depends only on its
// finished subtree.  range[] is stale for restructured nodes afterwards; count[] is kept right and is what the
// emit uses.
__device__ __forceinline__ Box load_box(const Box* leaf_box, const Box* node_box, uint32_t ref) {
  const volatile Box* b = (ref & kLeafFlag) ? leaf_box + (ref & ~kLeafFlag) : node_box + ref;
  Box r;
#pragma unroll
  for (int k = 0; k < 3; ++k) { r.lo[k] = b->lo[k]; r.hi[k] = b->hi[k]; }
  return r;
}
__device__ __forceinline__ Box unite(const Box& a, const Box& b) {
  Box u;
#pragma unroll
  for (int k = 0; k < 3; ++k) { u.lo[k] = fminf(a.lo[k], b.lo[k]); u.hi[k] = fmaxf(a.hi[k], b.hi[k]); }
  return u;
}
__device__ __forceinline__ float half_area(const Box& b) {
  const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
  return dx * dy + dy * dz + dz * dx;
}

__global__ void __launch_bounds__(256) lbvh_rotate_kernel(uint32_t n, uint2* children, uint32_t* parent_internal,
                                                          uint32_t* parent_leaf, const Box* __restrict__ leaf_box,
                                                          Box* node_box, uint32_t* count, uint32_t* __restrict__ arrived) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  volatile uint2* vchildren = children;
  volatile uint32_t* vcount = count;
  auto cnt = [&](uint32_t ref) -> uint32_t { return (ref & kLeafFlag) ? 1u : vcount[ref]; };
  auto set_parent = [&](uint32_t ref, uint32_t q) {
    if (ref & kLeafFlag) parent_leaf[ref & ~kLeafFlag] = q; else parent_internal[ref] = q;
  };
  uint32_t p = parent_leaf[i];
  while (p != 0xffffffffu) {
    __threadfence();  // this thread's stores before the flag
    if (atomicAdd(arrived + p, 1u) == 0u) return;
    __threadfence();  // the sibling subtree's stores after the flag
    const uint32_t L = vchildren[p].x, R = vchildren[p].y;
    const Box bl = load_box(leaf_box, node_box, L), br = load_box(leaf_box, node_box, R);
    const uint32_t nl = cnt(L), nr = cnt(R);
    float best = 0.0f;
    int pick = 0;
    Box nb = bl;
    uint32_t l1 = 0, l2 = 0, r1 = 0, r2 = 0;
    if (!(L & kLeafFlag) && nl > SRT_MAX_LEAF_TRIS) {
      l1 = vchildren[L].x; l2 = vchildren[L].y;
      const float old = half_area(bl);
      const Box a = unite(br, load_box(leaf_box, node_box, l2));  // R <-> L1
      if (nr + cnt(l2) > SRT_MAX_LEAF_TRIS && old - half_area(a) > best) { best = old - half_area(a); pick = 1; nb = a; }
      const Box b = unite(load_box(leaf_box, node_box, l1), br);  // R <-> L2
      if (cnt(l1) + nr > SRT_MAX_LEAF_TRIS && old - half_area(b) > best) { best = old - half_area(b); pick = 2; nb = b; }
    }
    if (!(R & kLeafFlag) && nr > SRT_MAX_LEAF_TRIS) {
      r1 = vchildren[R].x; r2 = vchildren[R].y;
      const float old = half_area(br);
      const Box a = unite(bl, load_box(leaf_box, node_box, r2));  // L <-> R1
      if (nl + cnt(r2) > SRT_MAX_LEAF_TRIS && old - half_area(a) > best) { best = old - half_area(a); pick = 3; nb = a; }
      const Box b = unite(load_box(leaf_box, node_box, r1), bl);  // L <-> R2
      if (cnt(r1) + nl > SRT_MAX_LEAF_TRIS && old - half_area(b) > best) { best = old - half_area(b); pick = 4; nb = b; }
    }
    if (pick == 1) {
      children[L] = make_uint2(R, l2); children[p] = make_uint2(L, l1);
      count[L] = nr + cnt(l2); node_box[L] = nb; set_parent(R, L); set_parent(l1, p);
    } else if (pick == 2) {
      children[L] = make_uint2(l1, R); children[p] = make_uint2(L, l2);
      count[L] = cnt(l1) + nr; node_box[L] = nb; set_parent(R, L); set_parent(l2, p);
    } else if (pick == 3) {
      children[R] = make_uint2(L, r2); children[p] = make_uint2(r1, R);
      count[R] = nl + cnt(r2); node_box[R] = nb; set_parent(L, R); set_parent(r1, p);
    } else if (pick == 4) {
      children[R] = make_uint2(r1, L); children[p] = make_uint2(r2, R);
      count[R] = cnt(r1) + nl; node_box[R] = nb; set_parent(L, R); set_parent(r2, p);
    }
    p = parent_internal[p];
  }
}

// ---- emit: the flat BVH2.  Children of radix node k at nodes 1 + 2k, 2 + 2k; subtrees of <= 4 triangles = leaves.
__device__ __forceinline__ void put_node(srt_bvh2_node_t* out, const Box& b, uint32_t left_first, uint32_t count) {
  float4* o = reinterpret_cast<float4*>(out);
  o[0] = make_float4(b.lo[0], b.lo[1], b.lo[2], __uint_as_float(left_first));
  o[1] = make_float4(b.hi[0], b.hi[1], b.hi[2], __uint_as_float(count));
}

__global__ void __launch_bounds__(256) lbvh_emit_kernel(uint32_t n, const uint2* __restrict__ children,
                                                        const uint2* __restrict__ range, const uint32_t* __restrict__ count,
                                                        const Box* __restrict__ leaf_box,
                                                        const Box* __restrict__ node_box, srt_bvh2_node_t* __restrict__ nodes) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n - 1) return;
  const uint32_t size = count[k];  // range[] is only trusted for subtrees that become leaves (never restructured)
  if (k == 0) {
    if (size <= SRT_MAX_LEAF_TRIS) put_node(nodes, node_box[0], 0u, size);
    else put_node(nodes, node_box[0], 1u, 0u);
  }
  if (size <= SRT_MAX_LEAF_TRIS) return;  // this subtree is one leaf of its parent
  const uint2 c = children[k];
  const uint32_t ch[2] = {c.x, c.y};
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    srt_bvh2_node_t* out = nodes + 1 + 2 * (size_t)k + s;
    if (ch[s] & kLeafFlag) {
      const uint32_t leaf = ch[s] & ~kLeafFlag;
      put_node(out, leaf_box[leaf], leaf, 1u);
    } else {
      const uint32_t csize = count[ch[s]];
      if (csize <= SRT_MAX_LEAF_TRIS) put_node(out, node_box[ch[s]], range[ch[s]].x, csize);
      else put_node(out, node_box[ch[s]], 1u + 2u * ch[s], 0u);
    }
  }
}

struct Scratch {  // one stream-ordered allocation carved into 256-byte aligned pieces
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

extern "C" int vx_cuda_rt_build_lbvh(vx_device_h hdevice, vx_buffer_h hsoup, uint32_t num_triangles, vx_buffer_h* hnodes,
                                     vx_buffer_h* htris, srt_lbvh_info_t* info) {
  if (nullptr == hdevice || nullptr == hsoup || nullptr == hnodes || nullptr == htris || nullptr == info) return -1;
  *hnodes = *htris = nullptr;
  memset(info, 0, sizeof(*info));
  vx_device* dev = static_cast<vx_device*>(hdevice);
  vx_buffer* sb = static_cast<vx_buffer*>(hsoup);
  const uint32_t n = num_triangles;
  if (n == 0 || n > (1u << 27) || sb->dev != dev || (uint64_t)n * sizeof(srt_triangle_t) > sb->size) return -1;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  cudaStream_t s = dev->stream;
  const srt_triangle_t* soup = static_cast<const srt_triangle_t*>(sb->dptr);
  uint32_t npow2 = kChunk;
  while (npow2 < n) npow2 <<= 1;
  const uint32_t num_nodes = n == 1 ? 1u : 2u * n - 1u;

  vx_buffer_h nb = nullptr, tb = nullptr;
  if (vx_mem_alloc(hdevice, (uint64_t)num_nodes * sizeof(srt_bvh2_node_t), VX_MEM_READ, &nb) != 0) return -1;
  if (vx_mem_alloc(hdevice, (uint64_t)n * sizeof(srt_tri_rec_t), VX_MEM_READ, &tb) != 0) { vx_mem_free(nb); return -1; }
  srt_bvh2_node_t* nodes = static_cast<srt_bvh2_node_t*>(static_cast<vx_buffer*>(nb)->dptr);
  float4* recs = static_cast<float4*>(static_cast<vx_buffer*>(tb)->dptr);

  Scratch tmp(s);
  tmp.plan<uint32_t>(8); tmp.plan<uint32_t>(n);
  const size_t head = tmp.planned;  // scene bounds + arrival flags: initialised below
  tmp.plan<unsigned long long>(npow2); tmp.plan<Box>(n); tmp.plan<Box>(n); tmp.plan<uint2>(n); tmp.plan<uint2>(n);
  tmp.plan<uint32_t>(n); tmp.plan<uint32_t>(n); tmp.plan<uint32_t>(n);
  if (!tmp.commit()) {
    cudaGetLastError();
    fprintf(stderr, "[VXDRV] lbvh build: out of device memory\n");
    vx_mem_free(nb); vx_mem_free(tb);
    return -1;
  }
  uint32_t* scene = tmp.take<uint32_t>(8);
  uint32_t* arrived = tmp.take<uint32_t>(n);
  unsigned long long* keys = tmp.take<unsigned long long>(npow2);
  Box* leaf_box = tmp.take<Box>(n);
  Box* node_box = tmp.take<Box>(n);
  uint2* children = tmp.take<uint2>(n);
  uint2* range = tmp.take<uint2>(n);
  uint32_t* parent_internal = tmp.take<uint32_t>(n);
  uint32_t* parent_leaf = tmp.take<uint32_t>(n);
  uint32_t* count = tmp.take<uint32_t>(n);
  // VX_DCR_RT_LBVH_ROTATE: 0 = default (2 rotation passes), k > 0 = k - 1 passes (1 = the plain radix tree)
  const uint32_t rot_dcr = vx_dev_dcr(dev, VX_DCR_RT_LBVH_ROTATE);
  const uint32_t rotate_passes = rot_dcr == 0 ? 2u : (rot_dcr - 1u > 8u ? 8u : rot_dcr - 1u);
  int rc = -1;
  do {
    if (cudaMemsetAsync(tmp.base, 0, head, s) != cudaSuccess) break;
    if (cudaMemsetAsync(scene, 0xff, 3 * sizeof(uint32_t), s) != cudaSuccess) break;  // lo = +max in the ordered encoding
    if (cudaMemsetAsync(nodes, 0, (size_t)num_nodes * sizeof(srt_bvh2_node_t), s) != cudaSuccess) break;
    const uint32_t blocks = (n + 255) / 256;
    lbvh_bounds_kernel<<<blocks < 1184 ? blocks : 1184, 256, 0, s>>>(soup, n, scene);
    lbvh_keys_kernel<<<(npow2 + 255) / 256, 256, 0, s>>>(soup, n, npow2, scene, keys);
    if (sort_keys(s, keys, npow2) != 0) break;
    lbvh_leaves_kernel<<<blocks, 256, 0, s>>>(soup, n, keys, recs, leaf_box);
    if (n == 1) {
      // a single leaf is the whole tree: box of the only triangle
      if (cudaMemcpyAsync(&nodes[0].bmin, &leaf_box[0].lo, 12, cudaMemcpyDeviceToDevice, s) != cudaSuccess) break;
      if (cudaMemcpyAsync(&nodes[0].bmax, &leaf_box[0].hi, 12, cudaMemcpyDeviceToDevice, s) != cudaSuccess) break;
      const uint32_t one = 1;
      if (cudaMemcpyAsync(&nodes[0].count, &one, 4, cudaMemcpyHostToDevice, s) != cudaSuccess) break;
      if (cudaStreamSynchronize(s) != cudaSuccess) break;  // `one` lives on this stack frame
    } else {
      lbvh_hierarchy_kernel<<<blocks, 256, 0, s>>>(keys, n, children, range, parent_internal, parent_leaf, count);
      lbvh_refit_kernel<<<blocks, 256, 0, s>>>(n, children, parent_internal, parent_leaf, leaf_box, node_box, arrived);
      bool ok = true;
      for (uint32_t pass = 0; pass < rotate_passes && ok; ++pass) {
        ok = cudaMemsetAsync(arrived, 0, (size_t)n * sizeof(uint32_t), s) == cudaSuccess;
        lbvh_rotate_kernel<<<blocks, 256, 0, s>>>(n, children, parent_internal, parent_leaf, leaf_box, node_box, count, arrived);
      }
      if (!ok) break;
      lbvh_emit_kernel<<<blocks, 256, 0, s>>>(n, children, range, count, leaf_box, node_box, nodes);
    }
    if (cudaGetLastError() != cudaSuccess) break;
    rc = 0;
  } while (0);
  if (rc != 0) {
    fprintf(stderr, "[VXDRV] lbvh build failed: %s\n", cudaGetErrorString(cudaGetLastError()));
    vx_mem_free(nb); vx_mem_free(tb);
    return -1;
  }
  info->num_nodes = num_nodes;
  info->num_tris = n;
  *hnodes = nb;
  *htris = tb;
  return 0;
}

// ---- the same build without any allocation (include/vortex_cuda.h: vx_cuda_rt_build_lbvh_noalloc) ----------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The caller passes the
// input AND both output buffers; the builder's scratch (sort keys, boxes, child / parent links, arrival flags) is
// carved from one file-scope device array sized for VX_CUDA_LBVH_NOALLOC_MAX_TRIS triangles, which every GPU that
// loads this library holds until the process exits.  All calls on one GPU share that array: it keeps the last
// call's contents (what must start at zero is cleared below, on the stream, at every call), and two calls that
// overlap on the GPU -- two vx_device handles of the same GPU, whose launch streams are independent -- race.
// The launch sequence is vx_cuda_rt_build_lbvh's, restated here so that the existing entry point stays untouched.
namespace srt {
namespace {

constexpr size_t lbvh_pad(size_t bytes) { return (bytes + 255) & ~(size_t)255; }
constexpr size_t lbvh_scratch_bytes(size_t n, size_t npow2) {
  return lbvh_pad(8 * sizeof(uint32_t)) + lbvh_pad(n * sizeof(uint32_t))            // scene bounds, arrival flags
         + lbvh_pad(npow2 * sizeof(unsigned long long))                             // sort keys
         + 2 * lbvh_pad(n * sizeof(Box)) + 2 * lbvh_pad(n * sizeof(uint2))          // leaf / node boxes, children / range
         + 3 * lbvh_pad(n * sizeof(uint32_t));                                      // parents (2), subtree sizes
}
constexpr uint32_t kNoAllocMaxTris = VX_CUDA_LBVH_NOALLOC_MAX_TRIS;
static_assert((kNoAllocMaxTris & (kNoAllocMaxTris - 1)) == 0 && kNoAllocMaxTris >= kChunk, "maximum is a padded sort size");
constexpr size_t kNoAllocScratchBytes = lbvh_scratch_bytes(kNoAllocMaxTris, kNoAllocMaxTris);  // 5.5 MiB at 65 536

}  // namespace
}  // namespace srt

// In host code this name is a host-side stand-in, not a device address: kernels get the pointer that
// cudaGetSymbolAddress returns for the current device (noalloc_scratch below), never &g_lbvh_scratch[0].
static __device__ __align__(256) uint8_t g_lbvh_scratch[srt::kNoAllocScratchBytes];

namespace srt {
namespace {

constexpr int kMaxCudaDevices = 64;
std::mutex g_scratch_mutex;
uint8_t* g_scratch_ptr[kMaxCudaDevices];  // per CUDA device ordinal; filled at the first call on that device

// Device address of the scratch array on `cuda_dev`, which must be the current device.
uint8_t* noalloc_scratch(int cuda_dev) {
  if (cuda_dev < 0 || cuda_dev >= kMaxCudaDevices) return nullptr;
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_ptr[cuda_dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_lbvh_scratch);
    if (e != cudaSuccess || p == nullptr) {
      fprintf(stderr, "[VXDRV] vx_cuda_rt_build_lbvh_noalloc: cudaGetSymbolAddress on device %d: %s\n", cuda_dev,
              cudaGetErrorString(e));
      cudaGetLastError();
      return nullptr;
    }
    g_scratch_ptr[cuda_dev] = static_cast<uint8_t*>(p);
  }
  return g_scratch_ptr[cuda_dev];
}

struct Carve {  // hands out 256-byte aligned pieces of the fixed array, in Scratch's order
  uint8_t* base;
  size_t used = 0;
  explicit Carve(uint8_t* b) : base(b) {}
  template <class T> T* take(size_t count) {
    T* q = reinterpret_cast<T*>(base + used);
    used += lbvh_pad((count ? count : 1) * sizeof(T));
    return q;
  }
};

// count = 1 for the single-leaf tree; static so that the asynchronous copy never reads a dead stack frame
const uint32_t kOneTriangle = 1;

}  // namespace
}  // namespace srt

extern "C" int vx_cuda_rt_build_lbvh_noalloc(vx_device_h hdevice, vx_buffer_h hsoup, uint32_t num_triangles,
                                             vx_buffer_h hnodes, vx_buffer_h htris, srt_lbvh_info_t* info) {
  if (nullptr == hdevice || nullptr == hsoup || nullptr == hnodes || nullptr == htris || nullptr == info) return -1;
  memset(info, 0, sizeof(*info));
  vx_device* dev = static_cast<vx_device*>(hdevice);
  vx_buffer* sb = static_cast<vx_buffer*>(hsoup);
  vx_buffer* nb = static_cast<vx_buffer*>(hnodes);
  vx_buffer* tb = static_cast<vx_buffer*>(htris);
  const uint32_t n = num_triangles;
  if (n > kNoAllocMaxTris) {
    fprintf(stderr, "[VXDRV] vx_cuda_rt_build_lbvh_noalloc: %u triangles exceed the fixed scratch (maximum %u = "
                    "VX_CUDA_LBVH_NOALLOC_MAX_TRIS); use vx_cuda_rt_build_lbvh\n", n, kNoAllocMaxTris);
    return -3;
  }
  const uint32_t num_nodes = n == 1 ? 1u : 2u * n - 1u;
  if (n == 0 || sb->dev != dev || nb->dev != dev || tb->dev != dev || nb == tb || nb == sb || tb == sb ||
      (uint64_t)n * sizeof(srt_triangle_t) > sb->size || (uint64_t)num_nodes * sizeof(srt_bvh2_node_t) > nb->size ||
      (uint64_t)n * sizeof(srt_tri_rec_t) > tb->size) {
    fprintf(stderr, "[VXDRV] vx_cuda_rt_build_lbvh_noalloc: bad arguments (%u triangles need %llu node bytes and "
                    "%llu record bytes in buffers of this device)\n", n,
            (unsigned long long)num_nodes * sizeof(srt_bvh2_node_t), (unsigned long long)n * sizeof(srt_tri_rec_t));
    return -1;
  }
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  uint8_t* scratch = noalloc_scratch(dev->cuda_dev);
  if (!scratch) return -1;
  cudaStream_t s = dev->stream;
  const srt_triangle_t* soup = static_cast<const srt_triangle_t*>(sb->dptr);
  srt_bvh2_node_t* nodes = static_cast<srt_bvh2_node_t*>(nb->dptr);
  float4* recs = static_cast<float4*>(tb->dptr);
  uint32_t npow2 = kChunk;
  while (npow2 < n) npow2 <<= 1;

  Carve tmp(scratch);
  uint32_t* scene = tmp.take<uint32_t>(8);
  uint32_t* arrived = tmp.take<uint32_t>(n);
  const size_t head = tmp.used;  // scene bounds + arrival flags: cleared at every call
  unsigned long long* keys = tmp.take<unsigned long long>(npow2);
  Box* leaf_box = tmp.take<Box>(n);
  Box* node_box = tmp.take<Box>(n);
  uint2* children = tmp.take<uint2>(n);
  uint2* range = tmp.take<uint2>(n);
  uint32_t* parent_internal = tmp.take<uint32_t>(n);
  uint32_t* parent_leaf = tmp.take<uint32_t>(n);
  uint32_t* count = tmp.take<uint32_t>(n);
  if (tmp.used > kNoAllocScratchBytes) return -1;  // cannot happen for n <= the maximum (lbvh_scratch_bytes)
  const uint32_t rot_dcr = vx_dev_dcr(dev, VX_DCR_RT_LBVH_ROTATE);
  const uint32_t rotate_passes = rot_dcr == 0 ? 2u : (rot_dcr - 1u > 8u ? 8u : rot_dcr - 1u);
  int rc = -1;
  do {
    if (cudaMemsetAsync(scratch, 0, head, s) != cudaSuccess) break;
    if (cudaMemsetAsync(scene, 0xff, 3 * sizeof(uint32_t), s) != cudaSuccess) break;
    if (cudaMemsetAsync(nodes, 0, (size_t)num_nodes * sizeof(srt_bvh2_node_t), s) != cudaSuccess) break;
    const uint32_t blocks = (n + 255) / 256;
    lbvh_bounds_kernel<<<blocks < 1184 ? blocks : 1184, 256, 0, s>>>(soup, n, scene);
    lbvh_keys_kernel<<<(npow2 + 255) / 256, 256, 0, s>>>(soup, n, npow2, scene, keys);
    if (sort_keys(s, keys, npow2) != 0) break;
    lbvh_leaves_kernel<<<blocks, 256, 0, s>>>(soup, n, keys, recs, leaf_box);
    if (n == 1) {
      if (cudaMemcpyAsync(&nodes[0].bmin, &leaf_box[0].lo, 12, cudaMemcpyDeviceToDevice, s) != cudaSuccess) break;
      if (cudaMemcpyAsync(&nodes[0].bmax, &leaf_box[0].hi, 12, cudaMemcpyDeviceToDevice, s) != cudaSuccess) break;
      if (cudaMemcpyAsync(&nodes[0].count, &kOneTriangle, 4, cudaMemcpyHostToDevice, s) != cudaSuccess) break;
    } else {
      lbvh_hierarchy_kernel<<<blocks, 256, 0, s>>>(keys, n, children, range, parent_internal, parent_leaf, count);
      lbvh_refit_kernel<<<blocks, 256, 0, s>>>(n, children, parent_internal, parent_leaf, leaf_box, node_box, arrived);
      bool ok = true;
      for (uint32_t pass = 0; pass < rotate_passes && ok; ++pass) {
        ok = cudaMemsetAsync(arrived, 0, (size_t)n * sizeof(uint32_t), s) == cudaSuccess;
        lbvh_rotate_kernel<<<blocks, 256, 0, s>>>(n, children, parent_internal, parent_leaf, leaf_box, node_box, count, arrived);
      }
      if (!ok) break;
      lbvh_emit_kernel<<<blocks, 256, 0, s>>>(n, children, range, count, leaf_box, node_box, nodes);
    }
    if (cudaGetLastError() != cudaSuccess) break;
    rc = 0;
  } while (0);
  if (rc != 0) {
    fprintf(stderr, "[VXDRV] lbvh build (no allocation) failed: %s\n", cudaGetErrorString(cudaGetLastError()));
    return -1;
  }
  info->num_nodes = num_nodes;
  info->num_tris = n;
  return 0;
}
