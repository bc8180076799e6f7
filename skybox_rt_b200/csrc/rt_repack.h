// rt_repack.h -- device layout of the wide BVH and the host repack entry point (rt_repack.cpp).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "skybox_rt.h"

namespace srt {

// 128 bytes = 8 x float4 = one cache line: a 4-wide node with full-precision child boxes in SoA
// planes, so that one 16-byte load brings the same plane of all four children.
//   q[0] = lo.x[4]  q[1] = hi.x[4]  q[2] = lo.y[4]  q[3] = hi.y[4]  q[4] = lo.z[4]  q[5] = hi.z[4]
//   q[6] = child refs[4] (as raw bits)   q[7] = unused
// Child ref: bit 31 clear = index of another Node4; bit 31 set = ONE triangle, bits 0-30 = its record
// index in the builder's own, unpermuted record array (the child box is then that triangle's box, so
// the ray/triangle test only runs for triangles whose box the ray enters).
// Empty slots hold the inverted box lo = +inf, hi = -inf: entry = +inf > exit = -inf for every direction sign and
// every [tmin, tmax], tmax = +inf included (1 / d is clamped to a finite value, so no inf * 0).
struct Node4 {
  float plane[6][4];
  uint32_t ref[4];
  uint32_t pad[4];
};
static_assert(sizeof(Node4) == 128, "Node4 must be 128 bytes");
constexpr uint32_t kTriBit = 0x80000000u;
constexpr uint32_t kEmptyRef = 0xFFFFFFFFu;

float scene_delta_host(const srt_bvh2_node_t& root);

// BVH2 -> Node4.  Triangle records are NOT permuted: triangle refs index the caller's record array.
// max_depth = number of Node4 levels (root = 1).  Returns 0 on success, -1 on a malformed BVH2
// (child / triangle range out of bounds, leaf larger than SRT_MAX_LEAF_TRIS, cycle).
int repack_bvh4(const srt_bvh2_node_t* nodes2, uint32_t num_nodes, const srt_tri_rec_t* tris, uint32_t num_tris,
                std::vector<Node4>& out_nodes, uint32_t* max_depth);

// The same on the device (rt_repack_device.cu): d_nodes2 / d_tris are device pointers, *d_out is cudaMalloc'ed and
// owned by the caller.  Byte-identical to repack_bvh4's output.
// per_level: the round-1 schedule (five launches and a read-back per tree level) instead of one cooperative launch.
int repack_bvh4_device(cudaStream_t stream, const srt_bvh2_node_t* d_nodes2, uint32_t num_nodes,
                       const srt_tri_rec_t* d_tris, uint32_t num_tris, float4** d_out, uint32_t* num_nodes4,
                       uint32_t* max_depth, bool per_level = false);

}  // namespace srt
