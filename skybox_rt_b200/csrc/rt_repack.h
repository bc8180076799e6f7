// rt_repack.h -- device layout of the wide BVH and the host repack entry point (rt_repack.cpp).
#pragma once

#include <stdint.h>

#include <vector>

#include "skybox_rt.h"

namespace srt {

// 80 bytes = 5 x 16: fetched as five float4 (ld.global.nc.v4).
struct Node8 {
  float p[3];         // frame origin (inflated node minimum)
  uint8_t e[3];       // biased exponents: scale_k = 2^(e[k]-127)
  uint8_t imask;      // bit s set: slot s holds an internal child
  uint32_t child_base;  // first internal child (children contiguous, slot order)
  uint32_t tri_base;    // first triangle record of the leaf children
  uint8_t meta[8];    // per slot: 0 empty | (7<<5)|rel child index | (count<<5)|triangle offset
  uint8_t q[6][8];    // qlo_x, qlo_y, qlo_z, qhi_x, qhi_y, qhi_z; byte lane = slot
};
static_assert(sizeof(Node8) == 80, "Node8 must be 80 bytes");

float scene_delta_host(const srt_bvh2_node_t& root);

// Returns 0 on success.  out_tris is a permutation of tris.
int repack_bvh8(const srt_bvh2_node_t* nodes2, uint32_t num_nodes, const srt_tri_rec_t* tris, uint32_t num_tris,
                std::vector<Node8>& out_nodes, std::vector<srt_tri_rec_t>& out_tris);

}  // namespace srt
