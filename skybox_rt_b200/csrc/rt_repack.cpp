// rt_repack.cpp -- host repack of the builder's BVH2 into the device layout (north_star: "the
// reference's BVH builder output is repacked into a 16B-aligned SoA wide-BVH node and
// precomputed-triangle layout").  Runs once per scene, on the first vx_start that sees it.
//
// Output layout: Node4 (rt_repack.h), 128 B = one cache line per 4-wide node, full-precision child boxes in SoA
// planes, one child slot per TRIANGLE with the triangle's own box.  Triangle records keep the builder's 48-byte
// (v0|prim, e1|albedo, e2|pad) form -- the operands of the intersection test must stay bit-identical -- and are
// not even re-ordered: triangle refs index the app's record array.
// (A compressed 8-wide node image after Ylitie, Karras & Laine was written in round 1 and never traversed; it was
// removed in round 2 -- DESIGN.md section 7 has the measurements that decided against it.)
//
// Boxes are rounded OUTWARD after inflating them by delta = 2^-18 * max|root coordinate|, so a child box always
// contains the builder's box plus the slack the oracle's slab test uses (DESIGN.md "Conservative box test").
// Nothing here affects which triangle a ray reports.
#include "rt_repack.h"

#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

namespace srt {

float scene_delta_host(const srt_bvh2_node_t& root) {
  float m = 0.0f;
  for (int k = 0; k < 3; ++k) {
    m = std::max(m, fabsf(root.bmin[k]));
    m = std::max(m, fabsf(root.bmax[k]));
  }
  return m * (1.0f / 262144.0f);
}

// ------------------------------------------------------------------------------------------------
// BVH2 -> Node4 (the layout the default traversal kernel reads, rt_wide.cuh)
// ------------------------------------------------------------------------------------------------
namespace {

inline float half_area(const srt_bvh2_node_t& n) {
  const float dx = n.bmax[0] - n.bmin[0], dy = n.bmax[1] - n.bmin[1], dz = n.bmax[2] - n.bmin[2];
  return dx * dy + dy * dz + dz * dx;
}

inline float round_down(double v) {
  float f = (float)v;
  if ((double)f > v) f = nextafterf(f, -INFINITY);
  return f;
}
inline float round_up(double v) {
  float f = (float)v;
  if ((double)f < v) f = nextafterf(f, INFINITY);
  return f;
}

// One prospective child of a wide node: a BVH2 node (inner, or a leaf not yet opened) or one triangle.
struct Ent {
  uint32_t id;
  bool tri;
  float area;
};

// Children of the wide node built from BVH2 node `root`.  Starting from root's two children, the
// entry with the largest surface area that still fits is opened: an inner node is replaced by its
// two children, a leaf by its (<= 4) triangles.  Every child ends up an inner node (or a leaf that
// did not fit and becomes a wide node of its own one level down) or a single triangle.
int collapse4(const srt_bvh2_node_t* n2, uint32_t num_nodes, uint32_t num_tris, uint32_t root, Ent out[4],
              int& count) {
  count = 0;
  auto add_node = [&](uint32_t n) -> int {
    if (n >= num_nodes || n == 0) return -1;
    // left_first is a triangle index only in a leaf; in an inner node it is a node index (checked when opened)
    if (n2[n].count > SRT_MAX_LEAF_TRIS || (n2[n].count > 0 && (uint64_t)n2[n].left_first + n2[n].count > num_tris)) return -1;
    out[count++] = Ent{n, false, half_area(n2[n])};
    return 0;
  };
  const srt_bvh2_node_t& r = n2[root];
  if (r.count > 0) {  // the wide node is a leaf's triangle fan
    if (r.count > SRT_MAX_LEAF_TRIS || (uint64_t)r.left_first + r.count > num_tris) return -1;
    for (uint32_t t = 0; t < r.count; ++t) out[count++] = Ent{r.left_first + t, true, 0.0f};
    return 0;
  }
  if (add_node(r.left_first) || add_node(r.left_first + 1)) return -1;
  for (;;) {
    int best = -1;
    for (int i = 0; i < count; ++i) {
      if (out[i].tri) continue;
      const srt_bvh2_node_t& c = n2[out[i].id];
      const int grow = c.count > 0 ? (int)c.count - 1 : 1;
      if (count + grow > 4) continue;
      if (best < 0 || out[i].area > out[best].area) best = i;
    }
    if (best < 0) break;
    const srt_bvh2_node_t c = n2[out[best].id];
    if (c.count > 0) {
      out[best] = Ent{c.left_first, true, 0.0f};
      for (uint32_t t = 1; t < c.count; ++t) out[count++] = Ent{c.left_first + t, true, 0.0f};
    } else {
      const uint32_t l = c.left_first;
      if (l + 1 >= num_nodes || l == 0) return -1;
      if (n2[l].count > SRT_MAX_LEAF_TRIS || (n2[l].count > 0 && (uint64_t)n2[l].left_first + n2[l].count > num_tris)) return -1;
      out[best] = Ent{l, false, half_area(n2[l])};
      if (add_node(l + 1)) return -1;
    }
  }
  return 0;
}

}  // namespace

int repack_bvh4(const srt_bvh2_node_t* n2, uint32_t num_nodes, const srt_tri_rec_t* tris, uint32_t num_tris,
                std::vector<Node4>& out, uint32_t* max_depth) {
  if (!n2 || !tris || num_nodes == 0 || num_tris == 0 || num_tris > 0x7FFFFFFFu) return -1;
  if (n2[0].count > SRT_MAX_LEAF_TRIS || (n2[0].count > 0 && n2[0].count > num_tris)) return -1;
  const double delta = (double)scene_delta_host(n2[0]);
  out.clear();
  out.reserve(num_nodes / 2 + 8);
  std::vector<uint32_t> src, depth;  // BVH2 node each wide node was built from, and its level
  src.reserve(num_nodes / 2 + 8);
  depth.reserve(num_nodes / 2 + 8);
  out.push_back(Node4());
  src.push_back(0);
  depth.push_back(1);
  uint32_t deepest = 1;
  for (size_t w = 0; w < out.size(); ++w) {
    Ent ch[4];
    int count;
    if (collapse4(n2, num_nodes, num_tris, src[w], ch, count) != 0) return -1;
    Node4 nd;
    for (int s = 0; s < 4; ++s) {
      for (int k = 0; k < 6; ++k) nd.plane[k][s] = (k & 1) ? -INFINITY : INFINITY;
      nd.ref[s] = kEmptyRef;
      nd.pad[s] = 0;
    }
    for (int s = 0; s < count; ++s) {
      double lo[3], hi[3];
      if (ch[s].tri) {
        const srt_tri_rec_t& t = tris[ch[s].id];
        for (int k = 0; k < 3; ++k) {
          const double a = (double)t.v0[k], b = a + (double)t.e1[k], c = a + (double)t.e2[k];
          lo[k] = std::min(a, std::min(b, c));
          hi[k] = std::max(a, std::max(b, c));
        }
        nd.ref[s] = kTriBit | ch[s].id;
      } else {
        const srt_bvh2_node_t& c = n2[ch[s].id];
        for (int k = 0; k < 3; ++k) { lo[k] = (double)c.bmin[k]; hi[k] = (double)c.bmax[k]; }
        if (out.size() >= (size_t)num_nodes + 1) return -1;  // more wide nodes than BVH2 nodes: a cycle
        nd.ref[s] = (uint32_t)out.size();
        out.push_back(Node4());
        src.push_back(ch[s].id);
        depth.push_back(depth[w] + 1);
        if (depth[w] + 1 > deepest) deepest = depth[w] + 1;
      }
      for (int k = 0; k < 3; ++k) {
        nd.plane[2 * k][s] = round_down(lo[k] - delta);
        nd.plane[2 * k + 1][s] = round_up(hi[k] + delta);
      }
    }
    out[w] = nd;
  }
  if (max_depth) *max_depth = deepest;
  return 0;
}

}  // namespace srt
