// rt_repack.cpp -- host repack of the builder's BVH2 into the device layout (north_star: "the
// reference's BVH builder output is repacked into a 16B-aligned SoA wide-BVH node and
// precomputed-triangle layout").  Runs once per scene, on the first vx_start that sees it.
//
// Output layout (all 16-byte aligned, fetched with 128-bit read-only loads):
//   Node8, 80 B = 5 x float4: an 8-wide node whose child boxes are stored as 8-bit offsets in a
//   per-node frame (origin p, power-of-two scale per axis), children in SoA byte lanes
//   (qlo_x[8] | qlo_y[8] | qlo_z[8] | qhi_x[8] | qhi_y[8] | qhi_z[8]).  Internal children are
//   contiguous from child_base in slot order; the triangles of leaf children are contiguous from
//   tri_base.  The format follows the published compressed wide BVH of Ylitie, Karras & Laine
//   (HPG 2017); meta encoding and slot assignment are this repo's.
//   Triangle records keep the builder's 48-byte (v0|prim, e1|albedo, e2|pad) form -- the operands
//   of the intersection test must stay bit-identical -- and are only re-ordered.
//
// Boxes are quantised OUTWARD after inflating them by delta = 2^-18 * max|root coordinate|, so a
// decoded child box always contains the builder's box plus the slack the oracle's slab test uses
// (DESIGN.md "Conservative box test").  Nothing here affects which triangle a ray reports.
#include "rt_repack.h"

#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

namespace srt {

namespace {

struct Child {
  uint32_t node;  // BVH2 node index
  float area;
};

inline float half_area(const srt_bvh2_node_t& n) {
  const float dx = n.bmax[0] - n.bmin[0], dy = n.bmax[1] - n.bmin[1], dz = n.bmax[2] - n.bmin[2];
  return dx * dy + dy * dz + dz * dx;
}

// children of wide node rooted at BVH2 node `root`: open the largest internal child until 8
void collapse(const srt_bvh2_node_t* n2, uint32_t root, Child out[8], int& count) {
  count = 0;
  if (n2[root].count > 0) {  // a leaf root (tiny scene): one leaf child
    out[count++] = Child{root, half_area(n2[root])};
    return;
  }
  out[count++] = Child{n2[root].left_first, half_area(n2[n2[root].left_first])};
  out[count++] = Child{n2[root].left_first + 1, half_area(n2[n2[root].left_first + 1])};
  while (count < 8) {
    int best = -1;
    for (int i = 0; i < count; ++i)
      if (n2[out[i].node].count == 0 && (best < 0 || out[i].area > out[best].area)) best = i;
    if (best < 0) break;
    const uint32_t l = n2[out[best].node].left_first;
    out[best] = Child{l, half_area(n2[l])};
    out[count++] = Child{l + 1, half_area(n2[l + 1])};
  }
}

// slot = octant of the child centre relative to the node centre, so that visiting slots in the
// order (slot ^ ray octant) approximates front-to-back.  Collisions take the nearest free slot.
void assign_slots(const srt_bvh2_node_t* n2, const srt_bvh2_node_t& parent, const Child ch[8], int count,
                  int slot_of[8]) {
  bool used[8] = {false, false, false, false, false, false, false, false};
  int want[8];
  for (int i = 0; i < count; ++i) {
    const srt_bvh2_node_t& c = n2[ch[i].node];
    int s = 0;
    for (int k = 0; k < 3; ++k) {
      const float cc = c.bmin[k] + c.bmax[k], pc = parent.bmin[k] + parent.bmax[k];
      if (cc > pc) s |= 1 << k;
    }
    want[i] = s;
    slot_of[i] = -1;
  }
  for (int i = 0; i < count; ++i)
    if (!used[want[i]]) { slot_of[i] = want[i]; used[want[i]] = true; }
  for (int i = 0; i < count; ++i) {
    if (slot_of[i] >= 0) continue;
    int best = -1, bd = 99;
    for (int s = 0; s < 8; ++s) {
      if (used[s]) continue;
      const int d = __builtin_popcount((unsigned)(s ^ want[i]));
      if (d < bd) { bd = d; best = s; }
    }
    slot_of[i] = best;
    used[best] = true;
  }
}

}  // namespace

float scene_delta_host(const srt_bvh2_node_t& root) {
  float m = 0.0f;
  for (int k = 0; k < 3; ++k) {
    m = std::max(m, fabsf(root.bmin[k]));
    m = std::max(m, fabsf(root.bmax[k]));
  }
  return m * (1.0f / 262144.0f);
}

int repack_bvh8(const srt_bvh2_node_t* n2, uint32_t num_nodes, const srt_tri_rec_t* tris, uint32_t num_tris,
                std::vector<Node8>& out_nodes, std::vector<srt_tri_rec_t>& out_tris) {
  if (!n2 || !tris || num_nodes == 0 || num_tris == 0) return -1;
  const double delta = (double)scene_delta_host(n2[0]);
  out_nodes.clear();
  out_tris.clear();
  out_tris.reserve(num_tris);
  out_nodes.reserve(num_nodes / 4 + 8);
  std::vector<uint32_t> src;  // BVH2 node each wide node was built from (BFS order)
  src.reserve(num_nodes / 4 + 8);
  out_nodes.push_back(Node8());
  src.push_back(0);
  for (size_t w = 0; w < out_nodes.size(); ++w) {
    const srt_bvh2_node_t& parent = n2[src[w]];
    Child ch[8];
    int count, slot_of[8];
    collapse(n2, src[w], ch, count);
    assign_slots(n2, parent, ch, count, slot_of);
    int child_in_slot[8] = {-1, -1, -1, -1, -1, -1, -1, -1};
    for (int i = 0; i < count; ++i) child_in_slot[slot_of[i]] = i;

    Node8 nd;
    memset(&nd, 0, sizeof(nd));
    // frame: origin = inflated parent min (rounded down), scale = smallest 2^e covering the extent
    int e[3];
    double p[3];
    for (int k = 0; k < 3; ++k) {
      float pf = (float)((double)parent.bmin[k] - delta);
      if ((double)pf > (double)parent.bmin[k] - delta) pf = nextafterf(pf, -INFINITY);
      nd.p[k] = pf;
      p[k] = (double)pf;
      const double ext = ((double)parent.bmax[k] + delta) - p[k];
      int ek = ext > 0.0 ? (int)ceil(log2(ext / 255.0)) : -100;
      if (ek < -100) ek = -100;
      while (ceil(ext / ldexp(1.0, ek)) > 255.0) ++ek;
      if (ek > 100) return -1;  // coordinates beyond what the frame can represent
      e[k] = ek;
      nd.e[k] = (uint8_t)(ek + 127);
    }
    nd.child_base = (uint32_t)out_nodes.size();
    nd.tri_base = (uint32_t)out_tris.size();
    uint32_t rel = 0, tri_off = 0;
    for (int s = 0; s < 8; ++s) {
      const int i = child_in_slot[s];
      if (i < 0) continue;  // meta 0 = empty slot; its boxes stay 0 and are never tested
      const srt_bvh2_node_t& c = n2[ch[i].node];
      for (int k = 0; k < 3; ++k) {
        const double sc = ldexp(1.0, e[k]);
        double lo = floor((((double)c.bmin[k] - delta) - p[k]) / sc);
        double hi = ceil((((double)c.bmax[k] + delta) - p[k]) / sc);
        lo = std::min(std::max(lo, 0.0), 255.0);
        hi = std::min(std::max(hi, 0.0), 255.0);
        nd.q[k][s] = (uint8_t)lo;
        nd.q[3 + k][s] = (uint8_t)hi;
      }
      if (c.count > 0) {
        if (c.count > 4 || tri_off + c.count > 32) return -1;
        nd.meta[s] = (uint8_t)((c.count << 5) | tri_off);
        for (uint32_t t = 0; t < c.count; ++t) out_tris.push_back(tris[c.left_first + t]);
        tri_off += c.count;
      } else {
        nd.meta[s] = (uint8_t)((7u << 5) | rel);
        nd.imask |= (uint8_t)(1u << s);
        ++rel;
        out_nodes.push_back(Node8());
        src.push_back(ch[i].node);
      }
    }
    out_nodes[w] = nd;
  }
  return out_tris.size() == num_tris ? 0 : -1;
}

// ------------------------------------------------------------------------------------------------
// BVH2 -> Node4 (the layout the default traversal kernel reads, rt_wide.cuh)
// ------------------------------------------------------------------------------------------------
namespace {

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
