/*
 * rt_bvh2.c -- host pre-pass of the raytrace app: triangle soup -> flat BVH2 + triangle records.
 *
 * Plays the role graphics::Binning plays for draw3d (reference sim/common/gfxutil.cpp:103-276,
 * called from tests/regression/draw3d/main.cpp:190): a host function that turns application
 * geometry into the flat device buffers the kernel consumes, serialised as arrays of PODs
 * (gfxutil.cpp:253-271).  The reference has no BVH builder, so
 * the algorithm is this repo's: top-down binned SAH, 16 bins on the centroid box, leaves of at
 * most SRT_MAX_LEAF_TRIS triangles, object-median fallback, children stored as adjacent pairs.
 * Deterministic: no threads race on the output, ties go to the lower bin / lower index.
 *
 * Plain C, no CUDA.  Boxes are exact fp32 min/max of vertices; nothing here needs to be
 * bit-reproduced elsewhere (closest-hit results do not depend on tree shape).
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "skybox_rt.h"

#define NBINS 16

typedef struct { float lo[3], hi[3]; } box_t;

typedef struct {
  const srt_triangle_t* soup;
  box_t* tb;        /* per input triangle bounds              */
  uint32_t* idx;    /* permutation, partitioned in place      */
  srt_bvh2_node_t* nodes;
  uint32_t num_nodes;
} builder_t;

static inline void box_reset(box_t* b) {
  b->lo[0] = b->lo[1] = b->lo[2] = 3.0e38f;
  b->hi[0] = b->hi[1] = b->hi[2] = -3.0e38f;
}
static inline void box_grow(box_t* b, const box_t* o) {
  int k;
  for (k = 0; k < 3; ++k) {
    if (o->lo[k] < b->lo[k]) b->lo[k] = o->lo[k];
    if (o->hi[k] > b->hi[k]) b->hi[k] = o->hi[k];
  }
}
static inline float box_half_area(const box_t* b) {
  float dx = b->hi[0] - b->lo[0], dy = b->hi[1] - b->lo[1], dz = b->hi[2] - b->lo[2];
  if (dx < 0.0f) return 0.0f;
  return dx * dy + dy * dz + dz * dx;
}
static inline float centroid(const box_t* b, int k) { return 0.5f * (b->lo[k] + b->hi[k]); }

/* object-median fallback: sort the range by (centroid[axis], index) -- a total order, so the
 * result is deterministic.  Rare (only when binning cannot separate the centroids). */
typedef struct { const box_t* tb; int axis; } sort_ctx_t;
static int key_cmp(const void* pa, const void* pb, void* pc) {
  const sort_ctx_t* c = (const sort_ctx_t*)pc;
  uint32_t a = *(const uint32_t*)pa, b = *(const uint32_t*)pb;
  float ca = centroid(&c->tb[a], c->axis), cb = centroid(&c->tb[b], c->axis);
  if (ca < cb) return -1;
  if (ca > cb) return 1;
  return a < b ? -1 : (a > b ? 1 : 0);
}
static void median_partition(builder_t* B, uint32_t first, uint32_t count, int axis) {
  sort_ctx_t c;
  c.tb = B->tb;
  c.axis = axis;
  qsort_r(B->idx + first, count, sizeof(uint32_t), key_cmp, &c);
}

typedef struct { uint32_t node, first, count; } work_t;

static int build(builder_t* B, uint32_t num_tris) {
  work_t* stack = (work_t*)malloc(sizeof(work_t) * 256);
  uint32_t sp = 0, cap = 256;
  if (!stack) return -1;
  stack[sp].node = 0; stack[sp].first = 0; stack[sp].count = num_tris; ++sp;
  B->num_nodes = 1;
  while (sp > 0) {
    work_t w = stack[--sp];
    srt_bvh2_node_t* n = B->nodes + w.node;
    box_t bounds, cb;
    uint32_t i, mid;
    int k, axis = -1;
    box_reset(&bounds);
    box_reset(&cb);
    for (i = w.first; i < w.first + w.count; ++i) {
      const box_t* t = &B->tb[B->idx[i]];
      box_grow(&bounds, t);
      for (k = 0; k < 3; ++k) {
        float c = centroid(t, k);
        if (c < cb.lo[k]) cb.lo[k] = c;
        if (c > cb.hi[k]) cb.hi[k] = c;
      }
    }
    for (k = 0; k < 3; ++k) { n->bmin[k] = bounds.lo[k]; n->bmax[k] = bounds.hi[k]; }
    if (w.count <= SRT_MAX_LEAF_TRIS) {
      n->left_first = w.first;
      n->count = w.count;
      continue;
    }
    /* --- binned SAH over the three axes --- */
    {
      float best_cost = 3.0e38f;
      int best_bin = -1;
      for (k = 0; k < 3; ++k) {
        box_t bb[NBINS], right_acc[NBINS];
        uint32_t bn[NBINS];
        box_t acc;
        uint32_t cnt;
        float ext = cb.hi[k] - cb.lo[k], scale;
        int b;
        if (!(ext > 0.0f)) continue;
        scale = (float)NBINS / ext;
        for (b = 0; b < NBINS; ++b) { box_reset(&bb[b]); bn[b] = 0; }
        for (i = w.first; i < w.first + w.count; ++i) {
          const box_t* t = &B->tb[B->idx[i]];
          int bi = (int)((centroid(t, k) - cb.lo[k]) * scale);
          if (bi > NBINS - 1) bi = NBINS - 1;
          if (bi < 0) bi = 0;
          box_grow(&bb[bi], t);
          bn[bi]++;
        }
        box_reset(&acc);
        for (b = NBINS - 1; b > 0; --b) { box_grow(&acc, &bb[b]); right_acc[b] = acc; }
        box_reset(&acc);
        cnt = 0;
        for (b = 0; b < NBINS - 1; ++b) { /* split between bin b and b+1 */
          float cost;
          box_grow(&acc, &bb[b]);
          cnt += bn[b];
          if (cnt == 0 || cnt == w.count) continue;
          cost = box_half_area(&acc) * (float)cnt + box_half_area(&right_acc[b + 1]) * (float)(w.count - cnt);
          if (cost < best_cost) { best_cost = cost; best_bin = b; axis = k; }
        }
      }
      if (axis >= 0) {
        float ext = cb.hi[axis] - cb.lo[axis], scale = (float)NBINS / ext;
        uint32_t a = w.first, z = w.first + w.count;
        while (a < z) {
          const box_t* t = &B->tb[B->idx[a]];
          int bi = (int)((centroid(t, axis) - cb.lo[axis]) * scale);
          if (bi > NBINS - 1) bi = NBINS - 1;
          if (bi < 0) bi = 0;
          if (bi <= best_bin) ++a;
          else { uint32_t tmp = B->idx[a]; B->idx[a] = B->idx[--z]; B->idx[z] = tmp; }
        }
        mid = a;
      } else {
        mid = w.first; /* force fallback */
      }
      if (mid == w.first || mid == w.first + w.count) {
        int ax = 0;
        float e = bounds.hi[0] - bounds.lo[0];
        if (bounds.hi[1] - bounds.lo[1] > e) { e = bounds.hi[1] - bounds.lo[1]; ax = 1; }
        if (bounds.hi[2] - bounds.lo[2] > e) ax = 2;
        median_partition(B, w.first, w.count, ax);
        mid = w.first + w.count / 2;
      }
    }
    n->left_first = B->num_nodes;
    n->count = 0;
    if (sp + 2 > cap) {
      work_t* ns = (work_t*)realloc(stack, sizeof(work_t) * cap * 2);
      if (!ns) { free(stack); return -1; }
      stack = ns;
      cap *= 2;
    }
    /* push right first so the left subtree is laid out right after its parent pair */
    stack[sp].node = B->num_nodes + 1; stack[sp].first = mid; stack[sp].count = w.first + w.count - mid; ++sp;
    stack[sp].node = B->num_nodes; stack[sp].first = w.first; stack[sp].count = mid - w.first; ++sp;
    B->num_nodes += 2;
  }
  free(stack);
  return 0;
}

int srt_bvh2_build(const srt_triangle_t* soup, uint32_t num_tris, srt_bvh2_t* out) {
  builder_t B;
  uint32_t i;
  int k;
  if (!out) return -1;
  memset(out, 0, sizeof(*out));
  if (!soup || num_tris == 0) return -1;
  B.soup = soup;
  B.tb = (box_t*)malloc(sizeof(box_t) * (size_t)num_tris);
  B.idx = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)num_tris);
  B.nodes = (srt_bvh2_node_t*)calloc(2 * (size_t)num_tris, sizeof(srt_bvh2_node_t));
  out->tris = (srt_tri_rec_t*)calloc(num_tris, sizeof(srt_tri_rec_t));
  if (!B.tb || !B.idx || !B.nodes || !out->tris) goto fail;
  for (i = 0; i < num_tris; ++i) {
    const srt_triangle_t* t = soup + i;
    B.idx[i] = i;
    for (k = 0; k < 3; ++k) {
      float a = t->v0[k], b = t->v1[k], c = t->v2[k];
      float mn = a < b ? a : b, mx = a > b ? a : b;
      B.tb[i].lo[k] = mn < c ? mn : c;
      B.tb[i].hi[k] = mx > c ? mx : c;
    }
  }
  if (build(&B, num_tris) != 0) goto fail;
  for (i = 0; i < num_tris; ++i) {
    const srt_triangle_t* s = soup + B.idx[i];
    srt_tri_rec_t* r = out->tris + i;
    for (k = 0; k < 3; ++k) {
      r->v0[k] = s->v0[k];
      r->e1[k] = s->v1[k] - s->v0[k];
      r->e2[k] = s->v2[k] - s->v0[k];
    }
    r->prim_id = B.idx[i];
    r->albedo = s->albedo;
    r->pad = 0;
  }
  out->nodes = (srt_bvh2_node_t*)realloc(B.nodes, sizeof(srt_bvh2_node_t) * (size_t)B.num_nodes);
  if (!out->nodes) out->nodes = B.nodes;
  out->num_nodes = B.num_nodes;
  out->num_tris = num_tris;
  free(B.tb);
  free(B.idx);
  return 0;
fail:
  free(B.tb); free(B.idx); free(B.nodes); free(out->tris);
  memset(out, 0, sizeof(*out));
  return -1;
}

void srt_bvh2_free(srt_bvh2_t* bvh) {
  if (!bvh) return;
  free(bvh->nodes);
  free(bvh->tris);
  memset(bvh, 0, sizeof(*bvh));
}
