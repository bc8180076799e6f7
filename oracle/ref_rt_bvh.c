/*
 * ref_rt_bvh.c -- oracle BVH2 builder, ray/triangle and ray/box tests, traversal.
 * TEST INFRASTRUCTURE ONLY (see ref_rt.c).  PARITY UNPINNED BY THE REFERENCE (no RT code
 * exists in /root/reference; SURVEY.md section 8a rows T1, T3, T4 are builder-defined).
 *
 * The builder here is deliberately NOT the product's builder: it splits at the spatial
 * median of the centroid box (object median as fallback), the product uses binned SAH.
 * Closest-hit results are defined to be independent of tree shape (min over (t, prim_id)
 * of all accepted triangles), so product-on-its-own-tree vs oracle-on-this-tree is a
 * stronger parity check than sharing one tree.
 */
#include "ref_rt.h"

/* ---------------------------------------------------------------- builder ---------- */
typedef struct {
  const srt_triangle_t* soup;
  uint32_t* idx;     /* permutation being partitioned */
  float* cent;       /* 3 floats per input triangle   */
  srt_bvh2_node_t* nodes;
  uint32_t num_nodes;
} build_ctx;

static void tri_bounds(const srt_triangle_t* t, float lo[3], float hi[3]) {
  int k;
  for (k = 0; k < 3; ++k) {
    float a = t->v0[k], b = t->v1[k], c = t->v2[k];
    float mn = a < b ? a : b, mx = a > b ? a : b;
    lo[k] = mn < c ? mn : c;
    hi[k] = mx > c ? mx : c;
  }
}

static void range_bounds(build_ctx* c, uint32_t first, uint32_t count, float lo[3], float hi[3],
                         float clo[3], float chi[3]) {
  uint32_t i;
  int k;
  for (k = 0; k < 3; ++k) { lo[k] = clo[k] = 3.0e38f; hi[k] = chi[k] = -3.0e38f; }
  for (i = first; i < first + count; ++i) {
    float tl[3], th[3];
    const float* ce = c->cent + 3 * (size_t)c->idx[i];
    tri_bounds(c->soup + c->idx[i], tl, th);
    for (k = 0; k < 3; ++k) {
      if (tl[k] < lo[k]) lo[k] = tl[k];
      if (th[k] > hi[k]) hi[k] = th[k];
      if (ce[k] < clo[k]) clo[k] = ce[k];
      if (ce[k] > chi[k]) chi[k] = ce[k];
    }
  }
}

static int cmp_axis;
static const float* cmp_cent;
static int cmp_idx(const void* pa, const void* pb) {
  uint32_t a = *(const uint32_t*)pa, b = *(const uint32_t*)pb;
  float ca = cmp_cent[3 * (size_t)a + cmp_axis], cb = cmp_cent[3 * (size_t)b + cmp_axis];
  if (ca < cb) return -1;
  if (ca > cb) return 1;
  return a < b ? -1 : (a > b ? 1 : 0);
}

static void build_node(build_ctx* c, uint32_t node, uint32_t first, uint32_t count) {
  float lo[3], hi[3], clo[3], chi[3];
  srt_bvh2_node_t* n = c->nodes + node;
  int k, axis = 0;
  uint32_t i, j, mid, left;
  float ext, split;
  range_bounds(c, first, count, lo, hi, clo, chi);
  for (k = 0; k < 3; ++k) { n->bmin[k] = lo[k]; n->bmax[k] = hi[k]; }
  if (count <= SRT_MAX_LEAF_TRIS) {
    n->left_first = first;
    n->count = count;
    return;
  }
  ext = chi[0] - clo[0];
  if (chi[1] - clo[1] > ext) { ext = chi[1] - clo[1]; axis = 1; }
  if (chi[2] - clo[2] > ext) { ext = chi[2] - clo[2]; axis = 2; }
  split = 0.5f * (clo[axis] + chi[axis]);
  i = first;
  j = first + count;
  while (i < j) {
    if (c->cent[3 * (size_t)c->idx[i] + axis] < split) ++i;
    else { uint32_t t = c->idx[i]; c->idx[i] = c->idx[--j]; c->idx[j] = t; }
  }
  mid = i;
  if (mid == first || mid == first + count) { /* all centroids on one side: object median */
    cmp_axis = axis;
    cmp_cent = c->cent;
    qsort(c->idx + first, count, sizeof(uint32_t), cmp_idx);
    mid = first + count / 2;
  }
  left = c->num_nodes;
  c->num_nodes += 2;
  n->left_first = left;
  n->count = 0;
  build_node(c, left, first, mid - first);
  build_node(c, left + 1, mid, first + count - mid);
}

int ort_bvh2_build(const srt_triangle_t* soup, uint32_t num_tris, srt_bvh2_t* out) {
  build_ctx c;
  uint32_t i;
  int k;
  memset(out, 0, sizeof(*out));
  if (num_tris == 0) return -1;
  c.soup = soup;
  c.idx = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)num_tris);
  c.cent = (float*)malloc(sizeof(float) * 3 * (size_t)num_tris);
  c.nodes = (srt_bvh2_node_t*)calloc(2 * (size_t)num_tris, sizeof(srt_bvh2_node_t));
  out->tris = (srt_tri_rec_t*)calloc(num_tris, sizeof(srt_tri_rec_t));
  if (!c.idx || !c.cent || !c.nodes || !out->tris) {
    free(c.idx); free(c.cent); free(c.nodes); free(out->tris);
    return -1;
  }
  for (i = 0; i < num_tris; ++i) {
    float lo[3], hi[3];
    c.idx[i] = i;
    tri_bounds(soup + i, lo, hi);
    for (k = 0; k < 3; ++k) c.cent[3 * (size_t)i + k] = 0.5f * (lo[k] + hi[k]);
  }
  c.num_nodes = 1;
  build_node(&c, 0, 0, num_tris);
  for (i = 0; i < num_tris; ++i) {
    const srt_triangle_t* s = soup + c.idx[i];
    srt_tri_rec_t* r = out->tris + i;
    for (k = 0; k < 3; ++k) {
      r->v0[k] = s->v0[k];
      r->e1[k] = s->v1[k] - s->v0[k];
      r->e2[k] = s->v2[k] - s->v0[k];
    }
    r->prim_id = c.idx[i];
    r->albedo = s->albedo;
    r->pad = 0;
  }
  out->nodes = c.nodes;
  out->num_nodes = c.num_nodes;
  out->num_tris = num_tris;
  free(c.idx);
  free(c.cent);
  return 0;
}

void ort_bvh2_free(srt_bvh2_t* b) {
  free(b->nodes);
  free(b->tris);
  memset(b, 0, sizeof(*b));
}

/* ---------------------------------------------------------------- ray / triangle ----
 * Moeller-Trumbore on the precomputed (v0, e1, e2).  Returns 1 when the supporting
 * plane is crossed inside the triangle; the caller applies the (tmin, tmax) window. */
int ort_intersect_tri(const srt_tri_rec_t* tr, v3 o, v3 d, float* t, float* u, float* v) {
  v3 e1 = vload(tr->e1), e2 = vload(tr->e2), v0 = vload(tr->v0);
  v3 p = vcross(d, e2);
  float det = vdot(e1, p);
  float inv, uu, vv;
  v3 tv, q;
  if (det == 0.0f) return 0;
  inv = 1.0f / det;
  tv = vsub(o, v0);
  uu = vdot(tv, p) * inv;
  if (!(uu >= 0.0f) || uu > 1.0f) return 0;
  q = vcross(tv, e1);
  vv = vdot(d, q) * inv;
  if (!(vv >= 0.0f) || uu + vv > 1.0f) return 0;
  *t = vdot(e2, q) * inv;
  *u = uu;
  *v = vv;
  return 1;
}

/* ---------------------------------------------------------------- ray / box --------
 * Conservative slab test: it may only cull boxes that contain no triangle the
 * Moeller-Trumbore test would accept, so that the traversal result equals the brute-force
 * minimum over all triangles.  Three safeguards:
 *   - zero direction components are replaced by +-1e-20, so no NaN can appear;
 *   - every box is inflated by `delta` = 2^-18 * (largest |coordinate| of the root box): rays
 *     that run exactly in a face plane of an axis-aligned box (or a hair outside it) stay inside;
 *   - the exit distance is widened by ORT_PAD = 1 + 2^-13 against rounding in the slab products. */
static inline float safe_inv(float d) {
  float a = d < 0.0f ? -d : d;
  float dd = a < 1e-20f ? (d < 0.0f ? -1e-20f : 1e-20f) : d;
  return 1.0f / dd;
}
#define ORT_STACK 128
static inline float fmin2(float a, float b) { return a < b ? a : b; }
static inline float fmax2(float a, float b) { return a > b ? a : b; }

static inline float scene_delta(const srt_bvh2_node_t* root) {
  float m = 0.0f;
  int k;
  for (k = 0; k < 3; ++k) {
    float a = root->bmin[k] < 0.0f ? -root->bmin[k] : root->bmin[k];
    float b = root->bmax[k] < 0.0f ? -root->bmax[k] : root->bmax[k];
    if (a > m) m = a;
    if (b > m) m = b;
  }
  return m * (1.0f / 262144.0f);
}

static inline int slab(const srt_bvh2_node_t* n, float dl, v3 o, v3 id, float tmin, float tmax, float* tnear) {
  float x0 = ((n->bmin[0] - dl) - o.x) * id.x, x1 = ((n->bmax[0] + dl) - o.x) * id.x;
  float y0 = ((n->bmin[1] - dl) - o.y) * id.y, y1 = ((n->bmax[1] + dl) - o.y) * id.y;
  float z0 = ((n->bmin[2] - dl) - o.z) * id.z, z1 = ((n->bmax[2] + dl) - o.z) * id.z;
  float tn = fmax2(fmax2(fmin2(x0, x1), fmin2(y0, y1)), fmax2(fmin2(z0, z1), tmin));
  float tf = fmin2(fmin2(fmax2(x0, x1), fmax2(y0, y1)), fmin2(fmax2(z0, z1), tmax));
  *tnear = tn;
  return tn <= tf * ORT_PAD;
}

/* closest hit = min over (t, prim_id) of accepted triangles with tmin < t < tmax */
static inline void consider(const srt_tri_rec_t* tr, v3 o, v3 d, float tmin, srt_hit_t* best) {
  float t, u, v;
  if (!ort_intersect_tri(tr, o, d, &t, &u, &v)) return;
  if (!(t > tmin)) return;
  if (t < best->t || (t == best->t && best->prim_id != SRT_MISS && tr->prim_id < best->prim_id)) {
    best->t = t; best->u = u; best->v = v; best->prim_id = tr->prim_id;
  }
}

void ort_closest_hit(const srt_bvh2_node_t* nodes, const srt_tri_rec_t* tris, const srt_ray_t* ray,
                     srt_hit_t* hit, ort_counters_t* ctr) {
  uint32_t stack[ORT_STACK];
  float stack_t[ORT_STACK];
  int sp = 0;
  v3 o = vload(ray->o), d = vload(ray->d);
  v3 id = V(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
  srt_hit_t best;
  float tn, dl = scene_delta(nodes);
  uint64_t nv = 1, tt = 0;
  best.t = ray->tmax; best.u = 0.0f; best.v = 0.0f; best.prim_id = SRT_MISS;
  if (slab(nodes, dl, o, id, ray->tmin, best.t, &tn)) { stack[0] = 0; stack_t[0] = tn; sp = 1; }
  while (sp > 0) {
    const srt_bvh2_node_t* n;
    --sp;
    if (stack_t[sp] > best.t * ORT_PAD) continue; /* best.t shrank since the push */
    n = nodes + stack[sp];
    if (n->count > 0) {
      uint32_t i;
      for (i = 0; i < n->count; ++i) consider(tris + n->left_first + i, o, d, ray->tmin, &best);
      tt += n->count;
    } else {
      float t0, t1;
      uint32_t l = n->left_first, r = l + 1;
      int h0 = slab(nodes + l, dl, o, id, ray->tmin, best.t, &t0);
      int h1 = slab(nodes + r, dl, o, id, ray->tmin, best.t, &t1);
      nv += 2;
      if (sp + 2 > ORT_STACK) abort();
      if (h0 && h1) { /* far child first, near child on top */
        if (t1 < t0) { uint32_t s = l; float ts = t0; l = r; r = s; t0 = t1; t1 = ts; }
        stack[sp] = r; stack_t[sp++] = t1;
        stack[sp] = l; stack_t[sp++] = t0;
      } else if (h0) { stack[sp] = l; stack_t[sp++] = t0; }
      else if (h1) { stack[sp] = r; stack_t[sp++] = t1; }
    }
  }
  *hit = best;
  if (ctr) { ctr->nodes_visited += nv; ctr->tris_tested += tt; }
}

int ort_any_hit(const srt_bvh2_node_t* nodes, const srt_tri_rec_t* tris, const srt_ray_t* ray,
                ort_counters_t* ctr) {
  uint32_t stack[ORT_STACK];
  int sp = 0, found = 0;
  v3 o = vload(ray->o), d = vload(ray->d);
  v3 id = V(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
  float tn, dl = scene_delta(nodes);
  uint64_t nv = 1, tt = 0;
  if (slab(nodes, dl, o, id, ray->tmin, ray->tmax, &tn)) stack[sp++] = 0;
  while (sp > 0 && !found) {
    const srt_bvh2_node_t* n = nodes + stack[--sp];
    if (n->count > 0) {
      uint32_t i;
      for (i = 0; i < n->count && !found; ++i) {
        float t, u, v;
        ++tt;
        if (ort_intersect_tri(tris + n->left_first + i, o, d, &t, &u, &v) && t > ray->tmin && t < ray->tmax)
          found = 1;
      }
    } else {
      uint32_t l = n->left_first;
      nv += 2;
      if (sp + 2 > ORT_STACK) abort();
      if (slab(nodes + l, dl, o, id, ray->tmin, ray->tmax, &tn)) stack[sp++] = l;
      if (slab(nodes + l + 1, dl, o, id, ray->tmin, ray->tmax, &tn)) stack[sp++] = l + 1;
    }
  }
  if (ctr) { ctr->nodes_visited += nv; ctr->tris_tested += tt; }
  return found;
}

/* ---------------------------------------------------------------- brute force ------ */
void ort_closest_hit_brute(const srt_tri_rec_t* tris, uint32_t num_tris, const srt_ray_t* ray, srt_hit_t* hit) {
  v3 o = vload(ray->o), d = vload(ray->d);
  srt_hit_t best;
  uint32_t i;
  best.t = ray->tmax; best.u = 0.0f; best.v = 0.0f; best.prim_id = SRT_MISS;
  for (i = 0; i < num_tris; ++i) consider(tris + i, o, d, ray->tmin, &best);
  *hit = best;
}

int ort_any_hit_brute(const srt_tri_rec_t* tris, uint32_t num_tris, const srt_ray_t* ray) {
  v3 o = vload(ray->o), d = vload(ray->d);
  uint32_t i;
  for (i = 0; i < num_tris; ++i) {
    float t, u, v;
    if (ort_intersect_tri(tris + i, o, d, &t, &u, &v) && t > ray->tmin && t < ray->tmax) return 1;
  }
  return 0;
}
