/*
 * draw3d_binning.c -- host pre-pass of the raster path: triangle setup + tile binning (plain C).
 *
 * Produces what the reference's graphics::Binning produces (sim/common/gfxutil.cpp:103-276): `primbuf`,
 * one 120-byte rast_prim_t per kept triangle (fixed-point edge equations + attribute planes), and
 * `tilebuf`, the non-empty 32x32 tiles in (tile_x, tile_y) order each with the ids of the primitives
 * whose bounding box touches it, in submission order.
 *
 * Built differently from the reference (which inserts into a std::map of vectors per primitive): one
 * setup pass that also counts primitives per tile on a dense grid, an exclusive scan, and one fill pass
 * -- O(prims + tile-prim pairs), no allocation per tile, and the layout a GPU binning pass would use.
 * All float expressions keep the reference's operation order (compile with -ffp-contract=off).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "skybox_draw3d.h"

typedef struct { float x, y, z, w; } vec4;
typedef struct { uint16_t tx0, tx1, ty0, ty1; } tile_rect_t; /* [tx0,tx1) x [ty0,ty1) */

/* cocogfx TFixed<F>(float): truncate(v * 2^F)  (third_party/cocogfx/include/fixed.hpp:97-100) */
static int32_t to_fixed(float v, float one) { return (int32_t)(v * one); }

typedef struct { float min_x, scale_x, min_y, scale_y, min_z, scale_z; } viewport_xf_t;

/* constants of ClipToHDC / NDCToScreen with (left, right, top, bottom) = (0, width, 0, height)
 * (third_party/cocogfx/include/math.hpp:1243-1250, :1267-1274) */
static viewport_xf_t make_xf(uint32_t width, uint32_t height, float near_, float far_) {
  viewport_xf_t x;
  x.min_x = (float)(int32_t)width / 2;
  x.scale_x = (float)(int32_t)width / 2;
  x.min_y = (float)(int32_t)height / 2;
  x.scale_y = (float)(int32_t)height / 2;
  x.min_z = (near_ + far_) / 2;
  x.scale_z = (far_ - near_) / 2;
  return x;
}

static vec4 load_pos(const srt_vertex_t* v) {
  vec4 p;
  p.x = v->pos[0]; p.y = v->pos[1]; p.z = v->pos[2]; p.w = v->pos[3];
  return p;
}

/* math.hpp:1234-1255 */
static vec4 to_hdc(vec4 c, const viewport_xf_t* x) {
  vec4 o;
  o.x = c.x * x->scale_x + c.w * x->min_x;
  o.y = c.y * x->scale_y + c.w * x->min_y;
  o.z = c.z * x->scale_z + c.w * x->min_z;
  o.w = c.w;
  return o;
}

/* math.hpp:1222-1230, :1258-1294 */
static vec4 to_screen(vec4 c, const viewport_xf_t* x) {
  float rhw = (c.w != 0.0f) ? (1.0f / c.w) : 0.0f;
  vec4 o;
  o.x = (c.x * rhw) * x->scale_x + x->min_x;
  o.y = (c.y * rhw) * x->scale_y + x->min_y;
  o.z = (c.z * rhw) * x->scale_z + x->min_z;
  o.w = rhw;
  return o;
}

static float max_f(float a, float b) { return a < b ? b : a; }
static float min_f(float a, float b) { return b < a ? b : a; }

static void plane(rast_attrib_t* d, float a0, float a1, float a2) { /* ATTRIBUTE_DELTA, gfxutil.cpp:189-192 */
  const float one = 16777216.0f;
  d->x = to_fixed(a0 - a2, one);
  d->y = to_fixed(a1 - a2, one);
  d->z = to_fixed(a2, one);
}

/* Triangle setup.  Returns 0 if the triangle is dropped. */
static int setup_prim(const srt_drawcall_t* draw, const srt_primitive_t* tri, const viewport_xf_t* xf, uint32_t width,
                      uint32_t height, uint32_t L, uint32_t tiles_x, uint32_t tiles_y, rast_prim_t* out, tile_rect_t* rect) {
  const srt_vertex_t *v0 = &draw->vertices[tri->i0], *v1 = &draw->vertices[tri->i1], *v2 = &draw->vertices[tri->i2];
  vec4 h0 = to_hdc(load_pos(v0), xf), h1 = to_hdc(load_pos(v1), xf), h2 = to_hdc(load_pos(v2), xf);
  vec4 s0, s1, s2;
  float e[3][3], det, m, scale;
  uint32_t left, right, top, bottom, tile = 1u << L;
  int i;
  /* edge equations in homogeneous 2D (gfxutil.cpp:35-75) */
  e[0][0] = (h1.y * h2.w) - (h2.y * h1.w);
  e[1][0] = (h2.y * h0.w) - (h0.y * h2.w);
  e[2][0] = (h0.y * h1.w) - (h1.y * h0.w);
  e[0][1] = (h2.x * h1.w) - (h1.x * h2.w);
  e[1][1] = (h0.x * h2.w) - (h2.x * h0.w);
  e[2][1] = (h1.x * h0.w) - (h0.x * h1.w);
  e[0][2] = (h1.x * h2.y) - (h2.x * h1.y);
  e[1][2] = (h2.x * h0.y) - (h0.x * h2.y);
  e[2][2] = (h0.x * h1.y) - (h1.x * h0.y);
  det = e[0][2] * h0.w + e[1][2] * h1.w + e[2][2] * h2.w;
  if (det == 0) return 0; /* degenerate, gfxutil.cpp:139-143 */
  if (det < 0)
    for (i = 0; i < 3; ++i) { e[i][0] *= -1.0f; e[i][1] *= -1.0f; e[i][2] *= -1.0f; }

  s0 = to_screen(load_pos(v0), xf); s1 = to_screen(load_pos(v1), xf); s2 = to_screen(load_pos(v2), xf);
  { /* bounding box clamped to the target (gfxutil.cpp:152-178); compared as uint32_t like rast_bbox_t */
    int l = (int)floorf(min_f(s0.x, min_f(s1.x, s2.x))), r = (int)ceilf(max_f(s0.x, max_f(s1.x, s2.x)));
    int t = (int)floorf(min_f(s0.y, min_f(s1.y, s2.y))), b = (int)ceilf(max_f(s0.y, max_f(s1.y, s2.y)));
    left = (uint32_t)(l > 0 ? l : 0);
    right = (uint32_t)(r < (int)width ? r : (int)width);
    top = (uint32_t)(t > 0 ? t : 0);
    bottom = (uint32_t)(b < (int)height ? b : (int)height);
    if (right <= left || bottom <= top) return 0;
  }
  for (i = 0; i < 3; ++i) e[i][2] += e[i][0] * 0.5f + e[i][1] * 0.5f; /* pixel centres, :195-198 */
  /* normalise by the largest |a|,|b| and convert to 16.16 (gfxutil.cpp:79-96) */
  m = fabsf(e[0][0]);
  m = max_f(m, fabsf(e[1][0])); m = max_f(m, fabsf(e[2][0]));
  m = max_f(m, fabsf(e[0][1])); m = max_f(m, fabsf(e[1][1])); m = max_f(m, fabsf(e[2][1]));
  scale = 1.0f / m;
  for (i = 0; i < 3; ++i) {
    out->edges[i].x = to_fixed(e[i][0] * scale, 65536.0f);
    out->edges[i].y = to_fixed(e[i][1] * scale, 65536.0f);
    out->edges[i].z = to_fixed(e[i][2] * scale, 65536.0f);
  }
  plane(&out->attribs.z, s0.z, s1.z, s2.z);
  plane(&out->attribs.r, v0->color[0], v1->color[0], v2->color[0]);
  plane(&out->attribs.g, v0->color[1], v1->color[1], v2->color[1]);
  plane(&out->attribs.b, v0->color[2], v1->color[2], v2->color[2]);
  plane(&out->attribs.a, v0->color[3], v1->color[3], v2->color[3]);
  plane(&out->attribs.u, v0->texcoord[0], v1->texcoord[0], v2->texcoord[0]);
  plane(&out->attribs.v, v0->texcoord[1], v1->texcoord[1], v2->texcoord[1]);
  { /* tiles touched (gfxutil.cpp:221-235).  A negative `right`/`bottom` (triangle wholly left of / above
       the target) passes the reference's unsigned reject test and makes its loop bounds wrap; such a
       triangle covers no pixel, so its tile range is clamped to the grid here: the primitive is kept in
       primbuf (ids stay the reference's) and lands in no tile that has a pixel. */
    uint32_t tx0 = left >> L, tx1 = (right + tile - 1) >> L, ty0 = top >> L, ty1 = (bottom + tile - 1) >> L;
    if ((int32_t)right < 0) tx1 = tx0;
    if ((int32_t)bottom < 0) ty1 = ty0;
    if (tx1 > tiles_x) tx1 = tiles_x;
    if (ty1 > tiles_y) ty1 = tiles_y;
    if (tx0 > tx1) tx0 = tx1;
    if (ty0 > ty1) ty0 = ty1;
    rect->tx0 = (uint16_t)tx0; rect->tx1 = (uint16_t)tx1; rect->ty0 = (uint16_t)ty0; rect->ty1 = (uint16_t)ty1;
  }
  return 1;
}

int srt_draw3d_binning(const srt_drawcall_t* draw, uint32_t width, uint32_t height, uint32_t L, uint32_t format,
                       srt_binning_t* out) {
  return srt_draw3d_binning_rank(draw, width, height, L, format, 0, 1, out);
}

/* Multi-GPU: tile rows are dealt round-robin, rank r keeps the tiles with tile_y % num_ranks == r
 * (tiles are disjoint by construction, gfxutil.cpp:237-250; the reference strides tiles over its own
 * cores the same way, gpu_sw.h:38).  primbuf is the full one on every rank, so pids mean the same. */
int srt_draw3d_binning_rank(const srt_drawcall_t* draw, uint32_t width, uint32_t height, uint32_t L, uint32_t format,
                            uint32_t rank, uint32_t num_ranks, srt_binning_t* out) {
  viewport_xf_t xf;
  uint32_t tiles_x, tiles_y, n, num_prims = 0, num_tiles = 0, tx, ty;
  uint64_t total = 0, ngrid, g;
  rast_prim_t* prims = NULL;
  tile_rect_t* rects = NULL;
  uint32_t *count = NULL, *cursor = NULL;
  uint8_t* tilebuf = NULL;
  size_t hdr_size = format == SRT_TILEBUF_WIDE ? sizeof(srt_tile_header_wide_t) : sizeof(rast_tile_header_t);
  int rc = -1;
  if (!draw || !out || width == 0 || height == 0 || L < 1 || L > 8 || width > 32768 || height > 32768 ||
      (format != SRT_TILEBUF_REFERENCE && format != SRT_TILEBUF_WIDE) || num_ranks == 0 || rank >= num_ranks)
    return -1;
  memset(out, 0, sizeof(*out));
  xf = make_xf(width, height, draw->viewport.near_, draw->viewport.far_);
  tiles_x = (width + (1u << L) - 1) >> L;
  tiles_y = (height + (1u << L) - 1) >> L;
  ngrid = (uint64_t)tiles_x * tiles_y;
  prims = (rast_prim_t*)malloc((draw->num_primitives ? draw->num_primitives : 1) * sizeof(rast_prim_t));
  rects = (tile_rect_t*)malloc((draw->num_primitives ? draw->num_primitives : 1) * sizeof(tile_rect_t));
  count = (uint32_t*)calloc(ngrid + 1, sizeof(uint32_t)); /* grid index = tx * tiles_y + ty: emission order */
  if (!prims || !rects || !count) goto done;

  /* pass 1: setup + per-tile counts */
  for (n = 0; n < draw->num_primitives; ++n) {
    if (!setup_prim(draw, &draw->primitives[n], &xf, width, height, L, tiles_x, tiles_y, &prims[num_prims], &rects[num_prims]))
      continue;
    for (tx = rects[num_prims].tx0; tx < rects[num_prims].tx1; ++tx)
      for (ty = rects[num_prims].ty0; ty < rects[num_prims].ty1; ++ty)
        if (ty % num_ranks == rank) ++count[(uint64_t)tx * tiles_y + ty];
    ++num_prims;
  }
  /* exclusive scan -> start of every tile's id list */
  cursor = (uint32_t*)malloc((ngrid + 1) * sizeof(uint32_t));
  if (!cursor) goto done;
  for (g = 0; g < ngrid; ++g) {
    cursor[g] = (uint32_t)total;
    total += count[g];
    if (count[g]) ++num_tiles;
  }
  if (total > 0xffffffffull) goto done;
  tilebuf = (uint8_t*)calloc(num_tiles * hdr_size + total * 4 + 1, 1);
  if (!tilebuf) goto done;
  { /* headers, x-major like std::map<pair<tile_x, tile_y>> iteration (gfxutil.cpp:243-262) */
    uint32_t* pids = (uint32_t*)(tilebuf + num_tiles * hdr_size);
    uint32_t t = 0;
    for (tx = 0; tx < tiles_x; ++tx)
      for (ty = 0; ty < tiles_y; ++ty) {
        g = (uint64_t)tx * tiles_y + ty;
        if (!count[g]) continue;
        if (format == SRT_TILEBUF_WIDE) {
          srt_tile_header_wide_t* h = (srt_tile_header_wide_t*)tilebuf + t;
          h->tile_x = (uint16_t)tx; h->tile_y = (uint16_t)ty;
          h->pids_count = count[g]; h->pids_offset = cursor[g]; h->reserved = 0;
        } else {
          rast_tile_header_t* h = (rast_tile_header_t*)tilebuf + t;
          /* offset is counted in u32s from the end of this header (gfxutil.cpp:252) */
          uint64_t off = (uint64_t)(num_tiles - 1 - t) * (sizeof(rast_tile_header_t) / 4) + cursor[g];
          if (off > 0xffff || count[g] > 0xffff) { rc = -2; goto done; }
          h->tile_x = (uint16_t)tx; h->tile_y = (uint16_t)ty;
          h->pids_offset = (uint16_t)off; h->pids_count = (uint16_t)count[g];
        }
        ++t;
      }
    /* pass 2: fill, primitives in submission order */
    for (n = 0; n < num_prims; ++n)
      for (tx = rects[n].tx0; tx < rects[n].tx1; ++tx)
        for (ty = rects[n].ty0; ty < rects[n].ty1; ++ty)
          if (ty % num_ranks == rank) pids[cursor[(uint64_t)tx * tiles_y + ty]++] = n;
  }
  out->tilebuf = tilebuf; tilebuf = NULL;
  out->tilebuf_size = num_tiles * hdr_size + total * 4;
  out->primbuf = (uint8_t*)prims; prims = NULL;
  out->primbuf_size = (uint64_t)num_prims * sizeof(rast_prim_t);
  out->num_tiles = num_tiles;
  out->num_prims = num_prims;
  out->num_tile_prims = total;
  out->format = format;
  rc = 0;
done:
  free(prims); free(rects); free(count); free(cursor); free(tilebuf);
  return rc;
}

void srt_binning_free(srt_binning_t* b) {
  if (!b) return;
  free(b->tilebuf);
  free(b->primbuf);
  memset(b, 0, sizeof(*b));
}
