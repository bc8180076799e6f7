/*
 * ref_draw3d.c -- CPU oracle for the raster hot path.  TEST INFRASTRUCTURE ONLY (see ref_draw3d.h).
 *
 * Plain-C restatement; every function cites the reference code it follows.  Paths are relative to
 * /root/reference.  Compile with -ffp-contract=off: the float expressions below must round exactly
 * like the reference's x86-64 build (no FMA).
 */
#include "ref_draw3d.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------------------------------------
 * Trace blob ("SKBXTRC1", tests/golden/make_raster_fixtures.py).  I/O only, no reference counterpart
 * (the reference reads Boost XML, third_party/cocogfx/src/cgltrace.cpp:146-165).
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  srt_states_t st;
  uint32_t texture_id;
  srt_viewport_t vp;
  uint32_t nv, np;
  const uint8_t* verts; /* nv * 40 B, unaligned */
  const uint8_t* prims; /* np * 12 B            */
} o_draw_t;

typedef struct {
  uint32_t id, format, width, height, size;
  const uint8_t* pixels;
} o_tex_t;

typedef struct {
  uint32_t nd, nt;
  o_draw_t* draws;
  o_tex_t* texs;
} o_trace_t;

static uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
static float rdf(const uint8_t* p) { float v; memcpy(&v, p, 4); return v; }

static int o_parse(o_trace_t* t, const uint8_t* blob, uint64_t size) {
  const uint8_t *p = blob, *end = blob + size;
  uint32_t i, k;
  memset(t, 0, sizeof(*t));
  if (size < 16 || memcmp(p, "SKBXTRC1", 8) != 0) return -1;
  t->nd = rd32(p + 8);
  t->nt = rd32(p + 12);
  p += 16;
  t->draws = (o_draw_t*)calloc(t->nd ? t->nd : 1, sizeof(o_draw_t));
  t->texs = (o_tex_t*)calloc(t->nt ? t->nt : 1, sizeof(o_tex_t));
  for (i = 0; i < t->nt; ++i) {
    if (p + 20 > end) return -1;
    t->texs[i].id = rd32(p); t->texs[i].format = rd32(p + 4); t->texs[i].width = rd32(p + 8);
    t->texs[i].height = rd32(p + 12); t->texs[i].size = rd32(p + 16);
    t->texs[i].pixels = p + 20;
    p += 20 + ((t->texs[i].size + 3) & ~3u);
    if (p > end) return -1;
  }
  for (i = 0; i < t->nd; ++i) {
    o_draw_t* d = &t->draws[i];
    uint32_t* f = (uint32_t*)&d->st;
    if (p + 24 * 4 + 16 + 4 + 24 + 8 > end) return -1;
    for (k = 0; k < 24; ++k) f[k] = rd32(p + 4 * k);
    p += 96;
    for (k = 0; k < 4; ++k) d->st.texture_envcolor[k] = rdf(p + 4 * k);
    p += 16;
    d->texture_id = rd32(p); p += 4;
    d->vp.left = (int32_t)rd32(p); d->vp.right = (int32_t)rd32(p + 4);
    d->vp.top = (int32_t)rd32(p + 8); d->vp.bottom = (int32_t)rd32(p + 12);
    d->vp.near_ = rdf(p + 16); d->vp.far_ = rdf(p + 20);
    p += 24;
    d->nv = rd32(p); d->np = rd32(p + 4); p += 8;
    d->verts = p; p += (uint64_t)d->nv * 40;
    d->prims = p; p += (uint64_t)d->np * 12;
    if (p > end) return -1;
  }
  return 0;
}

static void o_free(o_trace_t* t) { free(t->draws); free(t->texs); }

/* ------------------------------------------------------------------------------------------------
 * Host pre-pass
 * ---------------------------------------------------------------------------------------------- */
typedef struct { float x, y, z, w; } v4;
typedef struct { float x, y, z; } v3;

/* third_party/cocogfx/include/fixed.hpp:97-100 : TFixed(float) = static_cast<int32_t>(rhs * ONE) */
static int32_t fx16(float v) { return (int32_t)(v * 65536.0f); }
static int32_t fx24(float v) { return (int32_t)(v * 16777216.0f); }

/* third_party/cocogfx/include/math.hpp:1234-1255 */
static void clip_to_hdc(v4* o, v4 in, uint32_t width, uint32_t height, float near_, float far_) {
  float minX = (float)(int32_t)(0 + width) / 2, scaleX = (float)(int32_t)(width - 0) / 2;
  float minY = (float)(int32_t)(0 + height) / 2, scaleY = (float)(int32_t)(height - 0) / 2;
  float minZ = (near_ + far_) / 2, scaleZ = (far_ - near_) / 2;
  o->x = in.x * scaleX + in.w * minX;
  o->y = in.y * scaleY + in.w * minY;
  o->z = in.z * scaleZ + in.w * minZ;
  o->w = in.w;
}

/* math.hpp:1222-1230 (ClipToNDC), :1258-1279 (NDCToScreen), :1282-1294 (ClipToScreen) */
static void clip_to_screen(v4* o, v4 in, uint32_t width, uint32_t height, float near_, float far_) {
  float rhw = (in.w != 0.0f) ? (1.0f / in.w) : 0.0f;
  float nx = in.x * rhw, ny = in.y * rhw, nz = in.z * rhw;
  float minX = (float)(int32_t)(0 + width) / 2, scaleX = (float)(int32_t)(width - 0) / 2;
  float minY = (float)(int32_t)(0 + height) / 2, scaleY = (float)(int32_t)(height - 0) / 2;
  float minZ = (near_ + far_) / 2, scaleZ = (far_ - near_) / 2;
  o->x = nx * scaleX + minX;
  o->y = ny * scaleY + minY;
  o->z = nz * scaleZ + minZ;
  o->w = rhw;
}

/* sim/common/gfxutil.cpp:35-75 */
static int edge_equation(v3 e[3], v4 v0, v4 v1, v4 v2) {
  float a0 = (v1.y * v2.w) - (v2.y * v1.w);
  float a1 = (v2.y * v0.w) - (v0.y * v2.w);
  float a2 = (v0.y * v1.w) - (v1.y * v0.w);
  float b0 = (v2.x * v1.w) - (v1.x * v2.w);
  float b1 = (v0.x * v2.w) - (v2.x * v0.w);
  float b2 = (v1.x * v0.w) - (v0.x * v1.w);
  float c0 = (v1.x * v2.y) - (v2.x * v1.y);
  float c1 = (v2.x * v0.y) - (v0.x * v2.y);
  float c2 = (v0.x * v1.y) - (v1.x * v0.y);
  float det;
  int i;
  e[0].x = a0; e[0].y = b0; e[0].z = c0;
  e[1].x = a1; e[1].y = b1; e[1].z = c1;
  e[2].x = a2; e[2].y = b2; e[2].z = c2;
  det = c0 * v0.w + c1 * v1.w + c2 * v2.w;
  if (det < 0)
    for (i = 0; i < 3; ++i) { e[i].x *= -1.0f; e[i].y *= -1.0f; e[i].z *= -1.0f; }
  return det != 0;
}

static float fmax2(float a, float b) { return a < b ? b : a; } /* std::max */
static float fmin2(float a, float b) { return b < a ? b : a; } /* std::min */

/* sim/common/gfxutil.cpp:79-96 */
static void edge_to_fixed(rast_vec3e_t out[3], const v3 in[3]) {
  float m = fabsf(in[0].x), scale;
  int i;
  m = fmax2(m, fabsf(in[1].x)); m = fmax2(m, fabsf(in[2].x));
  m = fmax2(m, fabsf(in[0].y)); m = fmax2(m, fabsf(in[1].y)); m = fmax2(m, fabsf(in[2].y));
  scale = 1.0f / m;
  for (i = 0; i < 3; ++i) {
    out[i].x = fx16(in[i].x * scale);
    out[i].y = fx16(in[i].y * scale);
    out[i].z = fx16(in[i].z * scale);
  }
}

typedef struct { uint32_t key, pid; } o_pair_t; /* key = tile_x << 16 | tile_y : std::map<pair> order */

typedef struct {
  rast_prim_t* prims; uint32_t num_prims;
  uint8_t* tilebuf; uint64_t tilebuf_size; uint32_t num_tiles;
  /* the same tiles without the 16-bit limits of rast_tile_header_t (pids_offset / pids_count wrap beyond
   * 64 K entries, graphics.h:66-71): what the oracle's own kernel walks, so that it stays usable as a
   * checker at frame sizes the reference's buffer format cannot describe */
  uint32_t* tile_xy;    /* tile_x << 16 | tile_y, per tile */
  uint32_t* tile_start; /* num_tiles + 1 offsets into pid_list */
  uint32_t* pid_list;
} o_bins_t;

static void o_bins_free(o_bins_t* b) {
  free(b->tilebuf); free(b->prims); free(b->tile_xy); free(b->tile_start); free(b->pid_list);
}

static void load_vertex(const o_draw_t* d, uint32_t i, v4* pos, float col[4], float uv[2]) {
  const uint8_t* p = d->verts + (uint64_t)i * 40;
  int k;
  pos->x = rdf(p); pos->y = rdf(p + 4); pos->z = rdf(p + 8); pos->w = rdf(p + 12);
  for (k = 0; k < 4; ++k) col[k] = rdf(p + 16 + 4 * k);
  uv[0] = rdf(p + 32); uv[1] = rdf(p + 36);
}

#define ATTR_DELTA(d, x0, x1, x2) do { (d).x = fx24((x0) - (x2)); (d).y = fx24((x1) - (x2)); (d).z = fx24(x2); } while (0)

/* sim/common/gfxutil.cpp:103-276 graphics::Binning */
static int o_binning(o_bins_t* out, const o_draw_t* d, uint32_t width, uint32_t height, uint32_t L) {
  uint32_t n, npairs = 0, cap = 1024, i;
  o_pair_t* pairs = (o_pair_t*)malloc(cap * sizeof(o_pair_t));
  rast_prim_t* prims = (rast_prim_t*)malloc((d->np ? d->np : 1) * sizeof(rast_prim_t));
  uint32_t num_prims = 0;
  memset(out, 0, sizeof(*out));
  for (n = 0; n < d->np; ++n) {
    v4 p0, p1, p2, ph0, ph1, ph2, ps0, ps1, ps2;
    float c0[4], c1[4], c2[4], t0[2], t1[2], t2[2];
    v3 edges[3];
    int32_t left, right, top, bottom;
    rast_prim_t rp;
    uint32_t p, tx, ty, minTX, maxTX, minTY, maxTY, tileSize = 1u << L;
    load_vertex(d, rd32(d->prims + 12 * (uint64_t)n), &p0, c0, t0);
    load_vertex(d, rd32(d->prims + 12 * (uint64_t)n + 4), &p1, c1, t1);
    load_vertex(d, rd32(d->prims + 12 * (uint64_t)n + 8), &p2, c2, t2);
    clip_to_hdc(&ph0, p0, width, height, d->vp.near_, d->vp.far_);
    clip_to_hdc(&ph1, p1, width, height, d->vp.near_, d->vp.far_);
    clip_to_hdc(&ph2, p2, width, height, d->vp.near_, d->vp.far_);
    if (!edge_equation(edges, ph0, ph1, ph2)) continue; /* degenerate, :139-143 */
    clip_to_screen(&ps0, p0, width, height, d->vp.near_, d->vp.far_);
    clip_to_screen(&ps1, p1, width, height, d->vp.near_, d->vp.far_);
    clip_to_screen(&ps2, p2, width, height, d->vp.near_, d->vp.far_);
    { /* :152-184 bounding box (math.hpp:1299-1310), clamp to the render target, reject */
      float l = fmin2(ps0.x, fmin2(ps1.x, ps2.x)), r = fmax2(ps0.x, fmax2(ps1.x, ps2.x));
      float t = fmin2(ps0.y, fmin2(ps1.y, ps2.y)), b = fmax2(ps0.y, fmax2(ps1.y, ps2.y));
      int tl = (int)floorf(l), tr = (int)ceilf(r), tt = (int)floorf(t), tb = (int)ceilf(b);
      left = tl > 0 ? tl : 0;
      right = tr < (int32_t)width ? tr : (int32_t)width;
      top = tt > 0 ? tt : 0;
      bottom = tb < (int32_t)height ? tb : (int32_t)height;
      /* the reference compares these as uint32_t (rast_bbox_t, graphics.h:52-57) */
      if ((uint32_t)right <= (uint32_t)left || (uint32_t)bottom <= (uint32_t)top) continue;
    }
    /* :195-198 half-pixel offset */
    edges[0].z += edges[0].x * 0.5f + edges[0].y * 0.5f;
    edges[1].z += edges[1].x * 0.5f + edges[1].y * 0.5f;
    edges[2].z += edges[2].x * 0.5f + edges[2].y * 0.5f;
    edge_to_fixed(rp.edges, edges);
    ATTR_DELTA(rp.attribs.z, ps0.z, ps1.z, ps2.z);
    ATTR_DELTA(rp.attribs.r, c0[0], c1[0], c2[0]);
    ATTR_DELTA(rp.attribs.g, c0[1], c1[1], c2[1]);
    ATTR_DELTA(rp.attribs.b, c0[2], c1[2], c2[2]);
    ATTR_DELTA(rp.attribs.a, c0[3], c1[3], c2[3]);
    ATTR_DELTA(rp.attribs.u, t0[0], t1[0], t2[0]);
    ATTR_DELTA(rp.attribs.v, t0[1], t1[1], t2[1]);
    p = num_prims;
    prims[num_prims++] = rp;
    /* :221-235 tiles the bounding box touches */
    minTX = (uint32_t)left >> L; maxTX = ((uint32_t)right + tileSize - 1) >> L;
    minTY = (uint32_t)top >> L;  maxTY = ((uint32_t)bottom + tileSize - 1) >> L;
    for (ty = minTY; ty < maxTY; ++ty)
      for (tx = minTX; tx < maxTX; ++tx) {
        if (npairs == cap) { cap *= 2; pairs = (o_pair_t*)realloc(pairs, cap * sizeof(o_pair_t)); }
        pairs[npairs].key = ((tx & 0xffff) << 16) | (ty & 0xffff); /* map key is pair<uint16,uint16> */
        pairs[npairs].pid = p;
        ++npairs;
      }
  }
  { /* :243-262 emit headers in map order, then the pid lists; stable counting pass per key */
    uint32_t *order = (uint32_t*)malloc((npairs ? npairs : 1) * sizeof(uint32_t)), ntiles = 0, j;
    o_pair_t* sorted = (o_pair_t*)malloc((npairs ? npairs : 1) * sizeof(o_pair_t));
    /* stable LSD radix sort on the 32-bit key, 4 passes of 8 bits */
    o_pair_t *src = pairs, *dst = sorted, *tmp;
    int pass;
    for (pass = 0; pass < 4; ++pass) {
      uint32_t cnt[257] = {0};
      for (i = 0; i < npairs; ++i) cnt[((src[i].key >> (8 * pass)) & 0xff) + 1]++;
      for (j = 0; j < 256; ++j) cnt[j + 1] += cnt[j];
      for (i = 0; i < npairs; ++i) dst[cnt[(src[i].key >> (8 * pass)) & 0xff]++] = src[i];
      tmp = src; src = dst; dst = tmp;
    }
    /* after 4 passes the sorted data is back in `pairs` (src) */
    for (i = 0; i < npairs; ++i)
      if (i == 0 || src[i].key != src[i - 1].key) ++ntiles;
    out->tilebuf_size = (uint64_t)ntiles * sizeof(rast_tile_header_t) + (uint64_t)npairs * 4;
    out->tilebuf = (uint8_t*)calloc(out->tilebuf_size ? out->tilebuf_size : 1, 1);
    {
      rast_tile_header_t* hdr = (rast_tile_header_t*)out->tilebuf;
      uint32_t* pids = (uint32_t*)(out->tilebuf + (uint64_t)ntiles * sizeof(rast_tile_header_t));
      uint32_t t = 0, start = 0;
      for (i = 0; i <= npairs; ++i) {
        if (i == npairs || (i > 0 && src[i].key != src[i - 1].key)) {
          if (i > start) {
            rast_tile_header_t h;
            h.tile_x = (uint16_t)(src[start].key >> 16);
            h.tile_y = (uint16_t)(src[start].key & 0xffff);
            /* :252  (pids_buffer - (tile_header + 1)) / 4, truncated to 16 bits */
            h.pids_offset = (uint16_t)((((uint8_t*)(pids + start)) - ((uint8_t*)(hdr + t + 1))) / 4);
            h.pids_count = (uint16_t)(i - start);
            hdr[t++] = h;
          }
          start = i;
        }
        if (i < npairs) pids[i] = src[i].pid;
      }
    }
    out->num_tiles = ntiles;
    out->tile_xy = (uint32_t*)malloc((ntiles + 1) * sizeof(uint32_t));
    out->tile_start = (uint32_t*)malloc((ntiles + 1) * sizeof(uint32_t));
    out->pid_list = (uint32_t*)malloc((npairs ? npairs : 1) * sizeof(uint32_t));
    {
      uint32_t t = 0;
      for (i = 0; i < npairs; ++i) {
        if (i == 0 || src[i].key != src[i - 1].key) { out->tile_xy[t] = src[i].key; out->tile_start[t] = i; ++t; }
        out->pid_list[i] = src[i].pid;
      }
      out->tile_start[t] = npairs;
    }
    free(order);
    if (src == pairs) { free(sorted); } else { free(pairs); pairs = sorted; }
  }
  free(pairs);
  out->prims = prims;
  out->num_prims = num_prims;
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Device side.  State = the kernel argument (DCR mirrors) + host pointers standing in for the
 * addresses the DCRs carry (gpu_sw.h:28-29, :83-89 rebuild pointers as dcr << 6).
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  draw3d_kernel_arg_t arg;
  const uint32_t *tile_xy, *tile_start, *pid_list; /* see o_bins_t */
  const uint8_t* tilebuf;
  const uint8_t* primbuf;
  const uint8_t* texbuf;
  uint8_t* zbuf;
  uint8_t* cbuf;
} o_dev_t;

#define RAS(a, r) DRAW3D_RASTER_DCR(a, r)
#define OM(a, r) DRAW3D_OM_DCR(a, r)
#define TEX(a, r) DRAW3D_TEX_DCR(a, 0, r)

/* ---- texture unit: sim/common/graphics.cpp:35-314 -------------------------------------------- */
#define TEX_MASK ((1 << VX_TEX_FXD_FRAC) - 1)

/* graphics.cpp:35-54 TextureWrap<23> */
static int32_t tex_wrap(int32_t d, uint32_t wrap) {
  int32_t ret;
  switch (wrap) {
  default:
  case VX_TEX_WRAP_CLAMP:
    ret = d & -(d >= 0);
    ret |= ((TEX_MASK - ret) >> 31);
    break;
  case VX_TEX_WRAP_REPEAT:
    ret = d;
    break;
  case VX_TEX_WRAP_MIRROR:
    ret = d ^ ((int32_t)((uint32_t)d << (31 - VX_TEX_FXD_FRAC)) >> 31);
    break;
  }
  return ret & TEX_MASK;
}

/* graphics.cpp:56-71 */
static uint32_t format_stride(uint32_t format) {
  switch (format) {
  default:
  case VX_TEX_FORMAT_A8R8G8B8: return 4;
  case VX_TEX_FORMAT_R5G6B5:
  case VX_TEX_FORMAT_A1R5G5B5:
  case VX_TEX_FORMAT_A4R4G4B4:
  case VX_TEX_FORMAT_A8L8: return 2;
  case VX_TEX_FORMAT_L8:
  case VX_TEX_FORMAT_A8: return 1;
  }
}

/* graphics.cpp:73-124 */
static void unpack8888(uint32_t format, uint32_t texel, uint32_t* lo, uint32_t* hi) {
  uint32_t r, g, b, a;
  switch (format) {
  default:
  case VX_TEX_FORMAT_A8R8G8B8:
    r = (texel >> 16) & 0xff; g = (texel >> 8) & 0xff; b = texel & 0xff; a = texel >> 24;
    break;
  case VX_TEX_FORMAT_R5G6B5:
    r = ((texel >> 8) & 0xf8) | ((texel >> 13) & 0x07);
    g = ((texel >> 3) & 0xfc) | ((texel >> 9) & 0x03);
    b = ((texel << 3) & 0xf8) | ((texel >> 2) & 0x07);
    a = 0xff;
    break;
  case VX_TEX_FORMAT_A1R5G5B5:
    r = ((texel >> 7) & 0xf8) | ((texel >> 12) & 0x07);
    g = ((texel >> 2) & 0xf8) | ((texel >> 7) & 0x07);
    b = ((texel << 3) & 0xf8) | ((texel >> 2) & 0x07);
    a = (uint32_t)(((int32_t)(texel << 16)) >> 31) & 0xff;
    break;
  case VX_TEX_FORMAT_A4R4G4B4:
    r = ((texel >> 4) & 0xf0) | ((texel >> 8) & 0x0f);
    g = ((texel >> 0) & 0xf0) | ((texel >> 4) & 0x0f);
    b = ((texel << 4) & 0xf0) | ((texel >> 0) & 0x0f);
    a = ((texel >> 8) & 0xf0) | ((texel >> 12) & 0x0f);
    break;
  case VX_TEX_FORMAT_A8L8:
    r = texel & 0xff; g = r; b = r; a = (texel >> 8) & 0xff;
    break;
  case VX_TEX_FORMAT_L8:
    r = texel & 0xff; g = r; b = r; a = 0xff;
    break;
  case VX_TEX_FORMAT_A8:
    r = 0xff; g = 0xff; b = 0xff; a = texel & 0xff;
    break;
  }
  *lo = (r << 16) + b;
  *hi = (a << 16) + g;
}

/* sim/common/graphics.h:82-86 */
static uint32_t lerp8888(uint32_t a, uint32_t b, uint32_t f) {
  uint32_t p = a * (0xff - f) + b * f + 0x00800080;
  uint32_t q = (p >> 8) & 0x00ff00ff;
  return ((p + q) >> 8) & 0x00ff00ff;
}

static uint32_t fetch_texel(const uint8_t* base, uint32_t offset, uint32_t stride) {
  /* gpu_sw.h:183-205 memory_cb */
  const uint8_t* p = base + (uint64_t)offset * stride;
  switch (stride) {
  case 4: { uint32_t v; memcpy(&v, p, 4); return v; }
  case 2: { uint16_t v; memcpy(&v, p, 2); return v; }
  default: return *p;
  }
}

/* graphics.cpp:253-314 TextureSampler::read, stage 0 */
static uint32_t tex_read(const draw3d_kernel_arg_t* a, const uint8_t* texbuf, int32_t u, int32_t v, uint32_t lod) {
  uint32_t mip_off = TEX(a, VX_DCR_TEX_MIPOFF(lod));
  uint32_t logdim = TEX(a, VX_DCR_TEX_LOGDIM), format = TEX(a, VX_DCR_TEX_FORMAT);
  uint32_t filter = TEX(a, VX_DCR_TEX_FILTER), wrap = TEX(a, VX_DCR_TEX_WRAP);
  const uint8_t* base = texbuf + mip_off;
  int32_t lw = (int32_t)(logdim & 0xffff) - (int32_t)lod, lh = (int32_t)(logdim >> 16) - (int32_t)lod;
  uint32_t log_width = lw > 0 ? lw : 0, log_height = lh > 0 ? lh : 0;
  uint32_t wrapu = wrap & 0xffff, wrapv = wrap >> 16, stride = format_stride(format);
  if (filter == VX_TEX_FILTER_BILINEAR) {
    /* graphics.cpp:126-167 TexAddressLinear */
    int32_t dx = (1 << (VX_TEX_FXD_FRAC - 1)) >> log_width, dy = (1 << (VX_TEX_FXD_FRAC - 1)) >> log_height;
    uint32_t u0 = tex_wrap((int32_t)((uint32_t)u - (uint32_t)dx), wrapu), u1 = tex_wrap((int32_t)((uint32_t)u + (uint32_t)dx), wrapu);
    uint32_t v0 = tex_wrap((int32_t)((uint32_t)v - (uint32_t)dy), wrapv), v1 = tex_wrap((int32_t)((uint32_t)v + (uint32_t)dy), wrapv);
    uint32_t shift_u = VX_TEX_FXD_FRAC - log_width, shift_v = VX_TEX_FXD_FRAC - log_height;
    uint32_t x0s = (u0 << 8) >> shift_u, y0s = (v0 << 8) >> shift_v;
    uint32_t x0 = x0s >> 8, y0 = y0s >> 8, x1 = u1 >> shift_u, y1 = v1 >> shift_v;
    uint32_t alpha = x0s & 0xff, beta = y0s & 0xff;
    uint32_t t00 = fetch_texel(base, x0 + (y0 << log_width), stride);
    uint32_t t01 = fetch_texel(base, x1 + (y0 << log_width), stride);
    uint32_t t10 = fetch_texel(base, x0 + (y1 << log_width), stride);
    uint32_t t11 = fetch_texel(base, x1 + (y1 << log_width), stride);
    /* graphics.cpp:188-226 TexFilterLinear */
    uint32_t c0l, c0h, c1l, c1h, c2l, c2h, c3l, c3h, c01l, c01h, c23l, c23h, cl, ch;
    unpack8888(format, t00, &c0l, &c0h);
    unpack8888(format, t01, &c1l, &c1h);
    c01l = lerp8888(c0l, c1l, alpha); c01h = lerp8888(c0h, c1h, alpha);
    unpack8888(format, t10, &c2l, &c2h);
    unpack8888(format, t11, &c3l, &c3h);
    c23l = lerp8888(c2l, c3l, alpha); c23h = lerp8888(c2h, c3h, alpha);
    cl = lerp8888(c01l, c23l, beta); ch = lerp8888(c01h, c23h, beta);
    return (ch << 8) | cl;
  } else {
    /* graphics.cpp:169-186 TexAddressPoint, :228-239 TexFilterPoint */
    uint32_t uu = tex_wrap(u, wrapu), vv = tex_wrap(v, wrapv);
    uint32_t x = uu >> (VX_TEX_FXD_FRAC - log_width), y = vv >> (VX_TEX_FXD_FRAC - log_height);
    uint32_t cl, ch;
    unpack8888(format, fetch_texel(base, x + (y << log_width), stride), &cl, &ch);
    return (ch << 8) | cl;
  }
}

/* ---- depth / stencil: graphics.cpp:320-364, :533-596 ------------------------------------------ */
static int do_compare(uint32_t func, uint32_t a, uint32_t b) {
  switch (func) {
  default:
  case VX_OM_DEPTH_FUNC_NEVER: return 0;
  case VX_OM_DEPTH_FUNC_LESS: return a < b;
  case VX_OM_DEPTH_FUNC_EQUAL: return a == b;
  case VX_OM_DEPTH_FUNC_LEQUAL: return a <= b;
  case VX_OM_DEPTH_FUNC_GREATER: return a > b;
  case VX_OM_DEPTH_FUNC_NOTEQUAL: return a != b;
  case VX_OM_DEPTH_FUNC_GEQUAL: return a >= b;
  case VX_OM_DEPTH_FUNC_ALWAYS: return 1;
  }
}

static uint32_t do_stencil_op(uint32_t op, uint32_t ref, uint32_t val) {
  switch (op) {
  default:
  case VX_OM_STENCIL_OP_KEEP: return val;
  case VX_OM_STENCIL_OP_ZERO: return 0;
  case VX_OM_STENCIL_OP_REPLACE: return ref;
  case VX_OM_STENCIL_OP_INCR: return (val < 0xff) ? (val + 1) : val;
  case VX_OM_STENCIL_OP_DECR: return (val > 0) ? (val - 1) : val;
  case VX_OM_STENCIL_OP_INVERT: return ~val;
  case VX_OM_STENCIL_OP_INCR_WRAP: return (val + 1) & 0xff;
  case VX_OM_STENCIL_OP_DECR_WRAP: return (val - 1) & 0xff;
  }
}

typedef struct {
  uint32_t depth_func, s_func, s_zpass, s_zfail, s_fail, s_mask, s_ref; /* front face only: draw3d passes face 0 */
  int depth_enabled, stencil_enabled;
} o_ds_t;

/* graphics.cpp:533-562 DepthTencil::configure */
static void ds_configure(o_ds_t* s, const draw3d_kernel_arg_t* a) {
  int depth_writemask = OM(a, VX_DCR_OM_DEPTH_WRITEMASK) & 0x1;
  s->depth_func = OM(a, VX_DCR_OM_DEPTH_FUNC);
  s->s_func = OM(a, VX_DCR_OM_STENCIL_FUNC) & 0xffff;
  s->s_zpass = OM(a, VX_DCR_OM_STENCIL_ZPASS) & 0xffff;
  s->s_zfail = OM(a, VX_DCR_OM_STENCIL_ZFAIL) & 0xffff;
  s->s_fail = OM(a, VX_DCR_OM_STENCIL_FAIL) & 0xffff;
  s->s_ref = OM(a, VX_DCR_OM_STENCIL_REF) & 0xffff;
  s->s_mask = OM(a, VX_DCR_OM_STENCIL_MASK) & 0xffff;
  s->depth_enabled = !((s->depth_func == VX_OM_DEPTH_FUNC_ALWAYS) && !depth_writemask);
  s->stencil_enabled = !((s->s_func == VX_OM_DEPTH_FUNC_ALWAYS) && (s->s_zpass == VX_OM_STENCIL_OP_KEEP) &&
                         (s->s_zfail == VX_OM_STENCIL_OP_KEEP));
}

/* graphics.cpp:564-596 DepthTencil::test (is_backface = 0) */
static int ds_test(const o_ds_t* s, uint32_t depth, uint32_t ds_val, uint32_t* ds_result) {
  uint32_t depth_val = ds_val & VX_OM_DEPTH_MASK, stencil_val = ds_val >> VX_OM_DEPTH_BITS;
  uint32_t depth_ref = depth & VX_OM_DEPTH_MASK;
  uint32_t ref_m = s->s_ref & s->s_mask, val_m = stencil_val & s->s_mask, op;
  int passed = do_compare(s->s_func, ref_m, val_m);
  if (passed) {
    passed = do_compare(s->depth_func, depth_ref, depth_val);
    op = passed ? s->s_zpass : s->s_zfail;
  } else {
    op = s->s_fail;
  }
  *ds_result = (do_stencil_op(op, s->s_ref, stencil_val) << VX_OM_DEPTH_BITS) | depth_ref;
  return passed;
}

/* ---- blender: graphics.cpp:366-524, :600-636 --------------------------------------------------- */
typedef struct { int a, r, g, b; } o_col_t;
static o_col_t col_from(uint32_t v) { o_col_t c = {(int)(v >> 24), (int)((v >> 16) & 0xff), (int)((v >> 8) & 0xff), (int)(v & 0xff)}; return c; }
static uint32_t col_value(o_col_t c) { return ((uint32_t)(c.a & 0xff) << 24) | ((uint32_t)(c.r & 0xff) << 16) | ((uint32_t)(c.g & 0xff) << 8) | (uint32_t)(c.b & 0xff); }
static o_col_t col4(int a, int r, int g, int b) { o_col_t c = {a, r, g, b}; return c; }

/* graphics.cpp:366-405 */
static uint32_t do_logic_op(uint32_t op, uint32_t src, uint32_t dst) {
  switch (op) {
  default:
  case 0: return 0;
  case 1: return src & dst;
  case 2: return src & ~dst;
  case 3: return src;
  case 4: return ~src & dst;
  case 5: return dst;
  case 6: return src ^ dst;
  case 7: return src | dst;
  case 8: return ~(src | dst);
  case 9: return ~(src ^ dst);
  case 10: return ~dst;
  case 11: return src | ~dst;
  case 12: return ~src;
  case 13: return ~src | dst;
  case 14: return ~(src & dst);
  case 15: return 0xffffffff;
  }
}

/* graphics.cpp:407-478 (ONE_MINUS_CONST_A returns the per-channel complement, as the reference does) */
static o_col_t do_blend_func(uint32_t func, o_col_t src, o_col_t dst, o_col_t cst) {
  switch (func) {
  default:
  case VX_OM_BLEND_FUNC_ZERO: return col4(0, 0, 0, 0);
  case VX_OM_BLEND_FUNC_ONE: return col4(0xff, 0xff, 0xff, 0xff);
  case VX_OM_BLEND_FUNC_SRC_RGB: return src;
  case VX_OM_BLEND_FUNC_ONE_MINUS_SRC_RGB: return col4(0xff - src.a, 0xff - src.r, 0xff - src.g, 0xff - src.b);
  case VX_OM_BLEND_FUNC_DST_RGB: return dst;
  case VX_OM_BLEND_FUNC_ONE_MINUS_DST_RGB: return col4(0xff - dst.a, 0xff - dst.r, 0xff - dst.g, 0xff - dst.b);
  case VX_OM_BLEND_FUNC_SRC_A: return col4(src.a, src.a, src.a, src.a);
  case VX_OM_BLEND_FUNC_ONE_MINUS_SRC_A: return col4(0xff - src.a, 0xff - src.a, 0xff - src.a, 0xff - src.a);
  case VX_OM_BLEND_FUNC_DST_A: return col4(dst.a, dst.a, dst.a, dst.a);
  case VX_OM_BLEND_FUNC_ONE_MINUS_DST_A: return col4(0xff - dst.a, 0xff - dst.a, 0xff - dst.a, 0xff - dst.a);
  case VX_OM_BLEND_FUNC_CONST_RGB: return cst;
  case VX_OM_BLEND_FUNC_ONE_MINUS_CONST_RGB: return col4(0xff - cst.a, 0xff - cst.r, 0xff - cst.g, 0xff - cst.b);
  case VX_OM_BLEND_FUNC_CONST_A: return col4(cst.a, cst.a, cst.a, cst.a);
  case VX_OM_BLEND_FUNC_ONE_MINUS_CONST_A: return col4(0xff - cst.a, 0xff - cst.r, 0xff - cst.g, 0xff - cst.b);
  case VX_OM_BLEND_FUNC_ALPHA_SAT: {
    int factor = src.a < 0xff - dst.a ? src.a : 0xff - dst.a;
    return col4(0xff, factor, factor, factor);
  }
  }
}

static int div255(int x) { return (x + (x >> 8)) >> 8; } /* third_party/cocogfx/include/math.hpp:243-245 */
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* graphics.cpp:480-524 */
static o_col_t do_blend_mode(uint32_t mode, uint32_t logic_op, o_col_t src, o_col_t dst, o_col_t s, o_col_t d) {
  switch (mode) {
  default:
  case VX_OM_BLEND_MODE_ADD:
    return col4(div255(imin(src.a * s.a + dst.a * d.a + 0x80, 0xFF00)), div255(imin(src.r * s.r + dst.r * d.r + 0x80, 0xFF00)),
                div255(imin(src.g * s.g + dst.g * d.g + 0x80, 0xFF00)), div255(imin(src.b * s.b + dst.b * d.b + 0x80, 0xFF00)));
  case VX_OM_BLEND_MODE_SUB:
    return col4(div255(imax(src.a * s.a - dst.a * d.a + 0x80, 0)), div255(imax(src.r * s.r - dst.r * d.r + 0x80, 0)),
                div255(imax(src.g * s.g - dst.g * d.g + 0x80, 0)), div255(imax(src.b * s.b - dst.b * d.b + 0x80, 0)));
  case VX_OM_BLEND_MODE_REV_SUB:
    return col4(div255(imax(dst.a * d.a - src.a * s.a + 0x80, 0)), div255(imax(dst.r * d.r - src.r * s.r + 0x80, 0)),
                div255(imax(dst.g * d.g - src.g * s.g + 0x80, 0)), div255(imax(dst.b * d.b - src.b * s.b + 0x80, 0)));
  case VX_OM_BLEND_MODE_MIN:
    return col4(imin(src.a, dst.a), imin(src.r, dst.r), imin(src.g, dst.g), imin(src.b, dst.b));
  case VX_OM_BLEND_MODE_MAX:
    return col4(imax(src.a, dst.a), imax(src.r, dst.r), imax(src.g, dst.g), imax(src.b, dst.b));
  case VX_OM_BLEND_MODE_LOGICOP:
    return col_from(do_logic_op(logic_op, col_value(src), col_value(dst)));
  }
}

typedef struct {
  uint32_t mode_rgb, mode_a, src_rgb, src_a, dst_rgb, dst_a, cst, logic_op;
  int enabled;
} o_blend_t;

/* graphics.cpp:603-620 Blender::configure */
static void blend_configure(o_blend_t* b, const draw3d_kernel_arg_t* a) {
  b->mode_rgb = OM(a, VX_DCR_OM_BLEND_MODE) & 0xffff;
  b->mode_a = OM(a, VX_DCR_OM_BLEND_MODE) >> 16;
  b->src_rgb = (OM(a, VX_DCR_OM_BLEND_FUNC) >> 0) & 0xff;
  b->src_a = (OM(a, VX_DCR_OM_BLEND_FUNC) >> 8) & 0xff;
  b->dst_rgb = (OM(a, VX_DCR_OM_BLEND_FUNC) >> 16) & 0xff;
  b->dst_a = (OM(a, VX_DCR_OM_BLEND_FUNC) >> 24) & 0xff;
  b->cst = OM(a, VX_DCR_OM_BLEND_CONST);
  b->logic_op = OM(a, VX_DCR_OM_LOGIC_OP);
  b->enabled = !((b->mode_rgb == VX_OM_BLEND_MODE_ADD) && (b->mode_a == VX_OM_BLEND_MODE_ADD) &&
                 (b->src_rgb == VX_OM_BLEND_FUNC_ONE) && (b->src_a == VX_OM_BLEND_FUNC_ONE) &&
                 (b->dst_rgb == VX_OM_BLEND_FUNC_ZERO) && (b->dst_a == VX_OM_BLEND_FUNC_ZERO));
}

/* graphics.cpp:622-636 Blender::blend */
static uint32_t blend(const o_blend_t* b, uint32_t srcColor, uint32_t dstColor) {
  o_col_t src = col_from(srcColor), dst = col_from(dstColor), cst = col_from(b->cst);
  o_col_t s_rgb = do_blend_func(b->src_rgb, src, dst, cst), s_a = do_blend_func(b->src_a, src, dst, cst);
  o_col_t d_rgb = do_blend_func(b->dst_rgb, src, dst, cst), d_a = do_blend_func(b->dst_a, src, dst, cst);
  o_col_t rgb = do_blend_mode(b->mode_rgb, b->logic_op, src, dst, s_rgb, d_rgb);
  o_col_t a = do_blend_mode(b->mode_a, b->logic_op, src, dst, s_a, d_a);
  return col_value(col4(a.a, rgb.r, rgb.g, rgb.b));
}

/* ---- output merger: tests/regression/draw3d/gpu_sw.h:78-168 ------------------------------------ */
typedef struct {
  o_ds_t ds;
  o_blend_t bl;
  uint32_t zbuf_pitch, cbuf_pitch, cbuf_writemask, stencil_writemask;
  int depth_writemask, color_read, color_write;
  uint8_t *zbuf, *cbuf;
} o_om_t;

static void om_configure(o_om_t* m, const o_dev_t* dev) {
  const draw3d_kernel_arg_t* a = &dev->arg;
  uint32_t wm = OM(a, VX_DCR_OM_CBUF_WRITEMASK) & 0xf;
  ds_configure(&m->ds, a);
  blend_configure(&m->bl, a);
  m->zbuf = dev->zbuf;
  m->zbuf_pitch = OM(a, VX_DCR_OM_ZBUF_PITCH);
  m->depth_writemask = OM(a, VX_DCR_OM_DEPTH_WRITEMASK) & 0x1;
  m->stencil_writemask = OM(a, VX_DCR_OM_STENCIL_WRITEMASK) & 0xffff;
  m->cbuf = dev->cbuf;
  m->cbuf_pitch = OM(a, VX_DCR_OM_CBUF_PITCH);
  m->cbuf_writemask = (((wm >> 0) & 1) * 0x000000ffu) | (((wm >> 1) & 1) * 0x0000ff00u) |
                      (((wm >> 2) & 1) * 0x00ff0000u) | (((wm >> 3) & 1) * 0xff000000u);
  m->color_read = (wm != 0xf);
  m->color_write = (wm != 0x0);
}

/* gpu_sw.h:103-168 OutputMerger::write (+ private read/write) */
static void om_write(const o_om_t* m, uint32_t x, uint32_t y, uint32_t color, uint32_t depth) {
  int blend_enabled = m->bl.enabled, depth_enabled = m->ds.depth_enabled, stencil_enabled = m->ds.stencil_enabled;
  uint32_t depthstencil = 0, dst_ds = 0, dst_color = 0;
  uint32_t* zp = (uint32_t*)(m->zbuf + (uint64_t)y * m->zbuf_pitch + (uint64_t)x * 4);
  uint32_t* cp = (uint32_t*)(m->cbuf + (uint64_t)y * m->cbuf_pitch + (uint64_t)x * 4);
  int ds_passed;
  uint32_t ds_writemask;
  if (depth_enabled || stencil_enabled) dst_ds = *zp;
  if (m->color_write && (m->color_read || blend_enabled)) dst_color = *cp;
  ds_passed = !(depth_enabled || stencil_enabled) || ds_test(&m->ds, depth, dst_ds, &depthstencil);
  if (blend_enabled && ds_passed) color = blend(&m->bl, color, dst_color);
  ds_writemask = ((depth_enabled && ds_passed && m->depth_writemask) ? VX_OM_DEPTH_MASK : 0) |
                 (stencil_enabled ? (m->stencil_writemask << VX_OM_DEPTH_BITS) : 0);
  if (ds_writemask != 0) *zp = (dst_ds & ~ds_writemask) | (depthstencil & ds_writemask);
  if (m->color_write && ds_passed) *cp = (dst_color & ~m->cbuf_writemask) | (color & m->cbuf_writemask);
}

/* ---- fragment shader: tests/regression/draw3d/kernel.cpp:37-71, :232-279 ----------------------- */
typedef struct {
  const o_dev_t* dev;
  o_om_t om;
  uint32_t scissor_left, scissor_right, scissor_top, scissor_bottom;
} o_ctx_t;

/* kernel.cpp:48-51 imadd with s = 3 */
static int32_t imadd(int32_t a, int32_t b, int32_t c) {
  int32_t p = (int32_t)(((int64_t)a * (int64_t)b) >> 24);
  return (int32_t)((uint32_t)p + (uint32_t)c);
}
static int32_t interp(rast_attrib_t s, int32_t dx, int32_t dy) { /* :56-59 */
  int32_t tmp = imadd(s.x, dx, s.z);
  return imadd(s.y, dy, tmp);
}
static uint32_t mul_shift24_u8(int32_t fx, uint32_t k) { /* :61-71: (data * k) >> 24 stored to uint8_t */
  return (((uint32_t)fx * k) >> 24) & 0xff;
}

/* test-side instrumentation (not in the reference): fragments handed to the output merger */
static int g_count_frags = 0;
static volatile uint64_t g_frags = 0;

static void shade_quad(const o_ctx_t* c, uint32_t pos_mask, const int32_t bc[4][3], uint32_t pid) {
  const draw3d_kernel_arg_t* a = &c->dev->arg;
  const rast_prim_t* prim = (const rast_prim_t*)(c->dev->primbuf + (uint64_t)pid * sizeof(rast_prim_t));
  const rast_attribs_t* at = &prim->attribs;
  uint32_t mask = pos_mask & 0xf;
  uint32_t qx = (pos_mask >> 4) & ((1 << (VX_RASTER_DIM_BITS - 1)) - 1);
  uint32_t qy = (pos_mask >> (4 + (VX_RASTER_DIM_BITS - 1))) & ((1 << (VX_RASTER_DIM_BITS - 1)) - 1);
  int i;
  if (g_count_frags) __sync_fetch_and_add(&g_frags, (uint64_t)__builtin_popcount(mask));
  for (i = 0; i < 4; ++i) {
    int32_t z = 0, r = 1 << 24, g = 1 << 24, b = 1 << 24, al = 1 << 24, u = 0, v = 0, dx, dy; /* DEFAULTS :15-22 */
    uint32_t out;
    if (!(mask & (1u << i))) continue; /* the reference shades all four and drops these at OUTPUT; no side effects */
    { /* GRADIENTS_SW :37-44 ; float(TFixed) = float(data) / (1 << 24), fixed.hpp:404-406 */
      float F0 = (float)bc[i][0] / 16777216.0f, F1 = (float)bc[i][1] / 16777216.0f, F2 = (float)bc[i][2] / 16777216.0f;
      float rr = 1.0f / (F0 + F1 + F2);
      dx = fx24(rr * F0);
      dy = fx24(rr * F1);
    }
    if (a->depth_enabled) z = interp(at->z, dx, dy);
    if (a->color_enabled) {
      r = interp(at->r, dx, dy); g = interp(at->g, dx, dy); b = interp(at->b, dx, dy); al = interp(at->a, dx, dy);
    }
    if (a->tex_enabled) { u = interp(at->u, dx, dy); v = interp(at->v, dx, dy); }
    if (a->tex_enabled) {
      /* TEXTURING :160-164: TFixed<23>(TFixed<24>) = data >> 1 (fixed.hpp:20-22), lod 0 */
      uint32_t t = tex_read(a, c->dev->texbuf, u >> 1, v >> 1, 0);
      if (a->tex_modulate) {
        out = (mul_shift24_u8(al, t >> 24) << 24) | (mul_shift24_u8(r, (t >> 16) & 0xff) << 16) |
              (mul_shift24_u8(g, (t >> 8) & 0xff) << 8) | mul_shift24_u8(b, t & 0xff);
      } else {
        out = t;
      }
    } else {
      out = (mul_shift24_u8(al, 255) << 24) | (mul_shift24_u8(r, 255) << 16) | (mul_shift24_u8(g, 255) << 8) |
            mul_shift24_u8(b, 255);
    }
    /* OUTPUT :81-87, :166-173 */
    om_write(&c->om, (qx << 1) + (i & 1), (qy << 1) + (i >> 1), out, (uint32_t)z);
  }
}

/* ---- rasterizer: sim/common/graphics.cpp:640-842 (fixed-point FloatE; int32 wrap-around arithmetic
 * as the compiled reference does, fixed.hpp:318-330, :352-362) ------------------------------------ */
typedef struct { int32_t dx[3], dy[3], ext[3]; } o_delta_t;
static int32_t wadd(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static int32_t wmul(int32_t a, uint32_t b) { return (int32_t)((uint32_t)a * b); }
static int32_t wshl(int32_t a, uint32_t s) { return (int32_t)((uint32_t)a << s); }

/* graphics.cpp:644-647 */
static int32_t calc_edge_extents(rast_vec3e_t e) {
  return (e.y >= 0) ? ((e.x >= 0) ? wadd(e.x, e.y) : e.y) : ((e.x >= 0) ? e.x : 0);
}

/* graphics.cpp:799-842 */
static void render_quad(const o_ctx_t* c, uint32_t x, uint32_t y, uint32_t pid, const int32_t e[3], const o_delta_t* d) {
  int32_t bc[4][3];
  uint32_t mask = 0;
  int i, j, k;
  for (k = 0; k < 3; ++k)
    if (wadd(e[k], wshl(d->ext[k], 1)) < 0) return;
  for (j = 0; j < 2; ++j)
    for (i = 0; i < 2; ++i) {
      uint32_t p = j * 2 + i;
      int cov = 1;
      for (k = 0; k < 3; ++k) {
        bc[p][k] = wadd(wadd(e[k], wmul(d->dx[k], i)), wmul(d->dy[k], j));
        cov = cov && (bc[p][k] >= 0);
      }
      cov = cov && (x + i) >= c->scissor_left && (x + i) < c->scissor_right && (y + j) >= c->scissor_top &&
            (y + j) < c->scissor_bottom;
      mask |= (uint32_t)cov << p;
    }
  if (mask) {
    uint32_t pos_mask = ((y / 2) << (4 + VX_RASTER_DIM_BITS - 1)) | ((x / 2) << 4) | mask;
    shade_quad(c, pos_mask, (const int32_t(*)[3])bc, pid);
  }
}

/* graphics.cpp:743-797 */
static void render_tile(const o_ctx_t* c, uint32_t L, uint32_t x, uint32_t y, uint32_t pid, const int32_t e[3], const o_delta_t* d) {
  int k;
  for (k = 0; k < 3; ++k)
    if (wadd(e[k], wshl(d->ext[k], L)) < 0) return;
  if (L > 1) {
    int32_t s[3];
    uint32_t sub;
    --L;
    sub = 1u << L;
    render_tile(c, L, x, y, pid, e, d);
    for (k = 0; k < 3; ++k) s[k] = wadd(e[k], wshl(d->dx[k], L));
    render_tile(c, L, x + sub, y, pid, s, d);
    for (k = 0; k < 3; ++k) s[k] = wadd(e[k], wshl(d->dy[k], L));
    render_tile(c, L, x, y + sub, pid, s, d);
    for (k = 0; k < 3; ++k) s[k] = wadd(wadd(e[k], wshl(d->dx[k], L)), wshl(d->dy[k], L));
    render_tile(c, L, x + sub, y + sub, pid, s, d);
  } else {
    render_quad(c, x, y, pid, e, d);
  }
}

/* graphics.cpp:715-741 renderPrimitive + gpu_sw.h:34-62 Rasterizer::render (one tile) */
static void render_one_tile(const o_ctx_t* c, uint32_t cur_tile, uint32_t tile_logsize) {
  const draw3d_kernel_arg_t* a = &c->dev->arg;
  uint32_t x = (c->dev->tile_xy[cur_tile] >> 16) << tile_logsize, y = (c->dev->tile_xy[cur_tile] & 0xffff) << tile_logsize;
  uint32_t count = c->dev->tile_start[cur_tile + 1] - c->dev->tile_start[cur_tile], stride = RAS(a, VX_DCR_RASTER_PBUF_STRIDE);
  const uint32_t* pid_buf = c->dev->pid_list + c->dev->tile_start[cur_tile]; /* == (hdr + 1) + hdr->pids_offset */
  while (count--) {
    uint32_t pid = *pid_buf++;
    const rast_prim_t* prim = (const rast_prim_t*)(c->dev->primbuf + (uint64_t)pid * stride);
    o_delta_t d;
    int32_t value[3];
    int k;
    for (k = 0; k < 3; ++k) {
      d.dx[k] = prim->edges[k].x;
      d.dy[k] = prim->edges[k].y;
      d.ext[k] = calc_edge_extents(prim->edges[k]);
      value[k] = wadd(wadd(wmul(prim->edges[k].x, x), wmul(prim->edges[k].y, y)), prim->edges[k].z); /* :640-642 */
    }
    render_tile(c, tile_logsize, x, y, pid, value, &d);
  }
}

typedef struct {
  const o_ctx_t* ctx;
  uint32_t num_tiles, tile_logsize;
  volatile uint32_t next;
} o_job_t;

static void* tile_worker(void* arg) {
  o_job_t* j = (o_job_t*)arg;
  for (;;) {
    uint32_t t = __sync_fetch_and_add(&j->next, 1);
    if (t >= j->num_tiles) break;
    render_one_tile(j->ctx, t, j->tile_logsize);
  }
  return NULL;
}

/* kernel.cpp:286-299 main + vx_spawn_threads (kernel/src/vx_spawn.c:54-81): tiles are disjoint, so
 * the tile loop is split over host threads */
static void run_kernel(const o_dev_t* dev, uint32_t tile_logsize, uint32_t num_threads) {
  o_ctx_t ctx;
  o_job_t job;
  pthread_t th[256];
  uint32_t i, n = num_threads ? num_threads : 1;
  const draw3d_kernel_arg_t* a = &dev->arg;
  if (n > 256) n = 256;
  ctx.dev = dev;
  om_configure(&ctx.om, dev);
  /* graphics.cpp:708-713 Rasterizer::configure */
  ctx.scissor_left = RAS(a, VX_DCR_RASTER_SCISSOR_X) & 0xffff;
  ctx.scissor_right = RAS(a, VX_DCR_RASTER_SCISSOR_X) >> 16;
  ctx.scissor_top = RAS(a, VX_DCR_RASTER_SCISSOR_Y) & 0xffff;
  ctx.scissor_bottom = RAS(a, VX_DCR_RASTER_SCISSOR_Y) >> 16;
  job.ctx = &ctx;
  job.num_tiles = RAS(a, VX_DCR_RASTER_TILE_COUNT);
  job.tile_logsize = tile_logsize;
  job.next = 0;
  for (i = 1; i < n; ++i) pthread_create(&th[i], NULL, tile_worker, &job);
  tile_worker(&job);
  for (i = 1; i < n; ++i) pthread_join(th[i], NULL);
}

/* ------------------------------------------------------------------------------------------------
 * Host flow: tests/regression/draw3d/main.cpp:171-389 (render) and :392-517 (main)
 * ---------------------------------------------------------------------------------------------- */
static uint32_t log2ceil_u32(uint32_t v) { /* sim/common/bitmanip.h:31-33 */
  uint32_t n = 0;
  while ((1ull << n) < v) ++n;
  return n;
}

/* sim/common/gfxutil.cpp:280-296, :298-314, :316-330, :332-350 */
static uint32_t to_vx_format(uint32_t f) {
  switch (f) {
  case SRT_FORMAT_A8R8G8B8: return VX_TEX_FORMAT_A8R8G8B8;
  case SRT_FORMAT_R5G6B5: return VX_TEX_FORMAT_R5G6B5;
  case SRT_FORMAT_A1R5G5B5: return VX_TEX_FORMAT_A1R5G5B5;
  case SRT_FORMAT_A4R4G4B4: return VX_TEX_FORMAT_A4R4G4B4;
  case SRT_FORMAT_A8L8: return VX_TEX_FORMAT_A8L8;
  case SRT_FORMAT_L8: return VX_TEX_FORMAT_L8;
  case SRT_FORMAT_A8: return VX_TEX_FORMAT_A8;
  default: return 0xffffffffu;
  }
}
static uint32_t to_vx_compare(uint32_t c) {
  static const uint32_t m[8] = {VX_OM_DEPTH_FUNC_NEVER, VX_OM_DEPTH_FUNC_LESS, VX_OM_DEPTH_FUNC_EQUAL, VX_OM_DEPTH_FUNC_LEQUAL,
                                VX_OM_DEPTH_FUNC_GREATER, VX_OM_DEPTH_FUNC_NOTEQUAL, VX_OM_DEPTH_FUNC_GEQUAL, VX_OM_DEPTH_FUNC_ALWAYS};
  return m[c & 7];
}
static uint32_t to_vx_stencil_op(uint32_t o) {
  static const uint32_t m[6] = {VX_OM_STENCIL_OP_KEEP, VX_OM_STENCIL_OP_REPLACE, VX_OM_STENCIL_OP_INCR,
                                VX_OM_STENCIL_OP_DECR, VX_OM_STENCIL_OP_ZERO, VX_OM_STENCIL_OP_INVERT};
  return m[o < 6 ? o : 0];
}
static uint32_t to_vx_blend_func(uint32_t o) {
  static const uint32_t m[11] = {VX_OM_BLEND_FUNC_ZERO, VX_OM_BLEND_FUNC_ONE, VX_OM_BLEND_FUNC_SRC_RGB,
                                 VX_OM_BLEND_FUNC_ONE_MINUS_SRC_RGB, VX_OM_BLEND_FUNC_SRC_A, VX_OM_BLEND_FUNC_ONE_MINUS_SRC_A,
                                 VX_OM_BLEND_FUNC_DST_A, VX_OM_BLEND_FUNC_ONE_MINUS_DST_A, VX_OM_BLEND_FUNC_DST_RGB,
                                 VX_OM_BLEND_FUNC_ONE_MINUS_DST_RGB, VX_OM_BLEND_FUNC_ALPHA_SAT};
  return m[o < 11 ? o : 0];
}

/* main.cpp:214-346: the DCR writes and the boolean switches of one draw call.  `arg` persists across
 * draws (the reference's global kernel_arg), so state a draw does not write keeps its last value. */
/* Test hook (not in main.cpp, which always programs MODE_ADD): OM DCR values for draws whose trace state enables
 * blending -- blend mode, blend func, blend constant, logic op (same hook as oracle/ref_draw3d_harness.cpp). */
static uint32_t g_om_override[4];
static int g_om_override_on = 0;

static void program_draw(draw3d_kernel_arg_t* arg, const srt_states_t* st, const o_tex_t* tex, uint32_t num_tiles,
                         uint32_t width, uint32_t height) {
  RAS(arg, VX_DCR_RASTER_TILE_COUNT) = num_tiles;
  RAS(arg, VX_DCR_RASTER_PBUF_STRIDE) = sizeof(rast_prim_t);
  RAS(arg, VX_DCR_RASTER_SCISSOR_X) = (width << 16) | 0;
  RAS(arg, VX_DCR_RASTER_SCISSOR_Y) = (height << 16) | 0;
  OM(arg, VX_DCR_OM_CBUF_PITCH) = width * 4;
  OM(arg, VX_DCR_OM_CBUF_WRITEMASK) = st->color_writemask;
  if (st->depth_test || st->stencil_test) OM(arg, VX_DCR_OM_ZBUF_PITCH) = width * 4;
  if (st->depth_test) {
    OM(arg, VX_DCR_OM_DEPTH_FUNC) = to_vx_compare(st->depth_func);
    OM(arg, VX_DCR_OM_DEPTH_WRITEMASK) = st->depth_writemask;
  } else {
    OM(arg, VX_DCR_OM_DEPTH_FUNC) = VX_OM_DEPTH_FUNC_ALWAYS;
    OM(arg, VX_DCR_OM_DEPTH_WRITEMASK) = 0;
  }
  if (st->stencil_test) {
    OM(arg, VX_DCR_OM_STENCIL_FUNC) = to_vx_compare(st->stencil_func);
    OM(arg, VX_DCR_OM_STENCIL_ZPASS) = to_vx_stencil_op(st->stencil_zpass);
    OM(arg, VX_DCR_OM_STENCIL_ZPASS) = to_vx_stencil_op(st->stencil_zfail); /* sic, main.cpp:260 */
    OM(arg, VX_DCR_OM_STENCIL_FAIL) = to_vx_stencil_op(st->stencil_fail);
    OM(arg, VX_DCR_OM_STENCIL_REF) = st->stencil_ref;
    OM(arg, VX_DCR_OM_STENCIL_MASK) = st->stencil_mask;
    OM(arg, VX_DCR_OM_STENCIL_WRITEMASK) = st->stencil_writemask;
  } else {
    OM(arg, VX_DCR_OM_STENCIL_FUNC) = VX_OM_DEPTH_FUNC_ALWAYS;
    OM(arg, VX_DCR_OM_STENCIL_ZPASS) = VX_OM_STENCIL_OP_KEEP;
    OM(arg, VX_DCR_OM_STENCIL_FAIL) = VX_OM_STENCIL_OP_KEEP;
    OM(arg, VX_DCR_OM_STENCIL_REF) = 0;
    OM(arg, VX_DCR_OM_STENCIL_MASK) = VX_OM_STENCIL_MASK;
    OM(arg, VX_DCR_OM_STENCIL_WRITEMASK) = 0;
  }
  OM(arg, VX_DCR_OM_BLEND_MODE) = (VX_OM_BLEND_MODE_ADD << 16) | (VX_OM_BLEND_MODE_ADD << 0);
  if (st->blend_enabled && g_om_override_on) { /* test hook, not in main.cpp (see ord_set_om_override) */
    OM(arg, VX_DCR_OM_BLEND_MODE) = g_om_override[0];
    OM(arg, VX_DCR_OM_BLEND_FUNC) = g_om_override[1];
    OM(arg, VX_DCR_OM_BLEND_CONST) = g_om_override[2];
    OM(arg, VX_DCR_OM_LOGIC_OP) = g_om_override[3];
  } else if (st->blend_enabled) {
    uint32_t s = to_vx_blend_func(st->blend_src), d = to_vx_blend_func(st->blend_dst);
    OM(arg, VX_DCR_OM_BLEND_FUNC) = (d << 24) | (d << 16) | (s << 8) | (s << 0);
  } else {
    OM(arg, VX_DCR_OM_BLEND_FUNC) = (VX_OM_BLEND_FUNC_ZERO << 24) | (VX_OM_BLEND_FUNC_ZERO << 16) |
                                    (VX_OM_BLEND_FUNC_ONE << 8) | (VX_OM_BLEND_FUNC_ONE << 0);
  }
  if (st->texture_enabled && tex) {
    int tex_filter = (st->texture_magfilter != SRT_FILTER_NEAREST) || (st->texture_magfilter != SRT_FILTER_NEAREST);
    uint32_t wrapU = (st->texture_addressU == SRT_ADDRESS_WRAP), wrapV = (st->texture_addressU == SRT_ADDRESS_WRAP); /* sic :304 */
    arg->tex_stage = 0;
    TEX(arg, VX_DCR_TEX_LOGDIM) = (log2ceil_u32(tex->height) << 16) | log2ceil_u32(tex->width);
    TEX(arg, VX_DCR_TEX_FORMAT) = to_vx_format(tex->format);
    TEX(arg, VX_DCR_TEX_WRAP) = (wrapV << 16) | wrapU;
    TEX(arg, VX_DCR_TEX_FILTER) = tex_filter ? VX_TEX_FILTER_BILINEAR : VX_TEX_FILTER_POINT;
    TEX(arg, VX_DCR_TEX_MIPOFF(0)) = 0; /* level 0 only: draw3d never samples another (kernel.cpp:160-164) */
  }
  arg->depth_enabled = st->depth_test != 0;
  arg->color_enabled = st->color_enabled != 0;
  arg->tex_enabled = st->texture_enabled != 0;
  arg->tex_modulate = (st->texture_enabled && st->texture_envmode == SRT_ENVMODE_MODULATE);
  if (arg->tex_modulate && !arg->color_enabled) arg->tex_modulate = 0;
  if (arg->tex_enabled && arg->color_enabled && !arg->tex_modulate) arg->color_enabled = 0;
}

static double now_s(void) {
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec;
}

void ref_draw3d_set_om_override(const uint32_t* ov) {
  int i;
  g_om_override_on = ov != 0;
  if (ov) for (i = 0; i < 4; ++i) g_om_override[i] = ov[i];
}

uint32_t ref_draw3d_kernel_arg_size(void) { return (uint32_t)sizeof(draw3d_kernel_arg_t); }
uint32_t ref_draw3d_prim_size(void) { return (uint32_t)sizeof(rast_prim_t); }

int64_t ref_draw3d_binning(const uint8_t* blob, uint64_t blob_size, uint32_t draw, uint32_t width, uint32_t height,
                           uint32_t tile_logsize, uint8_t* tilebuf_out, uint64_t* tilebuf_size, uint8_t* primbuf_out,
                           uint64_t* primbuf_size) {
  o_trace_t t;
  o_bins_t b;
  int64_t rc = -1;
  if (o_parse(&t, blob, blob_size) != 0 || draw >= t.nd) { o_free(&t); return -1; }
  o_binning(&b, &t.draws[draw], width, height, tile_logsize);
  if (b.tilebuf_size <= *tilebuf_size && (uint64_t)b.num_prims * sizeof(rast_prim_t) <= *primbuf_size) {
    memcpy(tilebuf_out, b.tilebuf, b.tilebuf_size);
    memcpy(primbuf_out, b.prims, (uint64_t)b.num_prims * sizeof(rast_prim_t));
    *tilebuf_size = b.tilebuf_size;
    *primbuf_size = (uint64_t)b.num_prims * sizeof(rast_prim_t);
    rc = b.num_tiles;
  }
  o_bins_free(&b);
  o_free(&t);
  return rc;
}

int ref_draw3d_render(const uint8_t* blob, uint64_t blob_size, uint32_t width, uint32_t height, uint32_t num_threads,
                      uint32_t tile_logsize, uint32_t* cbuf_out, uint32_t* zbuf_out, double* kernel_seconds) {
  o_trace_t t;
  o_dev_t dev;
  uint64_t npix = (uint64_t)width * height, i;
  uint32_t d;
  double ksec = 0.0;
  if (o_parse(&t, blob, blob_size) != 0) { o_free(&t); return -1; }
  memset(&dev, 0, sizeof(dev));
  dev.zbuf = (uint8_t*)malloc(npix * 4);
  dev.cbuf = (uint8_t*)malloc(npix * 4);
  for (i = 0; i < npix; ++i) { ((uint32_t*)dev.zbuf)[i] = 0xffffffffu; ((uint32_t*)dev.cbuf)[i] = 0xff000000u; } /* main.cpp:45-46,:478-490 */
  dev.arg.log_num_tasks = 6;
  dev.arg.sw_tex = dev.arg.sw_rast = dev.arg.sw_om = 1;
  for (d = 0; d < t.nd; ++d) {
    const o_draw_t* dc = &t.draws[d];
    const o_tex_t* tex = NULL;
    o_bins_t b;
    double t0;
    o_binning(&b, dc, width, height, tile_logsize);
    if (b.num_tiles == 0) { o_bins_free(&b); continue; } /* main.cpp:190-191 */
    if (dc->st.texture_enabled) {
      uint32_t k;
      for (k = 0; k < t.nt; ++k)
        if (t.texs[k].id == dc->texture_id) tex = &t.texs[k];
      if (!tex) { o_bins_free(&b); o_free(&t); free(dev.zbuf); free(dev.cbuf); return -3; }
      dev.texbuf = tex->pixels;
    }
    program_draw(&dev.arg, &dc->st, tex, b.num_tiles, width, height);
    dev.tilebuf = b.tilebuf;
    dev.tile_xy = b.tile_xy; dev.tile_start = b.tile_start; dev.pid_list = b.pid_list;
    dev.primbuf = (const uint8_t*)b.prims;
    t0 = now_s();
    run_kernel(&dev, tile_logsize, num_threads);
    ksec += now_s() - t0;
    o_bins_free(&b);
  }
  if (cbuf_out) memcpy(cbuf_out, dev.cbuf, npix * 4);
  if (zbuf_out) memcpy(zbuf_out, dev.zbuf, npix * 4);
  if (kernel_seconds) *kernel_seconds = ksec;
  free(dev.zbuf); free(dev.cbuf);
  o_free(&t);
  return 0;
}

/* Fragments one frame sends to the output merger (the unit bench.py's draw3d metric counts). */
uint64_t ref_draw3d_count_fragments(const uint8_t* blob, uint64_t blob_size, uint32_t width, uint32_t height,
                                    uint32_t num_threads) {
  uint64_t n;
  g_frags = 0;
  g_count_frags = 1;
  if (ref_draw3d_render(blob, blob_size, width, height, num_threads, RASTER_TILE_LOGSIZE, NULL, NULL, NULL) != 0) g_frags = 0;
  g_count_frags = 0;
  n = g_frags;
  return n;
}

/* ---- sub-stage probes ------------------------------------------------------------------------- */
uint32_t ref_draw3d_blend(const draw3d_kernel_arg_t* arg, uint32_t src, uint32_t dst) {
  o_blend_t b;
  blend_configure(&b, arg);
  return blend(&b, src, dst);
}
int ref_draw3d_depth_stencil(const draw3d_kernel_arg_t* arg, uint32_t depth, uint32_t dst_ds, uint32_t* out_ds) {
  o_ds_t s;
  ds_configure(&s, arg);
  return ds_test(&s, depth, dst_ds, out_ds);
}
uint32_t ref_draw3d_tex(const draw3d_kernel_arg_t* arg, const uint8_t* texels, int32_t u, int32_t v) {
  return tex_read(arg, texels, u, v, 0);
}

/* The reference's texture-unit regression kernel (tests/regression/tex/kernel.cpp:42-134) over the texture state of
 * stage 0 in `arg`: a dst_width x dst_height image whose pixel (x, y) samples the texture at ((x + 0.5) / W,
 * (y + 0.5) / H), the coordinates accumulated in fp32 task by task exactly as the kernel does (fu += deltaX), at the
 * mip level its main() derives from the minification.  `trilinear` is the kernel's `filter == 2` branch (two levels
 * blended by `frac`; the reference app never sets that field, tex/main.cpp:231-252).  out: row 0 first. */
void ref_draw3d_tex_image(const draw3d_kernel_arg_t* arg, const uint8_t* texels, uint32_t dst_width, uint32_t dst_height,
                          uint32_t num_tasks, int trilinear, uint32_t* out) {
  if (num_tasks == 0) num_tasks = 1;
  if (num_tasks > dst_height) num_tasks = dst_height; /* tex/main.cpp:235 */
  /* kernel.cpp:106-128 main() */
  const uint32_t tile_height = (dst_height + num_tasks - 1) / num_tasks;
  const float deltaX = 1.0f / dst_width, deltaY = 1.0f / dst_height;
  const uint32_t logdim = TEX(arg, VX_DCR_TEX_LOGDIM);
  const float width_ratio = (float)(1 << (logdim & 0xffff)) / dst_width;
  const float height_ratio = (float)(1 << (logdim >> 16)) / dst_height;
  float minification = width_ratio > height_ratio ? width_ratio : height_ratio;
  if (minification < 1.0f) minification = 1.0f;
  const int32_t j = (int32_t)(minification * 65536); /* fixed16_t(float): fixed.hpp:97-100 */
  uint32_t lod = (uint32_t)(31 - __builtin_clz((uint32_t)j)) - 16;
  if (lod > VX_TEX_LOD_MAX) lod = VX_TEX_LOD_MAX;
  const uint32_t frac = (uint32_t)(j - (1 << (lod + 16))) >> (lod + 16 - 8);
  const uint32_t lodn = lod + 1 < VX_TEX_LOD_MAX ? lod + 1 : VX_TEX_LOD_MAX;
  /* kernel.cpp:42-103 kernel_body, one call per task */
  for (uint32_t task = 0; task < num_tasks; ++task) {
    const uint32_t y_start = task * tile_height;
    const uint32_t y_end = y_start + tile_height < dst_height ? y_start + tile_height : dst_height;
    float fv = (y_start + 0.5f) * deltaY;
    for (uint32_t y = y_start; y < y_end; ++y) {
      float fu = (0 + 0.5f) * deltaX;
      for (uint32_t x = 0; x < dst_width; ++x) {
        const int32_t xu = (int32_t)(fu * (1 << VX_TEX_FXD_FRAC)), xv = (int32_t)(fv * (1 << VX_TEX_FXD_FRAC));
        uint32_t color = tex_read(arg, texels, xu, xv, lod);
        if (trilinear) {
          const uint32_t texel1 = tex_read(arg, texels, xu, xv, lodn);
          const uint32_t cl = lerp8888(color & 0x00ff00ff, texel1 & 0x00ff00ff, frac);
          const uint32_t ch = lerp8888((color >> 8) & 0x00ff00ff, (texel1 >> 8) & 0x00ff00ff, frac);
          color = (ch << 8) | cl;
        }
        out[(uint64_t)y * dst_width + x] = color;
        fu += deltaX;
      }
      fv += deltaY;
    }
  }
}
