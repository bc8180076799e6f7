/*
 * draw3d_cgltrace.c -- reader for the reference's own trace files (plain C, no Boost).
 *
 * A .cgltrace is a Boost.Serialization XML archive of cocogfx::CGLTrace
 * (third_party/cocogfx/src/cgltrace.cpp:24-165).  Boost writes the members in serialize() order with one
 * element per member, so a forward-only scan for the next `<name ...>text<` is enough: no DOM, no
 * attribute handling, no entity decoding (the archive holds only numbers and base64).
 *
 *   cgltrace
 *     drawcalls: count, item x count { states{24 scalars, texture_envcolor{r,g,b,a} after texture_enabled},
 *                texture_id, vertices: count, item x count { first, second{pos{x,y,z,w}, color{r,g,b,a},
 *                texcoord{u,v}} }, primitives: count, item x count {i0,i1,i2}, viewport{left,right,top,bottom,
 *                near,far} }
 *     textures: count, item x count { first, second{format, width, height, size, pixels(base64)} }
 *
 * The result is re-encoded into the same flat SKBXTRC1 blob tests/golden/make_raster_fixtures.py writes
 * (vertices dense in ascending key order, primitive indices remapped) and handed to the common parser, so
 * both routes give byte-identical traces (tests/test_oracle_draw3d.py checks that).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include "skybox_draw3d.h"

typedef struct { const char* p; const char* end; int ok; } cur_t;

/* Advance to the text of the next element called `name`; returns its start, *len = its length. */
static const char* next_text(cur_t* c, const char* name, size_t* len) {
  size_t n = strlen(name);
  const char* p = c->p;
  while (c->ok && p + n + 2 < c->end) {
    p = (const char*)memchr(p, '<', (size_t)(c->end - p));
    if (!p || p + n + 2 >= c->end) break;
    if (memcmp(p + 1, name, n) == 0 && (p[1 + n] == '>' || p[1 + n] == ' ')) {
      const char* t = (const char*)memchr(p, '>', (size_t)(c->end - p));
      const char* e;
      if (!t) break;
      ++t;
      e = (const char*)memchr(t, '<', (size_t)(c->end - t));
      if (!e) break;
      c->p = e;
      *len = (size_t)(e - t);
      return t;
    }
    ++p;
  }
  c->ok = 0;
  *len = 0;
  return NULL;
}

static uint32_t next_u32(cur_t* c, const char* name) {
  size_t len;
  const char* t = next_text(c, name, &len);
  char buf[32];
  if (!t || len == 0 || len >= sizeof(buf)) { c->ok = 0; return 0; }
  memcpy(buf, t, len);
  buf[len] = 0;
  return (uint32_t)strtoull(buf, NULL, 10);
}

static int32_t next_i32(cur_t* c, const char* name) {
  size_t len;
  const char* t = next_text(c, name, &len);
  char buf[32];
  if (!t || len == 0 || len >= sizeof(buf)) { c->ok = 0; return 0; }
  memcpy(buf, t, len);
  buf[len] = 0;
  return (int32_t)strtol(buf, NULL, 10);
}

static float next_f32(cur_t* c, const char* name) {
  size_t len;
  const char* t = next_text(c, name, &len);
  char buf[64];
  if (!t || len == 0 || len >= sizeof(buf)) { c->ok = 0; return 0.0f; }
  memcpy(buf, t, len);
  buf[len] = 0;
  return strtof(buf, NULL);
}

/* growable output blob */
typedef struct { uint8_t* d; size_t n, cap; int ok; } out_t;
static void put(out_t* o, const void* src, size_t n) {
  if (!o->ok) return;
  if (o->n + n > o->cap) {
    size_t cap = o->cap ? o->cap : 1 << 16;
    uint8_t* d;
    while (cap < o->n + n) cap *= 2;
    d = (uint8_t*)realloc(o->d, cap);
    if (!d) { o->ok = 0; return; }
    o->d = d;
    o->cap = cap;
  }
  memcpy(o->d + o->n, src, n);
  o->n += n;
}
static void put_u32(out_t* o, uint32_t v) { put(o, &v, 4); }
static void put_f32(out_t* o, float v) { put(o, &v, 4); }

static int b64_value(int ch) {
  if (ch >= 'A' && ch <= 'Z') return ch - 'A';
  if (ch >= 'a' && ch <= 'z') return ch - 'a' + 26;
  if (ch >= '0' && ch <= '9') return ch - '0' + 52;
  if (ch == '+') return 62;
  if (ch == '/') return 63;
  return -1; /* whitespace, '=' */
}

/* boost::archive binary_object: base64 with line breaks; decodes exactly `size` bytes */
static int b64_decode(const char* t, size_t len, uint8_t* dst, uint32_t size) {
  uint32_t acc = 0, bits = 0, n = 0;
  size_t i;
  for (i = 0; i < len && n < size; ++i) {
    int v = b64_value((unsigned char)t[i]);
    if (v < 0) continue;
    acc = (acc << 6) | (uint32_t)v;
    bits += 6;
    if (bits >= 8) {
      bits -= 8;
      dst[n++] = (uint8_t)(acc >> bits);
    }
  }
  return n == size ? 0 : -1;
}

typedef struct { uint32_t key; float v[10]; } vert_t;
static int by_key(const void* a, const void* b) {
  uint32_t x = ((const vert_t*)a)->key, y = ((const vert_t*)b)->key;
  return x < y ? -1 : x > y;
}
static int find_key(const vert_t* v, uint32_t n, uint32_t key, uint32_t* pos) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    uint32_t mid = (lo + hi) / 2;
    if (v[mid].key < key) lo = mid + 1; else hi = mid;
  }
  if (lo < n && v[lo].key == key) { *pos = lo; return 0; }
  return -1;
}

static const char* const kStateNames[24] = {
    "color_enabled", "color_format", "color_writemask", "depth_test", "depth_writemask", "depth_format", "depth_func",
    "stencil_test", "stencil_func", "stencil_zpass", "stencil_zfail", "stencil_fail", "stencil_ref", "stencil_mask",
    "stencil_writemask", "texture_enabled", "texture_envmode", "texture_minfilter", "texture_magfilter",
    "texture_addressU", "texture_addressV", "blend_enabled", "blend_src", "blend_dst"};

static uint8_t* read_all(const char* path, size_t* size) {
  gzFile f = gzopen(path, "rb");
  size_t cap = 1 << 20, n = 0;
  uint8_t* buf;
  int got;
  if (!f) return NULL;
  buf = (uint8_t*)malloc(cap);
  while (buf && (got = gzread(f, buf + n, (unsigned)(cap - n))) > 0) {
    n += (size_t)got;
    if (n == cap) {
      uint8_t* nb = (uint8_t*)realloc(buf, cap *= 2);
      if (!nb) { free(buf); buf = NULL; }
      else buf = nb;
    }
  }
  gzclose(f);
  *size = n;
  return buf;
}

/* XML text -> SKBXTRC1 blob (malloc'ed).  Returns NULL on malformed input. */
static uint8_t* cgltrace_to_blob(const char* xml, size_t xml_size, uint64_t* blob_size) {
  cur_t c = {xml, xml + xml_size, 1};
  out_t draws = {NULL, 0, 0, 1}, texs = {NULL, 0, 0, 1}, all = {NULL, 0, 0, 1};
  uint32_t nd, nt, d, i, k;
  size_t len;
  if (!next_text(&c, "cgltrace", &len)) goto fail;
  nd = next_u32(&c, "count");
  if (!c.ok || nd > (1u << 20)) goto fail;
  for (d = 0; d < nd; ++d) {
    uint32_t st[24], nv, np, texture_id;
    float env[4];
    vert_t* verts;
    for (k = 0; k < 24; ++k) {
      st[k] = next_u32(&c, kStateNames[k]);
      if (k == 15) { /* texture_envcolor sits between texture_enabled and texture_envmode */
        env[0] = next_f32(&c, "r"); env[1] = next_f32(&c, "g"); env[2] = next_f32(&c, "b"); env[3] = next_f32(&c, "a");
      }
    }
    texture_id = next_u32(&c, "texture_id");
    if (!next_text(&c, "vertices", &len)) goto fail;
    nv = next_u32(&c, "count");
    if (!c.ok || nv > (1u << 26)) goto fail;
    verts = (vert_t*)malloc((nv ? nv : 1) * sizeof(vert_t));
    if (!verts) goto fail;
    for (i = 0; i < nv && c.ok; ++i) {
      static const char* const f[10] = {"x", "y", "z", "w", "r", "g", "b", "a", "u", "v"};
      verts[i].key = next_u32(&c, "first");
      for (k = 0; k < 10; ++k) verts[i].v[k] = next_f32(&c, f[k]);
    }
    qsort(verts, nv, sizeof(vert_t), by_key);
    if (!next_text(&c, "primitives", &len)) { free(verts); goto fail; }
    np = next_u32(&c, "count");
    if (!c.ok || np > (1u << 26)) { free(verts); goto fail; }
    /* flat layout: states, envcolor, texture_id, viewport, counts, vertices, primitives */
    for (k = 0; k < 24; ++k) put_u32(&draws, st[k]);
    for (k = 0; k < 4; ++k) put_f32(&draws, env[k]);
    put_u32(&draws, texture_id);
    {
      size_t viewport_at = draws.n;
      uint32_t zero[6] = {0, 0, 0, 0, 0, 0};
      put(&draws, zero, 24); /* patched below: the viewport follows the primitives in the archive */
      put_u32(&draws, nv);
      put_u32(&draws, np);
      for (i = 0; i < nv; ++i) put(&draws, verts[i].v, 40);
      for (i = 0; i < np && c.ok; ++i) {
        static const char* const f[3] = {"i0", "i1", "i2"};
        for (k = 0; k < 3; ++k) {
          uint32_t pos = 0;
          if (find_key(verts, nv, next_u32(&c, f[k]), &pos) != 0) c.ok = 0; /* vertices.at() would throw */
          put_u32(&draws, pos);
        }
      }
      free(verts);
      if (!next_text(&c, "viewport", &len) || !draws.ok) goto fail;
      {
        int32_t vi[4];
        float vf[2];
        vi[0] = next_i32(&c, "left"); vi[1] = next_i32(&c, "right"); vi[2] = next_i32(&c, "top"); vi[3] = next_i32(&c, "bottom");
        vf[0] = next_f32(&c, "near"); vf[1] = next_f32(&c, "far");
        memcpy(draws.d + viewport_at, vi, 16);
        memcpy(draws.d + viewport_at + 16, vf, 8);
      }
    }
    if (!c.ok) goto fail;
  }
  if (!next_text(&c, "textures", &len)) goto fail;
  nt = next_u32(&c, "count");
  if (!c.ok || nt > (1u << 20)) goto fail;
  for (i = 0; i < nt; ++i) {
    uint32_t hdr[5], pad = 0;
    const char* t;
    uint8_t* pix;
    hdr[0] = next_u32(&c, "first");
    hdr[1] = next_u32(&c, "format"); hdr[2] = next_u32(&c, "width"); hdr[3] = next_u32(&c, "height");
    hdr[4] = next_u32(&c, "size");
    t = next_text(&c, "pixels", &len);
    if (!c.ok || !t || hdr[4] > (1u << 30)) goto fail;
    pix = (uint8_t*)malloc(hdr[4] ? hdr[4] : 1);
    if (!pix || b64_decode(t, len, pix, hdr[4]) != 0) { free(pix); goto fail; }
    put(&texs, hdr, 20);
    put(&texs, pix, hdr[4]);
    put(&texs, &pad, (4 - (hdr[4] & 3)) & 3);
    free(pix);
  }
  put(&all, "SKBXTRC1", 8);
  put_u32(&all, nd);
  put_u32(&all, nt);
  put(&all, texs.d, texs.n);
  put(&all, draws.d, draws.n);
  if (!all.ok || !texs.ok || !draws.ok) goto fail;
  free(texs.d);
  free(draws.d);
  *blob_size = all.n;
  return all.d;
fail:
  free(texs.d);
  free(draws.d);
  free(all.d);
  return NULL;
}

/* CGLTrace::load (cgltrace.cpp:146-165) for the reference's own files. */
int srt_trace_load_cgltrace(const char* path, srt_trace_t* out) {
  size_t size = 0;
  uint64_t blob_size = 0;
  uint8_t *xml, *blob;
  int rc;
  if (!path || !out) return -1;
  memset(out, 0, sizeof(*out));
  xml = read_all(path, &size);
  if (!xml) {
    fprintf(stderr, "couldn't open file: %s!\n", path);
    return -1;
  }
  blob = cgltrace_to_blob((const char*)xml, size, &blob_size);
  free(xml);
  if (!blob) {
    fprintf(stderr, "failed to load archive: %s\n", path);
    return -1;
  }
  rc = srt_trace_from_blob(blob, blob_size, out);
  free(blob);
  return rc;
}
