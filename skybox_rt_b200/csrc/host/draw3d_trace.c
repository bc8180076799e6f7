/*
 * draw3d_trace.c -- loader for draw-call traces (host library, plain C).
 *
 * Takes the place of cocogfx::CGLTrace::load (reference third_party/cocogfx/src/cgltrace.cpp:146-165),
 * which needs Boost.Serialization.  The on-disk form here is the flat "SKBXTRC1" layout produced by
 * tests/golden/make_raster_fixtures.py from the reference's own .cgltrace files (same fields, same
 * order as the reference's serialize() functions, cgltrace.cpp:24-142); gzip'ed files are accepted
 * (zlib's gzread passes plain files through).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include "skybox_draw3d.h"

static int parse(srt_trace_t* t) {
  const uint8_t* p = t->blob;
  const uint8_t* end = t->blob + t->blob_size;
  uint32_t i;
  if (t->blob_size < 16 || memcmp(p, "SKBXTRC1", 8) != 0) return -1;
  memcpy(&t->num_drawcalls, p + 8, 4);
  memcpy(&t->num_textures, p + 12, 4);
  p += 16;
  if (t->num_drawcalls > (1u << 24) || t->num_textures > (1u << 24)) return -1;
  t->drawcalls = (srt_drawcall_t*)calloc(t->num_drawcalls ? t->num_drawcalls : 1, sizeof(srt_drawcall_t));
  t->textures = (srt_texture_t*)calloc(t->num_textures ? t->num_textures : 1, sizeof(srt_texture_t));
  if (!t->drawcalls || !t->textures) return -1;
  for (i = 0; i < t->num_textures; ++i) {
    srt_texture_t* x = &t->textures[i];
    if (end - p < 20) return -1;
    memcpy(x, p, 20); /* id, format, width, height, size */
    p += 20;
    if ((uint64_t)(end - p) < (((uint64_t)x->size + 3) & ~3ull)) return -1;
    x->pixels = p;
    p += ((uint64_t)x->size + 3) & ~3ull;
    if ((uint64_t)x->width * x->height * srt_format_bpp(x->format) > x->size) return -1;
  }
  for (i = 0; i < t->num_drawcalls; ++i) {
    srt_drawcall_t* d = &t->drawcalls[i];
    uint32_t k;
    if (end - p < 96 + 16 + 4 + 24 + 8) return -1;
    memcpy(&d->states, p, 96 + 16); /* 24 u32 + envcolor */
    p += 112;
    memcpy(&d->texture_id, p, 4);
    p += 4;
    memcpy(&d->viewport, p, 24);
    p += 24;
    memcpy(&d->num_vertices, p, 4);
    memcpy(&d->num_primitives, p + 4, 4);
    p += 8;
    if ((uint64_t)(end - p) < (uint64_t)d->num_vertices * sizeof(srt_vertex_t)) return -1;
    d->vertices = (const srt_vertex_t*)p; /* offsets are multiples of 4: blob is malloc-aligned */
    p += (uint64_t)d->num_vertices * sizeof(srt_vertex_t);
    if ((uint64_t)(end - p) < (uint64_t)d->num_primitives * sizeof(srt_primitive_t)) return -1;
    d->primitives = (const srt_primitive_t*)p;
    p += (uint64_t)d->num_primitives * sizeof(srt_primitive_t);
    for (k = 0; k < d->num_primitives; ++k)
      if (d->primitives[k].i0 >= d->num_vertices || d->primitives[k].i1 >= d->num_vertices ||
          d->primitives[k].i2 >= d->num_vertices)
        return -1; /* the reference's vertices.at(i) would throw */
  }
  return 0;
}

int srt_trace_from_blob(const void* blob, uint64_t size, srt_trace_t* out) {
  if (!blob || !out) return -1;
  memset(out, 0, sizeof(*out));
  out->blob = (uint8_t*)malloc(size ? size : 1);
  if (!out->blob) return -1;
  memcpy(out->blob, blob, size);
  out->blob_size = size;
  if (parse(out) != 0) {
    srt_trace_free(out);
    return -1;
  }
  return 0;
}

int srt_trace_load(const char* path, srt_trace_t* out) {
  gzFile f;
  size_t plen = path ? strlen(path) : 0;
  if (plen > 9 && strcmp(path + plen - 9, ".cgltrace") == 0) return srt_trace_load_cgltrace(path, out);
  uint64_t cap = 1 << 20, n = 0;
  uint8_t* buf;
  int got, rc;
  if (!path || !out) return -1;
  memset(out, 0, sizeof(*out));
  f = gzopen(path, "rb");
  if (!f) {
    fprintf(stderr, "couldn't open file: %s!\n", path); /* cgltrace.cpp:149 */
    return -1;
  }
  buf = (uint8_t*)malloc(cap);
  while (buf && (got = gzread(f, buf + n, (unsigned)(cap - n))) > 0) {
    n += (uint64_t)got;
    if (n == cap) {
      cap *= 2;
      buf = (uint8_t*)realloc(buf, cap);
    }
  }
  gzclose(f);
  if (!buf) return -1;
  out->blob = buf;
  out->blob_size = n;
  rc = parse(out);
  if (rc != 0) {
    fprintf(stderr, "failed to load trace: %s\n", path);
    srt_trace_free(out);
  }
  return rc;
}

void srt_trace_free(srt_trace_t* t) {
  if (!t) return;
  free(t->drawcalls);
  free(t->textures);
  free(t->blob);
  memset(t, 0, sizeof(*t));
}

const srt_texture_t* srt_trace_texture(const srt_trace_t* t, uint32_t id) {
  uint32_t i;
  for (i = 0; t && i < t->num_textures; ++i)
    if (t->textures[i].id == id) return &t->textures[i];
  return NULL;
}

const srt_drawcall_t* srt_drawcall_at(const srt_trace_t* t, uint32_t index) {
  return (t && index < t->num_drawcalls) ? &t->drawcalls[index] : NULL;
}

/* Format::GetInfo(format).BytePerPixel for the formats the texture unit decodes (format.hpp) */
uint32_t srt_format_bpp(uint32_t format) {
  switch (format) {
  case SRT_FORMAT_A8R8G8B8: return 4;
  case SRT_FORMAT_R5G6B5:
  case SRT_FORMAT_A1R5G5B5:
  case SRT_FORMAT_A4R4G4B4:
  case SRT_FORMAT_A8L8: return 2;
  case SRT_FORMAT_L8:
  case SRT_FORMAT_A8: return 1;
  default: return 0;
  }
}

/* graphics::toVXFormat, sim/common/gfxutil.cpp:280-296 */
uint32_t srt_format_to_vx(uint32_t format) {
  switch (format) {
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
