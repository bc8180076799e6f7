/* png_io.c -- see png_io.h. */
#include "png_io.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

static void put32(uint8_t* p, uint32_t v) { p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v; }
static uint32_t get32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3]; }

static int chunk(FILE* f, const char* tag, const uint8_t* data, uint32_t n) {
  uint8_t hdr[8], crcb[4];
  uint32_t crc;
  put32(hdr, n);
  memcpy(hdr + 4, tag, 4);
  crc = crc32(0, hdr + 4, 4);
  if (n) crc = crc32(crc, data, n);
  put32(crcb, crc);
  return fwrite(hdr, 1, 8, f) == 8 && (n == 0 || fwrite(data, 1, n, f) == n) && fwrite(crcb, 1, 4, f) == 4 ? 0 : -1;
}

int png_save_argb(const char* path, const uint8_t* bits, uint32_t w, uint32_t h, int64_t pitch) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  const size_t row = (size_t)w * 4 + 1;
  uint8_t* raw = (uint8_t*)malloc(row * h);
  uLongf zn = compressBound(row * h);
  uint8_t* z = (uint8_t*)malloc(zn);
  uint8_t ihdr[13];
  FILE* f;
  uint32_t x, y;
  int rc = -1;
  if (!raw || !z) goto out;
  for (y = 0; y < h; ++y) {
    const uint32_t* src = (const uint32_t*)(bits + (int64_t)y * pitch);
    uint8_t* dst = raw + row * y;
    *dst++ = 0; /* filter: none */
    for (x = 0; x < w; ++x) {
      const uint32_t c = src[x];
      *dst++ = (c >> 16) & 255; *dst++ = (c >> 8) & 255; *dst++ = c & 255; *dst++ = c >> 24;
    }
  }
  if (compress2(z, &zn, raw, row * h, 6) != Z_OK) goto out;
  f = fopen(path, "wb");
  if (!f) goto out;
  put32(ihdr, w); put32(ihdr + 4, h);
  ihdr[8] = 8; ihdr[9] = 6; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;
  rc = (fwrite(sig, 1, 8, f) == 8 && chunk(f, "IHDR", ihdr, 13) == 0 && chunk(f, "IDAT", z, (uint32_t)zn) == 0 &&
        chunk(f, "IEND", 0, 0) == 0) ? 0 : -1;
  fclose(f);
out:
  free(raw); free(z);
  return rc;
}

static int paeth(int a, int b, int c) {
  int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

int png_load_argb(const char* path, uint32_t** pixels, uint32_t* width, uint32_t* height) {
  FILE* f = fopen(path, "rb");
  uint8_t *file = 0, *z = 0, *raw = 0;
  uint32_t* out = 0;
  long n;
  size_t pos = 8, zn = 0;
  uint32_t w = 0, h = 0, bpp = 0, x, y;
  int rc = -1;
  if (!f) return -1;
  fseek(f, 0, SEEK_END); n = ftell(f); fseek(f, 0, SEEK_SET);
  file = (uint8_t*)malloc(n > 0 ? n : 1);
  z = (uint8_t*)malloc(n > 0 ? n : 1);
  if (!file || !z || n < 8 || fread(file, 1, n, f) != (size_t)n) goto out;
  while (pos + 12 <= (size_t)n) {
    const uint32_t len = get32(file + pos);
    const uint8_t* tag = file + pos + 4;
    if (pos + 12 + len > (size_t)n) goto out;
    if (!memcmp(tag, "IHDR", 4)) {
      w = get32(tag + 4); h = get32(tag + 8);
      if (tag[12] != 8 || (tag[13] != 6 && tag[13] != 2) || tag[16] != 0) goto out; /* 8-bit RGB(A), no interlace */
      bpp = tag[13] == 6 ? 4 : 3;
    } else if (!memcmp(tag, "IDAT", 4)) {
      memcpy(z + zn, tag + 4, len); zn += len;
    }
    pos += 12 + len;
  }
  if (!w || !h || !bpp) goto out;
  {
    const size_t row = (size_t)w * bpp + 1;
    uLongf rn = row * h;
    raw = (uint8_t*)malloc(rn);
    out = (uint32_t*)malloc((size_t)w * h * 4);
    if (!raw || !out || uncompress(raw, &rn, z, zn) != Z_OK || rn != row * h) goto out;
    for (y = 0; y < h; ++y) {
      uint8_t* cur = raw + row * y + 1;
      const uint8_t* up = y ? raw + row * (y - 1) + 1 : 0;
      const int ft = raw[row * y];
      for (x = 0; x < w * bpp; ++x) {
        const int a = x >= bpp ? cur[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= bpp) ? up[x - bpp] : 0;
        int v = cur[x];
        if (ft == 1) v += a; else if (ft == 2) v += b; else if (ft == 3) v += (a + b) >> 1; else if (ft == 4) v += paeth(a, b, c);
        cur[x] = (uint8_t)v;
      }
      for (x = 0; x < w; ++x) {
        const uint8_t* p = cur + x * bpp;
        out[(size_t)y * w + x] = ((uint32_t)(bpp == 4 ? p[3] : 255) << 24) | (p[0] << 16) | (p[1] << 8) | p[2];
      }
    }
  }
  *pixels = out; *width = w; *height = h;
  out = 0;
  rc = 0;
out:
  fclose(f);
  free(file); free(z); free(raw); free(out);
  return rc;
}

int png_compare(const char* fa, const char* fb, int tol) {
  uint32_t *a = 0, *b = 0, wa, ha, wb, hb;
  int errors = 0;
  size_t i;
  if (png_load_argb(fa, &a, &wa, &ha) != 0 || png_load_argb(fb, &b, &wb, &hb) != 0 || wa != wb || ha != hb) {
    free(a); free(b);
    return -1;
  }
  for (i = 0; i < (size_t)wa * ha; ++i) {
    int k, bad = 0;
    for (k = 0; k < 32; k += 8)
      if (abs((int)((a[i] >> k) & 255) - (int)((b[i] >> k) & 255)) > tol) bad = 1;
    errors += bad;
  }
  free(a); free(b);
  return errors;
}
