/*
 * rt_scene.c -- synthetic inputs of the raytrace app.
 *
 * The reference's regression apps generate their own inputs on the host with a fixed seed
 * before uploading them (e.g. tests/regression/demo/main.cpp:60-223 with std::srand(50));
 * this file is that step for the RT app: a Cornell-style box (config 1), a value-noise
 * displaced grid of qx*qy*2 triangles (configs 2-5) and an incoherent ray stress generator
 * (config 5).  All arithmetic is fp32 with no FMA contraction (build flag -ffp-contract=off)
 * and uses only + - * / floorf sqrtf and a 32-bit integer hash, so every compiler produces
 * the same bytes.  Plain C, no CUDA.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "skybox_rt.h"

static uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du;
  x ^= x >> 15; x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
static float unit24(uint32_t h) { return (float)(h >> 8) * (1.0f / 16777216.0f); }
static float rnd4(uint32_t seed, uint32_t a, uint32_t b, uint32_t c) {
  return unit24(mix32(c + mix32(b + mix32(a + mix32(seed)))));
}

static void emit(srt_triangle_t* t, const float* a, const float* b, const float* c, uint32_t albedo) {
  memcpy(t->v0, a, 12);
  memcpy(t->v1, b, 12);
  memcpy(t->v2, c, 12);
  t->albedo = albedo;
}

/* quad (a,b,c,d) -> triangles (a,b,c) and (a,c,d) */
static srt_triangle_t* emit_quad(srt_triangle_t* t, const float* a, const float* b, const float* c,
                                 const float* d, uint32_t albedo) {
  emit(t, a, b, c, albedo);
  emit(t + 1, a, c, d, albedo);
  return t + 2;
}

static srt_triangle_t* emit_box(srt_triangle_t* t, const float lo[3], const float hi[3], uint32_t albedo) {
  float p[8][3];
  int i;
  for (i = 0; i < 8; ++i) { /* bit0 = x, bit1 = y, bit2 = z */
    p[i][0] = (i & 1) ? hi[0] : lo[0];
    p[i][1] = (i & 2) ? hi[1] : lo[1];
    p[i][2] = (i & 4) ? hi[2] : lo[2];
  }
  t = emit_quad(t, p[0], p[1], p[3], p[2], albedo); /* z = lo */
  t = emit_quad(t, p[4], p[6], p[7], p[5], albedo); /* z = hi */
  t = emit_quad(t, p[0], p[2], p[6], p[4], albedo); /* x = lo */
  t = emit_quad(t, p[1], p[5], p[7], p[3], albedo); /* x = hi */
  t = emit_quad(t, p[0], p[4], p[5], p[1], albedo); /* y = lo */
  t = emit_quad(t, p[2], p[3], p[7], p[6], albedo); /* y = hi */
  return t;
}

int srt_scene_cornell(srt_triangle_t** soup, uint32_t* num_tris) {
  static const float c[8][3] = {{-1, -1, -1}, {1, -1, -1}, {-1, 1, -1}, {1, 1, -1},
                                {-1, -1, 1},  {1, -1, 1},  {-1, 1, 1},  {1, 1, 1}};
  static const float tall_lo[3] = {-0.65f, -1.0f, 0.05f}, tall_hi[3] = {-0.10f, 0.25f, 0.60f};
  static const float small_lo[3] = {0.15f, -1.0f, -0.55f}, small_hi[3] = {0.70f, -0.45f, 0.00f};
  static const float l[4][3] = {{-0.25f, 0.98f, -0.25f}, {0.25f, 0.98f, -0.25f},
                                {0.25f, 0.98f, 0.25f},   {-0.25f, 0.98f, 0.25f}};
  srt_triangle_t *base, *t;
  if (!soup || !num_tris) return -1;
  base = t = (srt_triangle_t*)calloc(36, sizeof(*t));
  if (!t) return -1;
  t = emit_quad(t, c[0], c[1], c[5], c[4], 0xffbababau); /* floor   y=-1 */
  t = emit_quad(t, c[2], c[6], c[7], c[3], 0xffbababau); /* ceiling y=+1 */
  t = emit_quad(t, c[4], c[5], c[7], c[6], 0xffbababau); /* back    z=+1 */
  t = emit_quad(t, c[0], c[4], c[6], c[2], 0xffa60d0du); /* left    x=-1 (red)   */
  t = emit_quad(t, c[1], c[3], c[7], c[5], 0xff1f731fu); /* right   x=+1 (green) */
  t = emit_box(t, tall_lo, tall_hi, 0xffb3b3ccu);
  t = emit_box(t, small_lo, small_hi, 0xffccb38cu);
  t = emit_quad(t, l[0], l[1], l[2], l[3], 0xfffffffau);  /* lamp panel */
  *soup = base;
  *num_tris = (uint32_t)(t - base);
  return 0;
}

void srt_camera_cornell(kernel_arg_t* k) {
  static const float pos[3] = {0.0f, 0.0f, -3.6f}, fwd[3] = {0, 0, 1}, right[3] = {1, 0, 0}, up[3] = {0, 1, 0};
  static const float light[3] = {0.0f, 0.9f, 0.0f}, sky[3] = {0, 0, 0};
  memcpy(k->cam_pos, pos, 12); memcpy(k->cam_fwd, fwd, 12);
  memcpy(k->cam_right, right, 12); memcpy(k->cam_up, up, 12);
  memcpy(k->light_pos, light, 12); memcpy(k->sky, sky, 12);
  k->tan_half_fov = 0.36397023f; /* tan 20 deg */
  k->ambient = 0.15f;
  k->ray_eps = 1e-4f;
}

void srt_camera_terrain(kernel_arg_t* k) {
  static const float pos[3] = {0.0f, 38.0f, -78.0f}, fwd[3] = {0.0f, -0.44721360f, 0.89442719f};
  static const float right[3] = {1, 0, 0}, up[3] = {0.0f, 0.89442719f, 0.44721360f};
  static const float light[3] = {35.0f, 30.0f, -25.0f}, sky[3] = {0.55f, 0.70f, 0.90f};
  memcpy(k->cam_pos, pos, 12); memcpy(k->cam_fwd, fwd, 12);
  memcpy(k->cam_right, right, 12); memcpy(k->cam_up, up, 12);
  memcpy(k->light_pos, light, 12); memcpy(k->sky, sky, 12);
  k->tan_half_fov = 0.41421356f; /* tan 22.5 deg */
  k->ambient = 0.12f;
  k->ray_eps = 1e-3f;
}

/* ---- terrain: 5 octaves of smoothstep value noise ---------------------------------- */
static float corner(int32_t ix, int32_t iz, uint32_t seed) {
  return unit24(mix32((uint32_t)ix * 0x9E3779B1u + mix32((uint32_t)iz + mix32(seed))));
}
static float smooth(float t) { return (t * t) * (3.0f - 2.0f * t); }
static float noise2(float x, float z, uint32_t seed) {
  float bx = floorf(x), bz = floorf(z);
  int32_t ix = (int32_t)bx, iz = (int32_t)bz;
  float wx = smooth(x - bx), wz = smooth(z - bz);
  float c00 = corner(ix, iz, seed), c10 = corner(ix + 1, iz, seed);
  float c01 = corner(ix, iz + 1, seed), c11 = corner(ix + 1, iz + 1, seed);
  float lo = c00 + (c10 - c00) * wx;
  float hi = c01 + (c11 - c01) * wx;
  return lo + (hi - lo) * wz;
}
static float height_at(float x, float z, uint32_t seed) {
  float sum = 0.0f, amp = 14.0f, freq = 1.0f / 28.0f;
  uint32_t oct;
  for (oct = 0; oct < 5; ++oct) {
    sum = sum + amp * (noise2(x * freq, z * freq, seed + oct) - 0.5f);
    amp = amp * 0.5f;
    freq = freq * 2.0f;
  }
  return sum;
}
static uint32_t band_colour(float h, uint32_t id, uint32_t seed) {
  static const uint32_t bands[4][3] = {{60, 90, 170}, {190, 170, 120}, {70, 150, 60}, {215, 215, 220}};
  uint32_t jitter = mix32(id + mix32(seed ^ 0xA5A5A5A5u)) & 31u;
  int b = h < -3.0f ? 0 : (h < 1.0f ? 1 : (h < 6.0f ? 2 : 3));
  return 0xff000000u | ((bands[b][0] + jitter) << 16) | ((bands[b][1] + jitter) << 8) | (bands[b][2] + jitter);
}

int srt_scene_terrain(uint32_t qx, uint32_t qy, uint32_t seed, srt_triangle_t** soup, uint32_t* num_tris) {
  uint64_t total = (uint64_t)qx * qy * 2;
  size_t vx = (size_t)qx + 1, vy = (size_t)qy + 1, i, j;
  float *xs, *zs, *hs;
  srt_triangle_t* t;
  float cell, hx, hz;
  if (!soup || !num_tris || qx == 0 || qy == 0 || total > 0x7fffffffu) return -1;
  xs = (float*)malloc(vx * sizeof(float));
  zs = (float*)malloc(vy * sizeof(float));
  hs = (float*)malloc(vx * vy * sizeof(float));
  t = (srt_triangle_t*)malloc((size_t)total * sizeof(*t));
  if (!xs || !zs || !hs || !t) { free(xs); free(zs); free(hs); free(t); return -1; }
  cell = 100.0f / (float)qy;
  hx = 0.5f * (float)qx;
  hz = 0.5f * (float)qy;
  for (i = 0; i < vx; ++i) xs[i] = ((float)i - hx) * cell;
  for (j = 0; j < vy; ++j) zs[j] = ((float)j - hz) * cell;
  for (j = 0; j < vy; ++j)
    for (i = 0; i < vx; ++i) hs[j * vx + i] = height_at(xs[i], zs[j], seed);
  for (j = 0; j < qy; ++j) {
    for (i = 0; i < qx; ++i) {
      float a[3], b[3], c[3], d[3], mean;
      uint32_t id = (uint32_t)((j * qx + i) * 2);
      a[0] = xs[i];     a[1] = hs[j * vx + i];           a[2] = zs[j];
      b[0] = xs[i + 1]; b[1] = hs[j * vx + i + 1];       b[2] = zs[j];
      c[0] = xs[i + 1]; c[1] = hs[(j + 1) * vx + i + 1]; c[2] = zs[j + 1];
      d[0] = xs[i];     d[1] = hs[(j + 1) * vx + i];     d[2] = zs[j + 1];
      mean = 0.25f * (((a[1] + b[1]) + c[1]) + d[1]);
      emit(t + id, a, c, b, band_colour(mean, id, seed));
      emit(t + id + 1, a, d, c, band_colour(mean, id + 1, seed));
    }
  }
  free(xs); free(zs); free(hs);
  *soup = t;
  *num_tris = (uint32_t)total;
  return 0;
}

/* ---- config 5: incoherent rays --------------------------------------------------------- */
void srt_rays_random(const float lo[3], const float hi[3], uint32_t seed, uint64_t first, uint64_t count,
                     srt_ray_t* rays) {
  uint64_t n;
  for (n = 0; n < count; ++n) {
    uint64_t id = first + n;
    uint32_t ka = (uint32_t)id, kb = (uint32_t)(id >> 32) + 0x51ED270Bu, attempt;
    srt_ray_t* r = rays + n;
    int axis;
    for (axis = 0; axis < 3; ++axis) r->o[axis] = lo[axis] + (hi[axis] - lo[axis]) * rnd4(seed, ka, kb, (uint32_t)axis);
    r->d[0] = 0.0f; r->d[1] = 0.0f; r->d[2] = 1.0f;
    for (attempt = 0; attempt < 32; ++attempt) {
      float x = 2.0f * rnd4(seed, ka, kb, 3 + 3 * attempt) - 1.0f;
      float y = 2.0f * rnd4(seed, ka, kb, 4 + 3 * attempt) - 1.0f;
      float z = 2.0f * rnd4(seed, ka, kb, 5 + 3 * attempt) - 1.0f;
      float r2 = (x * x + y * y) + z * z;
      if (r2 <= 1.0f && r2 > 1e-4f) {
        float len = sqrtf(r2);
        r->d[0] = x / len; r->d[1] = y / len; r->d[2] = z / len;
        break;
      }
    }
    r->tmin = 0.0f;
    r->tmax = 1e30f;
  }
}
