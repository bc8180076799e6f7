/*
 * ref_rt.c -- CPU ORACLE for the ray-tracing hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this; the product (skybox_rt_b200/) never does.
 *
 * PARITY UNPINNED BY THE REFERENCE: /root/reference contains no ray-tracing extension, no
 * BVH code and no RT golden image, so there is no reference vector
 * this file can be checked against.  It is instead the *definition* of the RT path
 * (SURVEY.md section 8a rows T1-T5), written in the reference's regression-kernel idiom:
 *   - one POD kernel_arg_t shared with the host      (tests/regression/draw3d/common.h:18-34)
 *   - kernel_body(arg) called once per task with blockIdx.x = task id
 *                                                    (kernel/src/vx_spawn.c:54-81)
 *   - a task is a strip of rows, tile_height = ceil(H / num_tasks)
 *                                                    (tests/regression/tex/kernel.cpp:44-46,108)
 *   - pixels stored as A8R8G8B8 at cbuf + x*stride + y*pitch, row 0 = bottom of the image
 *                                                    (tests/regression/draw3d/main.cpp:380-387)
 *   - float -> byte by truncation                    (tests/regression/draw3d/kernel.cpp:94-98)
 * It pins itself three ways (tests/test_oracle_rt.py): analytic single-triangle / box
 * known answers, brute-force all-triangle search vs BVH traversal, and a frozen hash of the
 * 64x64 Cornell render (tests/golden/).
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC -pthread (oracle/Makefile).
 * Every float expression below is evaluated left to right exactly as parenthesised, in
 * IEEE binary32 with round-to-nearest, no FMA.
 */
#include <pthread.h>

#include "ref_rt.h"

/* ================================================================ scenes ========== */
static void put_tri(srt_triangle_t* t, v3 a, v3 b, v3 c, uint32_t albedo) {
  t->v0[0] = a.x; t->v0[1] = a.y; t->v0[2] = a.z;
  t->v1[0] = b.x; t->v1[1] = b.y; t->v1[2] = b.z;
  t->v2[0] = c.x; t->v2[1] = c.y; t->v2[2] = c.z;
  t->albedo = albedo;
}
static uint32_t put_quad(srt_triangle_t* t, v3 a, v3 b, v3 c, v3 d, uint32_t albedo) {
  put_tri(t + 0, a, b, c, albedo);
  put_tri(t + 1, a, c, d, albedo);
  return 2;
}
/* axis-aligned box [lo,hi]: 6 faces, 12 triangles */
static uint32_t put_box(srt_triangle_t* t, v3 lo, v3 hi, uint32_t albedo) {
  uint32_t n = 0;
  v3 p000 = V(lo.x, lo.y, lo.z), p100 = V(hi.x, lo.y, lo.z), p010 = V(lo.x, hi.y, lo.z),
     p110 = V(hi.x, hi.y, lo.z), p001 = V(lo.x, lo.y, hi.z), p101 = V(hi.x, lo.y, hi.z),
     p011 = V(lo.x, hi.y, hi.z), p111 = V(hi.x, hi.y, hi.z);
  n += put_quad(t + n, p000, p100, p110, p010, albedo); /* -z */
  n += put_quad(t + n, p001, p011, p111, p101, albedo); /* +z */
  n += put_quad(t + n, p000, p010, p011, p001, albedo); /* -x */
  n += put_quad(t + n, p100, p101, p111, p110, albedo); /* +x */
  n += put_quad(t + n, p000, p001, p101, p100, albedo); /* -y */
  n += put_quad(t + n, p010, p110, p111, p011, albedo); /* +y */
  return n;
}

/* Config 1: Cornell-style box, 36 triangles. */
int ort_scene_cornell(srt_triangle_t** soup, uint32_t* num_tris) {
  srt_triangle_t* t = (srt_triangle_t*)calloc(36, sizeof(*t));
  uint32_t n = 0;
  const uint32_t white = 0xffbababau, red = 0xffa60d0du, green = 0xff1f731fu,
                 lamp = 0xfffffffau, tall = 0xffb3b3ccu, small = 0xffccb38cu;
  if (!t) return -1;
  n += put_quad(t + n, V(-1, -1, -1), V(1, -1, -1), V(1, -1, 1), V(-1, -1, 1), white);  /* floor   */
  n += put_quad(t + n, V(-1, 1, -1), V(-1, 1, 1), V(1, 1, 1), V(1, 1, -1), white);      /* ceiling */
  n += put_quad(t + n, V(-1, -1, 1), V(1, -1, 1), V(1, 1, 1), V(-1, 1, 1), white);      /* back    */
  n += put_quad(t + n, V(-1, -1, -1), V(-1, -1, 1), V(-1, 1, 1), V(-1, 1, -1), red);    /* left    */
  n += put_quad(t + n, V(1, -1, -1), V(1, 1, -1), V(1, 1, 1), V(1, -1, 1), green);      /* right   */
  n += put_box(t + n, V(-0.65f, -1.0f, 0.05f), V(-0.10f, 0.25f, 0.60f), tall);
  n += put_box(t + n, V(0.15f, -1.0f, -0.55f), V(0.70f, -0.45f, 0.00f), small);
  n += put_quad(t + n, V(-0.25f, 0.98f, -0.25f), V(0.25f, 0.98f, -0.25f), V(0.25f, 0.98f, 0.25f),
                V(-0.25f, 0.98f, 0.25f), lamp);
  *soup = t;
  *num_tris = n;
  return 0;
}

void ort_camera_cornell(kernel_arg_t* a) {
  a->cam_pos[0] = 0.0f; a->cam_pos[1] = 0.0f; a->cam_pos[2] = -3.6f;
  a->cam_fwd[0] = 0.0f; a->cam_fwd[1] = 0.0f; a->cam_fwd[2] = 1.0f;
  a->cam_right[0] = 1.0f; a->cam_right[1] = 0.0f; a->cam_right[2] = 0.0f;
  a->cam_up[0] = 0.0f; a->cam_up[1] = 1.0f; a->cam_up[2] = 0.0f;
  a->tan_half_fov = 0.36397023f; /* tan(20 deg) */
  a->light_pos[0] = 0.0f; a->light_pos[1] = 0.9f; a->light_pos[2] = 0.0f;
  a->ambient = 0.15f;
  a->sky[0] = 0.0f; a->sky[1] = 0.0f; a->sky[2] = 0.0f;
  a->ray_eps = 1e-4f;
}

/* value noise on the integer lattice, smoothstep-interpolated; all fp32, no FMA */
static float lattice(int32_t ix, int32_t iz, uint32_t seed) {
  uint32_t h = hash32((uint32_t)ix * 0x9E3779B1u + hash32((uint32_t)iz + hash32(seed)));
  return (float)(h >> 8) * (1.0f / 16777216.0f);
}
static float vnoise(float x, float z, uint32_t seed) {
  float fx = floorf(x), fz = floorf(z);
  int32_t ix = (int32_t)fx, iz = (int32_t)fz;
  float tx = x - fx, tz = z - fz;
  float sx = (tx * tx) * (3.0f - 2.0f * tx);
  float sz = (tz * tz) * (3.0f - 2.0f * tz);
  float a = lattice(ix, iz, seed), b = lattice(ix + 1, iz, seed);
  float c = lattice(ix, iz + 1, seed), d = lattice(ix + 1, iz + 1, seed);
  float ab = a + (b - a) * sx;
  float cd = c + (d - c) * sx;
  return ab + (cd - ab) * sz;
}
static float terrain_height(float x, float z, uint32_t seed) {
  float h = 0.0f, amp = 14.0f, freq = 1.0f / 28.0f;
  int o;
  for (o = 0; o < 5; ++o) {
    h = h + amp * (vnoise(x * freq, z * freq, seed + (uint32_t)o) - 0.5f);
    amp = amp * 0.5f;
    freq = freq * 2.0f;
  }
  return h;
}
static uint32_t terrain_albedo(float h, uint32_t id, uint32_t seed) {
  uint32_t v = hash32(id + hash32(seed ^ 0xA5A5A5A5u)) & 31u;
  uint32_t r, g, b;
  if (h < -3.0f) { r = 60; g = 90; b = 170; }
  else if (h < 1.0f) { r = 190; g = 170; b = 120; }
  else if (h < 6.0f) { r = 70; g = 150; b = 60; }
  else { r = 215; g = 215; b = 220; }
  return 0xff000000u | ((r + v) << 16) | ((g + v) << 8) | (b + v);
}

/* Configs 2-5: displaced grid, quads_x * quads_y * 2 triangles, 100 units deep in z. */
int ort_scene_terrain(uint32_t qx, uint32_t qy, uint32_t seed, srt_triangle_t** soup, uint32_t* num_tris) {
  uint64_t n = (uint64_t)qx * qy * 2;
  srt_triangle_t* t;
  float cell, halfx, halfz;
  uint32_t i, j;
  float* row0;
  float* row1;
  if (qx == 0 || qy == 0 || n > 0x7fffffffu) return -1;
  t = (srt_triangle_t*)malloc(n * sizeof(*t));
  row0 = (float*)malloc(((size_t)qx + 1) * sizeof(float));
  row1 = (float*)malloc(((size_t)qx + 1) * sizeof(float));
  if (!t || !row0 || !row1) { free(t); free(row0); free(row1); return -1; }
  cell = 100.0f / (float)qy;
  halfx = 0.5f * (float)qx;
  halfz = 0.5f * (float)qy;
  for (i = 0; i <= qx; ++i) row0[i] = terrain_height(((float)i - halfx) * cell, (0.0f - halfz) * cell, seed);
  for (j = 0; j < qy; ++j) {
    float z0 = ((float)j - halfz) * cell, z1 = ((float)(j + 1) - halfz) * cell;
    for (i = 0; i <= qx; ++i) row1[i] = terrain_height(((float)i - halfx) * cell, z1, seed);
    for (i = 0; i < qx; ++i) {
      float x0 = ((float)i - halfx) * cell, x1 = ((float)(i + 1) - halfx) * cell;
      v3 a = V(x0, row0[i], z0), b = V(x1, row0[i + 1], z0), c = V(x1, row1[i + 1], z1), d = V(x0, row1[i], z1);
      uint32_t id = (j * qx + i) * 2;
      float hm = 0.25f * (((a.y + b.y) + c.y) + d.y);
      put_tri(t + id, a, c, b, terrain_albedo(hm, id, seed));
      put_tri(t + id + 1, a, d, c, terrain_albedo(hm, id + 1, seed));
    }
    { float* tmp = row0; row0 = row1; row1 = tmp; }
  }
  free(row0); free(row1);
  *soup = t;
  *num_tris = (uint32_t)n;
  return 0;
}

void ort_camera_terrain(kernel_arg_t* a) {
  /* eye (0,38,-78) looking at (0,-4,6): fwd = normalize(0,-42,84) = (0,-1,2)/sqrt(5) */
  a->cam_pos[0] = 0.0f; a->cam_pos[1] = 38.0f; a->cam_pos[2] = -78.0f;
  a->cam_fwd[0] = 0.0f; a->cam_fwd[1] = -0.44721360f; a->cam_fwd[2] = 0.89442719f;
  a->cam_right[0] = 1.0f; a->cam_right[1] = 0.0f; a->cam_right[2] = 0.0f;
  a->cam_up[0] = 0.0f; a->cam_up[1] = 0.89442719f; a->cam_up[2] = 0.44721360f;
  a->tan_half_fov = 0.41421356f; /* tan(22.5 deg) */
  a->light_pos[0] = 35.0f; a->light_pos[1] = 30.0f; a->light_pos[2] = -25.0f;
  a->ambient = 0.12f;
  a->sky[0] = 0.55f; a->sky[1] = 0.70f; a->sky[2] = 0.90f;
  a->ray_eps = 1e-3f;
}

/* Config 5 ray generator: ray i depends only on (seed, i). */
void ort_rays_random(const float lo[3], const float hi[3], uint32_t seed, uint64_t first, uint64_t count,
                     srt_ray_t* rays) {
  uint64_t k;
  for (k = 0; k < count; ++k) {
    uint64_t idx = first + k;
    uint32_t a = (uint32_t)idx, b = (uint32_t)(idx >> 32) + 0x51ED270Bu;
    srt_ray_t* r = rays + k;
    uint32_t j, tries = 0;
    float dx = 0.0f, dy = 0.0f, dz = 1.0f;
    for (j = 0; j < 3; ++j) {
      float u = rnd01(seed, a, b, j);
      r->o[j] = lo[j] + (hi[j] - lo[j]) * u;
    }
    for (tries = 0; tries < 32; ++tries) {
      float x = 2.0f * rnd01(seed, a, b, 3 + 3 * tries) - 1.0f;
      float y = 2.0f * rnd01(seed, a, b, 4 + 3 * tries) - 1.0f;
      float z = 2.0f * rnd01(seed, a, b, 5 + 3 * tries) - 1.0f;
      float r2 = (x * x + y * y) + z * z;
      if (r2 <= 1.0f && r2 > 1e-4f) {
        float len = sqrtf(r2);
        dx = x / len; dy = y / len; dz = z / len;
        break;
      }
    }
    r->d[0] = dx; r->d[1] = dy; r->d[2] = dz;
    r->tmin = 0.0f;
    r->tmax = ORT_TMAX;
  }
}
