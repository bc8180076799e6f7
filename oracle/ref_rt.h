/*
 * ref_rt.h -- internal helpers shared by the oracle's translation units.
 * TEST INFRASTRUCTURE ONLY (see ref_rt.c).  Float expressions: IEEE binary32, RN, no FMA.
 */
#ifndef ORACLE_REF_RT_H
#define ORACLE_REF_RT_H

#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/skybox_rt.h"

#define ORT_TMAX 1e30f
#define ORT_PAD 1.00012207031250f /* 1 + 2^-13: slack of the conservative slab test */

typedef struct { float x, y, z; } v3;

static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vscale(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 vdivs(v3 a, float s) { return V(a.x / s, a.y / s, a.z / s); }
static inline float vdot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline v3 vcross(v3 a, v3 b) {
  return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline v3 vload(const float* p) { return V(p[0], p[1], p[2]); }

/* ---------------------------------------------------------------- integer hash / RNG */
static inline uint32_t hash32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du;
  x ^= x >> 15; x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
/* uniform in [0,1): 24 random bits, keyed (seed, a, b, c) */
static inline float rnd01(uint32_t seed, uint32_t a, uint32_t b, uint32_t c) {
  uint32_t h = hash32(c + hash32(b + hash32(a + hash32(seed))));
  return (float)(h >> 8) * (1.0f / 16777216.0f);
}


/* per-run counters (rays by class; nodes visited / triangles tested by the BVH2 traversal) */
typedef struct {
  uint64_t rays_primary, rays_shadow, rays_bounce;
  uint64_t nodes_visited, tris_tested;
} ort_counters_t;

/* ref_rt.c */
int ort_scene_cornell(srt_triangle_t** soup, uint32_t* num_tris);
int ort_scene_terrain(uint32_t qx, uint32_t qy, uint32_t seed, srt_triangle_t** soup, uint32_t* num_tris);
void ort_camera_cornell(kernel_arg_t* a);
void ort_camera_terrain(kernel_arg_t* a);
void ort_rays_random(const float lo[3], const float hi[3], uint32_t seed, uint64_t first, uint64_t count,
                     srt_ray_t* rays);
/* ref_rt_bvh.c */
int ort_bvh2_build(const srt_triangle_t* soup, uint32_t num_tris, srt_bvh2_t* out);
void ort_bvh2_free(srt_bvh2_t* bvh);
int ort_intersect_tri(const srt_tri_rec_t* tr, v3 o, v3 d, float* t, float* u, float* v);
void ort_closest_hit(const srt_bvh2_node_t* nodes, const srt_tri_rec_t* tris, const srt_ray_t* ray,
                     srt_hit_t* hit, ort_counters_t* ctr);
int ort_any_hit(const srt_bvh2_node_t* nodes, const srt_tri_rec_t* tris, const srt_ray_t* ray,
                ort_counters_t* ctr);
void ort_closest_hit_brute(const srt_tri_rec_t* tris, uint32_t num_tris, const srt_ray_t* ray, srt_hit_t* hit);
int ort_any_hit_brute(const srt_tri_rec_t* tris, uint32_t num_tris, const srt_ray_t* ray);
/* ref_rt_kernel.c */
int ort_render(const kernel_arg_t* arg, uint32_t num_threads, ort_counters_t* ctr);

#endif
