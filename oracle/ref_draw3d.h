/*
 * ref_draw3d.h -- CPU oracle for the raster hot path (group R).
 *
 * TEST INFRASTRUCTURE ONLY: loaded by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs, never by the product (skybox_rt_b200/).
 *
 * Plain-C restatement of the reference's draw3d path: host pre-pass (graphics::Binning), per-draw DCR
 * programming (tests/regression/draw3d/main.cpp:171-346) and the device kernel with software raster,
 * texture and output-merger (tests/regression/draw3d/kernel.cpp, gpu_sw.h, sim/common/graphics.cpp).
 *
 * PARITY IS PINNED: this oracle reproduces all 17 golden PNGs of tests/regression/draw3d that have a
 * trace (tests/golden/raster, name_ref_size.npy) with zero differing pixels, and is compared framebuffer- and
 * depth-buffer-exact with the reference's own code compiled natively (oracle/_ref/libref_draw3d.so,
 * built by oracle/Makefile from the sources under /root/reference) at sizes the goldens do not cover.
 *
 * The two entry points have the same signatures as the _ref harness (oracle/ref_draw3d_harness.cpp)
 * so a test can swap one library for the other.
 */
#ifndef REF_DRAW3D_H
#define REF_DRAW3D_H

#include <stdint.h>

#include "../include/skybox_draw3d.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Whole trace: clear, then every draw call in order.  cbuf_out / zbuf_out: width*height u32, row 0 =
 * framebuffer row 0 (the PNG the reference saves is this, flipped).  Either may be NULL. */
int ref_draw3d_render(const uint8_t* blob, uint64_t blob_size, uint32_t width, uint32_t height,
                      uint32_t num_threads, uint32_t tile_logsize, uint32_t* cbuf_out, uint32_t* zbuf_out,
                      double* kernel_seconds);

/* graphics::Binning of draw call `draw`; buffers are caller-sized, sizes in/out.  Returns the tile count. */
int64_t ref_draw3d_binning(const uint8_t* blob, uint64_t blob_size, uint32_t draw, uint32_t width,
                           uint32_t height, uint32_t tile_logsize, uint8_t* tilebuf_out,
                           uint64_t* tilebuf_size, uint8_t* primbuf_out, uint64_t* primbuf_size);

uint64_t ref_draw3d_count_fragments(const uint8_t* blob, uint64_t blob_size, uint32_t width, uint32_t height,
                                    uint32_t num_threads);
/* Test hook: OM DCR values {blend mode, blend func, blend constant, logic op} programmed for every draw whose trace
 * state enables blending (the reference app itself only ever programs MODE_ADD, main.cpp:271-279); NULL clears it. */
void ref_draw3d_set_om_override(const uint32_t* ov);
uint32_t ref_draw3d_kernel_arg_size(void);
uint32_t ref_draw3d_prim_size(void);

/* Sub-stage probes for unit tests. */
uint32_t ref_draw3d_blend(const draw3d_kernel_arg_t* arg, uint32_t src, uint32_t dst);
int ref_draw3d_depth_stencil(const draw3d_kernel_arg_t* arg, uint32_t depth, uint32_t dst_ds, uint32_t* out_ds);
uint32_t ref_draw3d_tex(const draw3d_kernel_arg_t* arg, const uint8_t* texels, int32_t u, int32_t v);
/* tests/regression/tex/kernel.cpp restated: the texture of stage 0 resampled to dst_width x dst_height. */
void ref_draw3d_tex_image(const draw3d_kernel_arg_t* arg, const uint8_t* texels, uint32_t dst_width, uint32_t dst_height,
                          uint32_t num_tasks, int trilinear, uint32_t* out);

#ifdef __cplusplus
}
#endif
#endif
