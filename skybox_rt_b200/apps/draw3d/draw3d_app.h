/*
 * draw3d_app.h -- host side of the `draw3d` regression app, as a small C library.
 *
 * Follows the reference app tests/regression/draw3d/main.cpp: open -> caps check -> allocate and clear
 * depth/colour buffers (:461-490) -> per draw call: bin, upload tile/primitive/texture buffers, program
 * the raster / output-merger / texture DCRs (with the kernel_arg_t mirrors), upload kernel_arg_t,
 * vx_start, vx_ready_wait (:171-361) -> read the colour buffer back (:380-387).  Everything goes through
 * the vx_* C ABI (include/vortex.h); nothing here knows about CUDA.
 *
 * Two ways to drive it:
 *   srt_draw3d_render()             the reference's loop verbatim: everything per draw, synchronous;
 *   srt_draw3d_prepare() + _submit() the same work split so that a frame can be re-submitted with its
 *                                   buffers resident on the device (what bench.py times as `value`).
 */
#ifndef SRT_DRAW3D_APP_H
#define SRT_DRAW3D_APP_H

#include <stdint.h>

#include "skybox_draw3d.h"
#include "vortex.h"

#ifdef __cplusplus
extern "C" {
#endif

/* flags: the reference's command-line switches (main.cpp:110-118) plus this back-end's extension */
#define SRT_DRAW3D_SW_RAST 0x1 /* -x */
#define SRT_DRAW3D_SW_OM 0x2   /* -y */
#define SRT_DRAW3D_SW_TEX 0x4  /* -u */
#define SRT_DRAW3D_WIDE_TILES 0x8 /* 32-bit tile headers (needed beyond ~64 K tile-primitive pairs) */
#define SRT_DRAW3D_DEVICE_CLEAR 0x20 /* srt_draw3d_clear fills the targets on the device (vx_cuda_fill32) instead of
                                        uploading two host arrays of the clear values like the reference app */
#define SRT_DRAW3D_GPU_BINNING 0x10 /* triangle setup + binning on the device (vx_cuda_draw3d_binning): vertices and
                                       indices are uploaded instead of primbuf / tilebuf; same bytes, same frame */

#define SRT_DRAW3D_AUTO 0x40 /* pick per frame / per draw what is faster on this back-end, same pixels: device-side clear for
                                targets of >= 2^20 pixels, device binning for draws of >= 10^4 triangles, and no
                                vx_ready_wait between the draws of a frame (they queue on the device in order; buffers of
                                a finished draw are freed stream-ordered).  Without it the loop is the reference's.       */

typedef struct srt_draw3d srt_draw3d_t;

typedef struct {
  uint32_t draws_submitted;      /* draws with at least one tile                                  */
  uint64_t prims, tiles, tile_prims; /* totals over the submitted draws                          */
  uint64_t h2d_bytes, d2h_bytes; /* bytes moved by the last render / prepare / read              */
  double bin_ms, upload_ms, run_ms; /* host wall clock of the last render() by phase             */
} srt_draw3d_stats_t;

/* rank / num_ranks: this process renders the tile rows with (tile_y % num_ranks) == rank; 0,1 = all. */
int srt_draw3d_open(vx_device_h device, uint32_t width, uint32_t height, uint32_t flags, uint32_t rank,
                    uint32_t num_ranks, srt_draw3d_t** out);
void srt_draw3d_close(srt_draw3d_t* r);

/* Output-merger state the trace format cannot express.  A .cgltrace carries GL blend factors only, and the reference
 * app always programs VX_OM_BLEND_MODE_ADD with no logic op and no blend constant (main.cpp:271-279), although its
 * output merger implements six equations, sixteen logic ops and four CONST_* factors (sim/common/graphics.cpp:366-524).
 * With an override set, every draw whose trace state enables blending is programmed with these four DCR values
 * instead (VX_DCR_OM_BLEND_MODE / _FUNC / _CONST / VX_DCR_OM_LOGIC_OP, hw/rtl/VX_types.vh:439-458).  NULL clears it. */
typedef struct {
  uint32_t blend_mode;  /* mode_a << 16 | mode_rgb                       */
  uint32_t blend_func;  /* dst_a << 24 | dst_rgb << 16 | src_a << 8 | src_rgb */
  uint32_t blend_const; /* A8R8G8B8                                       */
  uint32_t logic_op;
} srt_om_override_t;
void srt_draw3d_set_om_override(srt_draw3d_t* r, const srt_om_override_t* ov);

/* Host upload of the clear values 0xff000000 / 0xffffffff (main.cpp:45-46, :478-490). */
int srt_draw3d_clear(srt_draw3d_t* r);

/* The reference's render(): draws [start_draw, end_draw] of the trace, each one binned, uploaded,
 * started and waited for in turn. */
int srt_draw3d_render(srt_draw3d_t* r, const srt_trace_t* trace, uint32_t start_draw, uint32_t end_draw);

/* Bin and upload every draw of the trace once and keep the buffers on the device. */
int srt_draw3d_prepare(srt_draw3d_t* r, const srt_trace_t* trace);
/* Re-issue the prepared draws: DCR writes + vx_start per draw, no host<->device traffic, no waiting
 * (call vx_ready_wait afterwards).  `stats` != 0 turns the fragment counters on. */
int srt_draw3d_submit(srt_draw3d_t* r);

/* Colour (and optionally depth) buffer read-back; width*height u32 each, row 0 = framebuffer row 0. */
int srt_draw3d_read(srt_draw3d_t* r, uint32_t* cbuf, uint32_t* zbuf);

/* Colour read-back that returns at once (vx_cuda_copy_from_dev_async): ordered after the draws queued so far, it runs
 * while the next frame renders into ANOTHER srt_draw3d_t; cbuf should be page-locked.  srt_draw3d_read_wait blocks
 * until the ticket's bytes are on the host.  This app object must not be cleared or drawn to before that. */
int srt_draw3d_read_async(srt_draw3d_t* r, uint32_t* cbuf, uint64_t* ticket);
int srt_draw3d_read_wait(srt_draw3d_t* r, uint64_t ticket);

const srt_draw3d_stats_t* srt_draw3d_stats(const srt_draw3d_t* r);
vx_buffer_h srt_draw3d_color_buffer(const srt_draw3d_t* r);
vx_buffer_h srt_draw3d_depth_buffer(const srt_draw3d_t* r);

#ifdef __cplusplus
}
#endif
#endif
