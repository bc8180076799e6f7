/*
 * draw3d -- the reference's 3D rendering regression app (tests/regression/draw3d/main.cpp) on the
 * CUDA back-end.  Same command line (main.cpp:86-134), same console messages, same exit code
 * convention (number of mismatching pixels when -r is given, main.cpp:505-514).
 *
 * Differences, all forced by what is absent here: the trace is the flat .trc(.gz) re-encoding of a
 * .cgltrace (no Boost), PNG I/O goes through apps/raytrace/png_io.c (no libpng), and the "kernel" is a
 * pipeline descriptor instead of kernel.vxbin (no RISC-V toolchain).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "../raytrace/png_io.h"
#include "draw3d_app.h"

static const char* trace_file = "triangle.trc.gz";
static const char* output_file = "output.png";
static const char* reference_file = NULL;
static int sw_rast = 0, sw_tex = 0, sw_om = 0, wide = 0, gpu_binning = 0;
static uint32_t start_draw = 0, end_draw = 0xffffffffu;
static uint32_t dst_width = 128, dst_height = 128;

static void show_usage(void) {
  printf("Vortex 3D Rendering Test.\n");
  printf("Usage: [-t trace] [-s startdraw] [-e enddraw] [-o output] [-r reference] [-w width] [-h height] "
         "[-x s/w rast] [-y s/w om] [-u s/w tex] [-W wide tile headers] [-g binning on the GPU]\n");
}

static void parse_args(int argc, char** argv) {
  int c;
  while ((c = getopt(argc, argv, "t:s:e:o:r:w:h:k:uxyzWg?")) != -1) {
    switch (c) {
    case 't': trace_file = optarg; break;
    case 's': start_draw = (uint32_t)atoi(optarg); break;
    case 'e': end_draw = (uint32_t)atoi(optarg); break;
    case 'o': output_file = optarg; break;
    case 'r': reference_file = optarg; break;
    case 'w': dst_width = (uint32_t)atoi(optarg); break;
    case 'h': dst_height = (uint32_t)atoi(optarg); break;
    case 'u': sw_tex = 1; break;
    case 'x': sw_rast = 1; break;
    case 'y': sw_om = 1; break;
    case 'W': wide = 1; break;
    case 'g': gpu_binning = 1; break;
    case 'k':
      if (atoi(optarg) != RASTER_TILE_LOGSIZE) {
        printf("Error: the tile size is fixed at 2^%d on this back-end\n", RASTER_TILE_LOGSIZE);
        exit(-1);
      }
      break;
    case 'z': break;
    case '?': show_usage(); exit(0);
    default: show_usage(); exit(-1);
    }
  }
  if (strcmp(output_file, "null") == 0 && reference_file) {
    printf("Error: the output file is missing for reference validation!\n");
    exit(1);
  }
}

int main(int argc, char* argv[]) {
  vx_device_h device = NULL;
  srt_draw3d_t* app = NULL;
  srt_trace_t trace;
  const srt_draw3d_stats_t* st;
  uint32_t d, depth = 0, stencil = 0, blend = 0;
  uint64_t verts = 0, prims = 0;
  int rc;

  parse_args(argc, argv);
  printf("open device connection\n");
  rc = vx_dev_open(&device);
  if (rc != 0) { printf("Error: 'vx_dev_open(&device)' returned %d!\n", rc); return -1; }
  if (srt_trace_load(trace_file, &trace) != 0) { vx_dev_close(device); return -1; }
  for (d = 0; d < trace.num_drawcalls; ++d) {
    depth |= trace.drawcalls[d].states.depth_test;
    stencil |= trace.drawcalls[d].states.stencil_test;
    blend |= trace.drawcalls[d].states.blend_enabled;
    verts += trace.drawcalls[d].num_vertices;
    prims += trace.drawcalls[d].num_primitives;
  }
  printf("CGL Trace: drawcalls=%u, vertices=%llu, primitives=%llu, textures=%u, depth=%u, stencil=%u, blend=%u\n",
         trace.num_drawcalls, (unsigned long long)verts, (unsigned long long)prims, trace.num_textures, depth != 0,
         stencil != 0, blend != 0);
  printf("upload program\n");
  rc = srt_draw3d_open(device, dst_width, dst_height,
                       (sw_rast ? SRT_DRAW3D_SW_RAST : 0) | (sw_om ? SRT_DRAW3D_SW_OM : 0) |
                           (sw_tex ? SRT_DRAW3D_SW_TEX : 0) | (wide ? SRT_DRAW3D_WIDE_TILES : 0) |
                               (gpu_binning ? SRT_DRAW3D_GPU_BINNING : 0),
                       0, 1, &app);
  if (rc == 0) {
    printf("clear depth buffer\nclear destination buffer\n");
    rc = srt_draw3d_clear(app);
  }
  if (rc == 0) {
    printf("render\n");
    rc = srt_draw3d_render(app, &trace, start_draw, end_draw);
  }
  if (rc == 0) {
    st = srt_draw3d_stats(app);
    printf("Binning allocated %llu tiles with %llu total primitives.\n", (unsigned long long)st->tiles,
           (unsigned long long)st->prims);
    printf("Total elapsed time: %lg ms (binning %lg, upload %lg, run %lg)\n", st->bin_ms + st->upload_ms + st->run_ms,
           st->bin_ms, st->upload_ms, st->run_ms);
    if (!getenv("VORTEX_PROFILING")) vx_dump_perf(device, stdout); /* with it set, vx_dev_close dumps */
  }
  if (rc == 0 && strcmp(output_file, "null") != 0) {
    uint32_t* pixels = (uint32_t*)malloc((size_t)dst_width * dst_height * 4);
    printf("save output image\n");
    rc = pixels ? srt_draw3d_read(app, pixels, NULL) : -1;
    if (rc == 0) {
      /* bottom row first: SaveImage(bits + (H-1)*pitch, W, H, -pitch), main.cpp:383-386 */
      const int64_t pitch = (int64_t)dst_width * 4;
      rc = png_save_argb(output_file, (const uint8_t*)pixels + (dst_height - 1) * pitch, dst_width, dst_height, -pitch);
    }
    free(pixels);
  }
  printf("cleanup\n");
  srt_draw3d_close(app);
  srt_trace_free(&trace);
  vx_dev_close(device);
  if (rc != 0) return -1;
  if (reference_file) {
    int errors = png_compare(output_file, reference_file, 1); /* tolerance 1, main.cpp:507 */
    if (errors == 0) {
      printf("PASSED!\n");
    } else {
      printf("FAILED! %d errors.\n", errors);
      return errors;
    }
  }
  return 0;
}
