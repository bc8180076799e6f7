/*
 * vortex_cuda.h -- CUDA-specific additions to the vx_* C ABI (no counterpart in the reference).
 *
 * The reference host API (runtime/include/vortex.h) has no notion of a device-side queue that a
 * second library could share.  On B200 the framebuffer gather between GPUs (NCCL) and device-side
 * timing (CUDA events) must be ordered with the kernels vx_start enqueues, so the back-end exposes
 * its launch stream.  Plain C: the stream is passed as void* (a cudaStream_t).
 */
#ifndef SKYBOX_RT_B200_VORTEX_CUDA_H
#define SKYBOX_RT_B200_VORTEX_CUDA_H

#include "vortex.h"

#ifdef __cplusplus
extern "C" {
#endif

/* The CUDA stream vx_start launches on and vx_copy_* are ordered with (owned by the device). */
int vx_cuda_stream(vx_device_h hdevice, void** stream);

/* Page-lock a host range the app reuses for vx_copy_to_dev / vx_copy_from_dev (cudaHostRegister), so
 * those copies run as direct DMA at link rate instead of through the driver's staging buffer.  Optional:
 * the copies work on any host pointer.  The range must stay mapped until vx_cuda_host_unregister. */
int vx_cuda_host_register(vx_device_h hdevice, void* ptr, uint64_t size);
int vx_cuda_host_unregister(vx_device_h hdevice, void* ptr);

/* Totals of the performance monitor (VX_DCR_BASE_MPM_CLASS != 0, include/skybox_rt.h) accumulated over every
 * profiled run since vx_dev_open: out[i] = slot i below for i < count.  vx_dump_perf prints from these. */
enum {
  VX_PERF_CYCLES = 0,   /* SM cycles (elapsed device time x SM clock)            */
  VX_PERF_NS,           /* elapsed device time, ns                                */
  VX_PERF_RUNS,         /* profiled vx_start calls                                */
  VX_PERF_RT_PRIMARY, VX_PERF_RT_SHADOW, VX_PERF_RT_BOUNCE, /* rays by class      */
  VX_PERF_RT_NODES,     /* BVH node records fetched                               */
  VX_PERF_RT_TRIS,      /* triangles tested                                       */
  VX_PERF_RT_NODE_BYTES,/* bytes of those node fetches                            */
  VX_PERF_RT_STORES,    /* pixels / hit records written                           */
  VX_PERF_FRAGMENTS,    /* fragments shaded                                       */
  VX_PERF_TESTS,        /* quad x primitive coverage tests                        */
  VX_PERF_TILES,        /* tiles rendered                                         */
  VX_PERF_TEX_READS,    /* texels fetched                                         */
  VX_PERF_OM_READS,     /* framebuffer words read                                 */
  VX_PERF_OM_WRITES,    /* framebuffer words written                              */
  VX_PERF_RASTER_READS, /* tile headers + primitive ids + primitive records read  */
  VX_PERF_COUNT
};
int vx_cuda_perf_totals(vx_device_h hdevice, uint64_t* out, uint32_t count);

/* CUDA device ordinal the handle was opened on. */
int vx_cuda_device(vx_device_h hdevice, int* ordinal);

#ifdef __cplusplus
}
#endif

#endif
