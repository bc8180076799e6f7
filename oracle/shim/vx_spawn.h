/* Native-host stand-in for kernel/include/vx_spawn.h:24-59.  vx_spawn_threads runs kernel_func once
 * per task with blockIdx.x = task id (kernel/src/vx_spawn.c:54-81), here on a pool of host threads.
 * TEST INFRASTRUCTURE ONLY (oracle/_ref build). */
#ifndef ORACLE_SHIM_VX_SPAWN_H
#define ORACLE_SHIM_VX_SPAWN_H
#include <stdint.h>
#include <vx_intrinsics.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef union {
  struct { uint32_t x, y, z; };
  uint32_t m[3];
} dim3_t;
extern __thread dim3_t blockIdx;
typedef void (*vx_kernel_func_cb)(void* arg);
int vx_spawn_threads(uint32_t dimension, const uint32_t* grid_dim, const uint32_t* block_dim,
                     vx_kernel_func_cb kernel_func, const void* arg);
#ifdef __cplusplus
}
#endif
#endif
