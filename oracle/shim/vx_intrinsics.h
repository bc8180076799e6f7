/* Native-host stand-in for the reference's kernel/include/vx_intrinsics.h (RISC-V inline asm,
 * :39-242) so that tests/regression/draw3d/kernel.cpp compiles with g++ for the host cores.
 * TEST INFRASTRUCTURE ONLY (oracle/_ref build); nothing under skybox_rt_b200/ includes this.
 * Only what draw3d/kernel.cpp touches is provided:
 *   csr_read(VX_CSR_MSCRATCH)  -> the kernel argument pointer  (reference: hw/rtl/VX_types.vh:60)
 *   vx_rast / vx_tex / vx_om   -> unreachable with -x -y -u (software raster/tex/om), abort otherwise */
#ifndef ORACLE_SHIM_VX_INTRINSICS_H
#define ORACLE_SHIM_VX_INTRINSICS_H
#include <stdint.h>
#include <stdlib.h>
#include <VX_types.h>

#define __UNIFORM__

#ifdef __cplusplus
extern "C" {
#endif
extern uintptr_t ref_shim_mscratch;
static inline uintptr_t ref_shim_csr_read(unsigned csr) {
  if (csr == VX_CSR_MSCRATCH) return ref_shim_mscratch;
  abort();
}
#define csr_read(csr) ref_shim_csr_read(csr)
static inline unsigned vx_tex(unsigned, unsigned, unsigned, unsigned) { abort(); }
static inline void vx_om(unsigned, unsigned, unsigned, unsigned, unsigned) { abort(); }
static inline unsigned vx_rast() { abort(); }
#ifdef __cplusplus
}
#endif
#endif
