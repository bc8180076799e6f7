/*
 * vortex.h -- host <-> device C ABI of the B200 back-end (libvortex.so / libvortex-cuda.so).
 *
 * This is the drop-in boundary.  Every entry point below has the name, argument
 * order, argument meaning and return convention (0 = success, non-zero = failure)
 * of the function with the same name in the reference host runtime
 * (/root/reference runtime/include/vortex.h:74-139); the per-function comments cite
 * the reference line each one replaces and the reference driver code whose
 * behaviour it reproduces.  An application written against the reference header
 * (open -> caps -> alloc -> copy -> start -> wait -> copy back) recompiles against
 * this one unchanged.
 *
 * Plain C: opaque handles, pointers and sizes only.  No CUDA or torch types.
 */
#ifndef SKYBOX_RT_B200_VORTEX_H
#define SKYBOX_RT_B200_VORTEX_H

#include <stddef.h>
#include <stdint.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

/* opaque handles (reference vortex.h:25-26) */
typedef void* vx_device_h;
typedef void* vx_buffer_h;

/* ---- vx_dev_caps() ids (reference vortex.h:29-36) -------------------------------- */
enum {
  VX_CAPS_VERSION         = 0x0,
  VX_CAPS_NUM_THREADS     = 0x1, /* lanes per warp: 32 on sm_100a                     */
  VX_CAPS_NUM_WARPS       = 0x2, /* resident warps per SM                             */
  VX_CAPS_NUM_CORES       = 0x3, /* SM count (148 on B200)                            */
  VX_CAPS_CACHE_LINE_SIZE = 0x4, /* 128                                               */
  VX_CAPS_GLOBAL_MEM_SIZE = 0x5, /* HBM bytes                                         */
  VX_CAPS_LOCAL_MEM_SIZE  = 0x6, /* shared memory per SM                              */
  VX_CAPS_ISA_FLAGS       = 0x7
};

/* ---- ISA flag bits reported by VX_CAPS_ISA_FLAGS ---------------------------------
 * Bit positions follow the reference's generated VX_types.h (hw/rtl/VX_types.vh
 * ISA_EXT_* numbering, used at reference vortex.h:50-58): extension bit k lives at
 * bit 32+k.  The CUDA back-end sets TEX|RASTER|OM only when the raster pipeline is
 * compiled in, and adds one bit of its own, VX_ISA_EXT_RT (bit 32+12, unused by the
 * reference), to advertise the ray-tracing pipeline. */
#define VX_ISA_STD_A        (1ull << 0)
#define VX_ISA_STD_C        (1ull << 2)
#define VX_ISA_STD_D        (1ull << 3)
#define VX_ISA_STD_F        (1ull << 5)
#define VX_ISA_STD_I        (1ull << 8)
#define VX_ISA_EXT_ICACHE   (1ull << (32 + 0))
#define VX_ISA_EXT_DCACHE   (1ull << (32 + 1))
#define VX_ISA_EXT_L2CACHE  (1ull << (32 + 2))
#define VX_ISA_EXT_L3CACHE  (1ull << (32 + 3))
#define VX_ISA_EXT_LMEM     (1ull << (32 + 4))
#define VX_ISA_EXT_ZICOND   (1ull << (32 + 5))
#define VX_ISA_EXT_TEX      (1ull << (32 + 6))
#define VX_ISA_EXT_RASTER   (1ull << (32 + 7))
#define VX_ISA_EXT_OM       (1ull << (32 + 8))
#define VX_ISA_EXT_RT       (1ull << (32 + 12))

/* memory types / access flags / timeout (reference vortex.h:60-71) */
#define VX_MEM_TYPE_GLOBAL  0
#define VX_MEM_TYPE_LOCAL   1
#define VX_MEM_READ         0x1
#define VX_MEM_WRITE        0x2
#define VX_MEM_READ_WRITE   0x3
#define VX_MAX_TIMEOUT      (24 * 60 * 60 * 1000) /* ms */

/* ---- device ------------------------------------------------------------------- */

/* reference vortex.h:74, stub/vortex.cpp:58-97.  Opens cuda device $VX_CUDA_DEVICE
 * (else $LOCAL_RANK, else 0), creates the launch stream.  One device per process. */
int vx_dev_open(vx_device_h* hdevice);

/* reference vortex.h:77, stub/vortex.cpp:99-104.  Dumps perf when VORTEX_PROFILING is
 * set, then frees every buffer still owned by the device (apps leak args buffers,
 * draw3d/main.cpp:346) and destroys the stream. */
int vx_dev_close(vx_device_h hdevice);

/* reference vortex.h:80; ids above.  Unknown id returns -1 (simx aborts instead,
 * runtime/simx/vortex.cpp:84-88). */
int vx_dev_caps(vx_device_h hdevice, uint32_t caps_id, uint64_t* value);

/* ---- memory --------------------------------------------------------------------- */

/* reference vortex.h:83, callbacks.inc:57-78.  From the device's stream-ordered pool
 * (cudaMallocAsync on the launch stream), 256 B aligned (>= the reference's 64 B block
 * alignment).  size 0 or NULL handle -> -1. */
int vx_mem_alloc(vx_device_h hdevice, uint64_t size, int flags, vx_buffer_h* hbuffer);

/* reference vortex.h:86.  Fixed-address reservations have no meaning for CUDA virtual
 * addresses; the call allocates `size` bytes and remembers `address` as the buffer's
 * *reported* address only for kernel images (used by vx_upload_kernel_bytes). */
int vx_mem_reserve(vx_device_h hdevice, uint64_t address, uint64_t size, int flags, vx_buffer_h* hbuffer);

/* reference vortex.h:89, callbacks.inc:98-110.  NULL is a successful no-op.  The free is
 * ordered behind the work already queued on the launch stream, so it neither waits for a
 * running kernel nor pulls the memory from under it. */
int vx_mem_free(vx_buffer_h hbuffer);

/* reference vortex.h:92.  Range-checked; access rights are recorded, not enforced. */
int vx_mem_access(vx_buffer_h hbuffer, uint64_t offset, uint64_t size, int flags);

/* reference vortex.h:95.  The device address IS the CUDA device pointer value. */
int vx_mem_address(vx_buffer_h hbuffer, uint64_t* address);

/* reference vortex.h:98 (bytes free / bytes used on the device). */
int vx_mem_info(vx_device_h hdevice, uint64_t* mem_free, uint64_t* mem_used);

/* reference vortex.h:101/104, callbacks.inc:146-169.  Synchronous memcpy semantics:
 * the host buffer is reusable on return.  offset+size > buffer size -> -1.  Ordered
 * after any running vx_start. */
int vx_copy_to_dev(vx_buffer_h hbuffer, const void* host_ptr, uint64_t dst_offset, uint64_t size);
int vx_copy_from_dev(void* host_ptr, vx_buffer_h hbuffer, uint64_t src_offset, uint64_t size);

/* ---- execution -------------------------------------------------------------------- */

/* reference vortex.h:107, simx/vortex.cpp:172-193.  ASYNCHRONOUS: enqueues the pipeline
 * named by the kernel descriptor in `hkernel` with the kernel_arg_t in `harguments` on
 * the device stream and returns.  A second vx_start queues behind the first. */
int vx_start(vx_device_h hdevice, vx_buffer_h hkernel, vx_buffer_h harguments);

/* reference vortex.h:110, simx/vortex.cpp:195-209.  0 when idle, -1 on timeout (ms) or
 * if the kernel faulted. */
int vx_ready_wait(vx_device_h hdevice, uint64_t timeout);

/* reference vortex.h:113/116.  DCR file kept on the host; raster/tex/om state is read
 * from it at vx_start.  Addresses outside the reference's DCR ranges return -1. */
int vx_dcr_read(vx_device_h hdevice, uint32_t addr, uint32_t* value);
int vx_dcr_write(vx_device_h hdevice, uint32_t addr, uint32_t value);

/* reference vortex.h:119, simx/vortex.cpp:224-236: the value of ONE core, no summing (vx_dump_perf adds them up,
 * stub/utils.cpp:159-805).  core_id = SM index; core_id >= VX_CAPS_NUM_CORES returns -1.  VX_CSR_MINSTRET is the
 * work started on that SM (performance monitor on) or the run's total on core 0 and 0 elsewhere (monitor off), so
 * the sum over cores is the total either way; VX_CSR_MCYCLE and the class CSRs 0xB03.. read the same on every
 * core (device-wide units, like the reference's shared raster / tex / om units); the back-end CSRs VX_CSR_RT_* /
 * VX_CSR_RASTER_* (skybox_rt.h, skybox_draw3d.h) are device totals of the last vx_start on every core. */
int vx_mpm_query(vx_device_h hdevice, uint32_t addr, uint32_t core_id, uint64_t* value);

/* ---- utilities (reference vortex.h:124-139, stub/utils.cpp:49-155, :159, :807) ---- */
int vx_upload_kernel_bytes(vx_device_h hdevice, const void* content, uint64_t size, vx_buffer_h* hbuffer);
int vx_upload_kernel_file(vx_device_h hdevice, const char* filename, vx_buffer_h* hbuffer);
int vx_upload_bytes(vx_device_h hdevice, const void* content, uint64_t size, vx_buffer_h* hbuffer);
int vx_upload_file(vx_device_h hdevice, const char* filename, vx_buffer_h* hbuffer);
int vx_check_occupancy(vx_device_h hdevice, uint32_t group_size, uint32_t* max_localmem);
int vx_dump_perf(vx_device_h hdevice, FILE* stream);

#ifdef __cplusplus
}
#endif

#endif /* SKYBOX_RT_B200_VORTEX_H */
