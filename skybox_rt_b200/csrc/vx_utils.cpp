// vx_utils.cpp -- the helper half of the host API plus the driver-plugin entry point.
//
// Helpers follow reference runtime/stub/utils.cpp: vx_upload_kernel_bytes (:49-85),
// vx_upload_kernel_file (:87-111), vx_upload_bytes (:113-132), vx_upload_file (:134-158),
// vx_dump_perf (:159-805, reduced to the totals line every app greps), vx_check_occupancy (:807-835).
// vx_dev_init fills the reference's plugin vtable (runtime/common/callbacks.h:23-73) so the same
// library also loads under the reference's own stub as VORTEX_DRIVER=cuda
// (runtime/stub/vortex.cpp:58-82).  Host-only C++: no CUDA calls here.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "skybox_rt.h"
#include "vortex.h"

namespace {

bool read_file(const char* filename, std::vector<char>& out) {
  FILE* f = fopen(filename, "rb");
  if (!f) {
    printf("error: %s not found\n", filename);
    return false;
  }
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out.resize(n > 0 ? (size_t)n : 0);
  size_t got = out.empty() ? 0 : fread(out.data(), 1, out.size(), f);
  fclose(f);
  return got == out.size();
}

}  // namespace

extern "C" {

int vx_upload_kernel_bytes(vx_device_h hdevice, const void* content, uint64_t size, vx_buffer_h* hbuffer) {
  if (nullptr == hdevice || nullptr == content || size <= 8 || nullptr == hbuffer) return -1;
  if (size < 16) return -1;
  uint64_t min_vma, max_vma;
  memcpy(&min_vma, content, 8);
  memcpy(&max_vma, (const char*)content + 8, 8);
  const uint64_t bin_size = size - 16;
  if (max_vma <= min_vma || max_vma - min_vma < bin_size) return -1;
  const uint64_t runtime_size = max_vma - min_vma;
  vx_buffer_h buf;
  int err = vx_mem_reserve(hdevice, min_vma, runtime_size, 0, &buf);
  if (err) return err;
  // image read-only, the rest (globals of a RISC-V image; unused here) read-write
  if ((err = vx_mem_access(buf, 0, bin_size, VX_MEM_READ)) != 0 ||
      (err = vx_mem_access(buf, bin_size, runtime_size - bin_size, VX_MEM_READ_WRITE)) != 0 ||
      (err = vx_copy_to_dev(buf, (const char*)content + 16, 0, bin_size)) != 0) {
    vx_mem_free(buf);
    return err;
  }
  *hbuffer = buf;
  return 0;
}

int vx_upload_kernel_file(vx_device_h hdevice, const char* filename, vx_buffer_h* hbuffer) {
  if (nullptr == hdevice || nullptr == filename || nullptr == hbuffer) return -1;
  std::vector<char> content;
  if (!read_file(filename, content)) return -1;
  return vx_upload_kernel_bytes(hdevice, content.data(), content.size(), hbuffer);
}

int vx_upload_bytes(vx_device_h hdevice, const void* content, uint64_t size, vx_buffer_h* hbuffer) {
  if (nullptr == hdevice || nullptr == content || 0 == size || nullptr == hbuffer) return -1;
  vx_buffer_h buf;
  int err = vx_mem_alloc(hdevice, size, VX_MEM_READ, &buf);
  if (err) return err;
  if ((err = vx_copy_to_dev(buf, content, 0, size)) != 0) {
    vx_mem_free(buf);
    return err;
  }
  *hbuffer = buf;
  return 0;
}

int vx_upload_file(vx_device_h hdevice, const char* filename, vx_buffer_h* hbuffer) {
  if (nullptr == hdevice || nullptr == filename || nullptr == hbuffer) return -1;
  std::vector<char> content;
  if (!read_file(filename, content)) return -1;
  return vx_upload_bytes(hdevice, content.data(), content.size(), hbuffer);
}

int vx_check_occupancy(vx_device_h hdevice, uint32_t group_size, uint32_t* max_localmem) {
  uint64_t warps_per_core, threads_per_warp;
  int err;
  if ((err = vx_dev_caps(hdevice, VX_CAPS_NUM_WARPS, &warps_per_core)) != 0) return err;
  if ((err = vx_dev_caps(hdevice, VX_CAPS_NUM_THREADS, &threads_per_warp)) != 0) return err;
  const uint64_t threads_per_core = warps_per_core * threads_per_warp;
  if (group_size > threads_per_core) {
    printf("Error: cannot schedule kernel with group_size > threads_per_core (%u,%lu)\n", group_size,
           (unsigned long)threads_per_core);
    return -1;
  }
  if (max_localmem) {
    const uint64_t warps_per_group = (group_size + threads_per_warp - 1) / threads_per_warp;
    const uint64_t groups_per_core = warps_per_group ? warps_per_core / warps_per_group : warps_per_core;
    uint64_t local_mem;
    if ((err = vx_dev_caps(hdevice, VX_CAPS_LOCAL_MEM_SIZE, &local_mem)) != 0) return err;
    *max_localmem = (uint32_t)(local_mem / (groups_per_core ? groups_per_core : 1));
  }
  return 0;
}

// The reference prints one "PERF: instrs=, cycles=, IPC=" totals line that its CI greps
// (stub/utils.cpp:800-803).  Here "instructions" are rays; the RT class lines follow.
int vx_dump_perf(vx_device_h hdevice, FILE* stream) {
  if (nullptr == hdevice || nullptr == stream) return -1;
  uint64_t cycles = 0, rays = 0, prim = 0, shadow = 0, bounce = 0, nodes = 0, tris = 0, ns = 0;
  vx_mpm_query(hdevice, VX_CSR_MCYCLE, 0, &cycles);
  vx_mpm_query(hdevice, VX_CSR_MINSTRET, 0, &rays);
  vx_mpm_query(hdevice, VX_CSR_RT_RAYS_PRIMARY, 0, &prim);
  vx_mpm_query(hdevice, VX_CSR_RT_RAYS_SHADOW, 0, &shadow);
  vx_mpm_query(hdevice, VX_CSR_RT_RAYS_BOUNCE, 0, &bounce);
  vx_mpm_query(hdevice, VX_CSR_RT_NODES, 0, &nodes);
  vx_mpm_query(hdevice, VX_CSR_RT_TRIS, 0, &tris);
  vx_mpm_query(hdevice, VX_CSR_RT_ELAPSED_NS, 0, &ns);
  const double ipc = cycles ? (double)rays / (double)cycles : 0.0;
  fprintf(stream, "PERF: rt: rays_primary=%lu, rays_shadow=%lu, rays_bounce=%lu, nodes=%lu, tris=%lu, elapsed_ns=%lu\n",
          (unsigned long)prim, (unsigned long)shadow, (unsigned long)bounce, (unsigned long)nodes,
          (unsigned long)tris, (unsigned long)ns);
  if (ns) fprintf(stream, "PERF: rt: Mrays/s=%f\n", (double)rays * 1e3 / (double)ns);
  fprintf(stream, "PERF: instrs=%lu, cycles=%lu, IPC=%f\n", (unsigned long)rays, (unsigned long)cycles, ipc);
  fflush(stream);
  return 0;
}

// ---- driver plugin entry (reference runtime/common/callbacks.h:23-73) ----------------------
typedef struct {
  int (*dev_open)(vx_device_h*);
  int (*dev_close)(vx_device_h);
  int (*dev_caps)(vx_device_h, uint32_t, uint64_t*);
  int (*mem_alloc)(vx_device_h, uint64_t, int, vx_buffer_h*);
  int (*mem_reserve)(vx_device_h, uint64_t, uint64_t, int, vx_buffer_h*);
  int (*mem_free)(vx_buffer_h);
  int (*mem_access)(vx_buffer_h, uint64_t, uint64_t, int);
  int (*mem_address)(vx_buffer_h, uint64_t*);
  int (*mem_info)(vx_device_h, uint64_t*, uint64_t*);
  int (*copy_to_dev)(vx_buffer_h, const void*, uint64_t, uint64_t);
  int (*copy_from_dev)(void*, vx_buffer_h, uint64_t, uint64_t);
  int (*start)(vx_device_h, vx_buffer_h, vx_buffer_h);
  int (*ready_wait)(vx_device_h, uint64_t);
  int (*dcr_read)(vx_device_h, uint32_t, uint32_t*);
  int (*dcr_write)(vx_device_h, uint32_t, uint32_t);
  int (*mpm_query)(vx_device_h, uint32_t, uint32_t, uint64_t*);
} callbacks_t;

int vx_dev_init(callbacks_t* cb) {
  if (nullptr == cb) return -1;
  cb->dev_open = vx_dev_open;
  cb->dev_close = vx_dev_close;
  cb->dev_caps = vx_dev_caps;
  cb->mem_alloc = vx_mem_alloc;
  cb->mem_reserve = vx_mem_reserve;
  cb->mem_free = vx_mem_free;
  cb->mem_access = vx_mem_access;
  cb->mem_address = vx_mem_address;
  cb->mem_info = vx_mem_info;
  cb->copy_to_dev = vx_copy_to_dev;
  cb->copy_from_dev = vx_copy_from_dev;
  cb->start = vx_start;
  cb->ready_wait = vx_ready_wait;
  cb->dcr_read = vx_dcr_read;
  cb->dcr_write = vx_dcr_write;
  cb->mpm_query = vx_mpm_query;
  return 0;
}

}  // extern "C"
