// vx_cuda.cu -- the vx_* host API over the CUDA runtime (libvortex.so for B200).
//
// Each function reproduces the argument checks, return codes and ordering rules of the reference
// path  stub forwarder -> callbacks_t lambda -> vx_device method :
//   runtime/stub/vortex.cpp:58-182, runtime/common/callbacks.inc:20-225,
//   runtime/simx/vortex.cpp:32-246 (the driver used as the behavioural template).
// No CPU fallback: without a usable CUDA device vx_dev_open fails.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <new>

#include "vortex_cuda.h"
#include "vx_internal.h"

namespace srt {
int rt_read_counters(vx_device* dev);
}

namespace {

inline vx_device* as_dev(vx_device_h h) { return reinterpret_cast<vx_device*>(h); }
inline vx_buffer* as_buf(vx_buffer_h h) { return reinterpret_cast<vx_buffer*>(h); }

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  if (!v || !*v) return dflt;
  return atoi(v);
}

// first OM DCR + 18 == end of every range the reference decodes (hw/rtl/VX_types.vh:22-28,
// :332-340, :352-359, :439-458 with VX_TEX_LOD_MAX = 15)
constexpr uint32_t kRefDcrBegin = 0x001;
constexpr uint32_t kRefDcrEnd = 0x006 + (6 + 16) + 6 + 18;

bool dcr_addr_ok(uint32_t addr) {
  return (addr >= kRefDcrBegin && addr < kRefDcrEnd) || (addr >= VX_DCR_RT_BASE && addr < VX_DCR_RT_END);
}

void unlink_buffer(vx_buffer* b) {
  vx_device* d = b->dev;
  if (b->prev) b->prev->next = b->next; else d->buffers = b->next;
  if (b->next) b->next->prev = b->prev;
  d->bytes_used -= b->size;
}

// Buffers come from the device's stream-ordered pool (cudaMallocAsync on the launch stream, release threshold
// raised in vx_dev_open): an allocation is a pool hit instead of a driver call, and a free is ordered behind
// the kernels and copies already queued, so vx_mem_free does not have to wait for them.  The per-draw
// alloc / upload / free pattern of the reference's apps (draw3d/main.cpp:198-211, :346) costs microseconds.
int destroy_buffer(vx_buffer* b) {
  srt::rt_buffer_freed(b->dev, b);
  unlink_buffer(b);
  cudaError_t e = cudaFreeAsync(b->dptr, b->dev->stream);
  free(b->shadow);
  delete b;
  return e == cudaSuccess ? 0 : -1;
}

int alloc_buffer(vx_device* dev, uint64_t size, int flags, uint64_t reported, vx_buffer** out) {
  void* p = nullptr;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  VX_CUDA_OK(cudaMallocAsync(&p, size, dev->stream));
  vx_buffer* b = new (std::nothrow) vx_buffer();
  if (!b) { cudaFreeAsync(p, dev->stream); return -1; }
  b->dev = dev;
  b->dptr = p;
  b->size = size;
  b->reported_addr = reported ? reported : (uint64_t)(uintptr_t)p;
  b->flags = flags;
  b->shadow = size <= kShadowMaxBytes ? (uint8_t*)calloc(1, size) : nullptr;
  b->generation = 0;
  b->uid = ++dev->next_uid;
  b->prev = nullptr;
  b->next = dev->buffers;
  if (dev->buffers) dev->buffers->prev = b;
  dev->buffers = b;
  dev->bytes_used += size;
  *out = b;
  return 0;
}

}  // namespace

extern "C" {

// ------------------------------------------------------------------------------- device
int vx_dev_open(vx_device_h* hdevice) {
  if (nullptr == hdevice) return -1;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
    fprintf(stderr, "[VXDRV] no CUDA device: the B200 back-end has no CPU fallback\n");
    return 1;  // same code the stub returns when the driver cannot be loaded (stub/vortex.cpp:67-78)
  }
  int ordinal = env_int("VX_CUDA_DEVICE", env_int("LOCAL_RANK", 0));
  if (ordinal < 0 || ordinal >= count) ordinal = ordinal % count;
  vx_device* dev = new (std::nothrow) vx_device();
  if (!dev) return -1;
  dev->cuda_dev = ordinal;
  dev->buffers = nullptr;
  dev->next_uid = 0;
  dev->bytes_used = 0;
  dev->rt = nullptr;
  dev->raster = nullptr;
  dev->copy_stream = nullptr;
  dev->copy_next_ticket = 0;
  dev->stage_ring = nullptr;
  dev->stage_head = dev->stage_count = 0;
  dev->stage_tail = 0;
  dev->last_raster_tiles = 0;
  dev->last_raster_fragments = dev->last_raster_tests = 0;
  dev->run_recorded = false;
  dev->last_rays_primary = 0;
  dev->last_counters_valid = false;
  dev->last_pipeline = 0;
  dev->last_launches = 0;
  dev->last_node_bytes = 0;
  dev->last_rt_stores = 0;
  memset(dev->last_counters, 0, sizeof(dev->last_counters));
  memset(dev->last_raster_cls, 0, sizeof(dev->last_raster_cls));
  memset(dev->last_instrs_sm, 0, sizeof(dev->last_instrs_sm));
  memset(dev->acc, 0, sizeof(dev->acc));
  memset(dev->acc_instrs_sm, 0, sizeof(dev->acc_instrs_sm));
  dev->harvested = true;
  dev->run_mpm_class = 0;
  if (cudaSetDevice(ordinal) != cudaSuccess || cudaGetDeviceProperties(&dev->prop, ordinal) != cudaSuccess ||
      cudaStreamCreateWithFlags(&dev->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&dev->ev_start) != cudaSuccess || cudaEventCreate(&dev->ev_stop) != cudaSuccess) {
    fprintf(stderr, "[VXDRV] cannot initialise CUDA device %d: %s\n", ordinal, cudaGetErrorString(cudaGetLastError()));
    delete dev;
    return -1;
  }
  {  // stream-ordered temporaries (raster_binning.cu) stay in the pool between calls instead of going back to the OS
    cudaMemPool_t pool;
    uint64_t keep = ~0ull;
    if (cudaDeviceGetDefaultMemPool(&pool, ordinal) == cudaSuccess)
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    cudaGetLastError();
  }
  *hdevice = dev;
  return 0;
}

int vx_dev_close(vx_device_h hdevice) {
  if (nullptr == hdevice) return -1;
  vx_device* dev = as_dev(hdevice);
  cudaSetDevice(dev->cuda_dev);
  cudaStreamSynchronize(dev->stream);
  if (dev->copy_stream) {
    cudaStreamSynchronize(dev->copy_stream);
    cudaEventDestroy(dev->copy_fork);
    for (cudaEvent_t e : dev->copy_done) cudaEventDestroy(e);
    cudaStreamDestroy(dev->copy_stream);
    dev->copy_stream = nullptr;
  }
  if (dev->stage_ring) {
    for (cudaEvent_t e : dev->stage_done) cudaEventDestroy(e);
    cudaFreeHost(dev->stage_ring);
    dev->stage_ring = nullptr;
  }
  if (env_int("VORTEX_PROFILING", 0) != 0) vx_dump_perf(hdevice, stdout);  // stub/vortex.cpp:99-104
  srt::rt_destroy(dev);
  srt::raster_destroy(dev);
  while (dev->buffers) destroy_buffer(dev->buffers);  // apps leak their per-draw args buffers
  cudaStreamSynchronize(dev->stream);                 // the frees are stream-ordered
  {  // hand the pool's cached memory back: another library in this process (torch) may want it
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev->cuda_dev) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    cudaGetLastError();
  }
  cudaEventDestroy(dev->ev_start);
  cudaEventDestroy(dev->ev_stop);
  cudaStreamDestroy(dev->stream);
  delete dev;
  return 0;
}

int vx_dev_caps(vx_device_h hdevice, uint32_t caps_id, uint64_t* value) {
  if (nullptr == hdevice || nullptr == value) return -1;
  const vx_device* dev = as_dev(hdevice);
  switch (caps_id) {
    case VX_CAPS_VERSION: *value = 100; break;  // sm_100a
    case VX_CAPS_NUM_THREADS: *value = (uint64_t)dev->prop.warpSize; break;
    case VX_CAPS_NUM_WARPS: *value = (uint64_t)(dev->prop.maxThreadsPerMultiProcessor / dev->prop.warpSize); break;
    case VX_CAPS_NUM_CORES: *value = (uint64_t)dev->prop.multiProcessorCount; break;
    case VX_CAPS_CACHE_LINE_SIZE: *value = 128; break;
    case VX_CAPS_GLOBAL_MEM_SIZE: *value = (uint64_t)dev->prop.totalGlobalMem; break;
    case VX_CAPS_LOCAL_MEM_SIZE: *value = (uint64_t)dev->prop.sharedMemPerMultiprocessor; break;
    case VX_CAPS_ISA_FLAGS:
      *value = VX_ISA_STD_A | VX_ISA_STD_F | VX_ISA_STD_D | VX_ISA_STD_I | VX_ISA_EXT_ICACHE | VX_ISA_EXT_DCACHE |
               VX_ISA_EXT_L2CACHE | VX_ISA_EXT_LMEM | VX_ISA_EXT_RT | VX_ISA_EXT_TEX | VX_ISA_EXT_RASTER | VX_ISA_EXT_OM;
      break;
    default:
      fprintf(stderr, "[VXDRV] invalid caps id: %u\n", caps_id);
      return -1;
  }
  return 0;
}

// ------------------------------------------------------------------------------- memory
int vx_mem_alloc(vx_device_h hdevice, uint64_t size, int flags, vx_buffer_h* hbuffer) {
  if (nullptr == hdevice || nullptr == hbuffer || 0 == size) return -1;
  vx_buffer* b;
  if (alloc_buffer(as_dev(hdevice), size, flags, 0, &b) != 0) return -1;
  *hbuffer = b;
  return 0;
}

int vx_mem_reserve(vx_device_h hdevice, uint64_t address, uint64_t size, int flags, vx_buffer_h* hbuffer) {
  if (nullptr == hdevice || nullptr == hbuffer || 0 == size) return -1;
  vx_buffer* b;
  if (alloc_buffer(as_dev(hdevice), size, flags, address, &b) != 0) return -1;
  *hbuffer = b;
  return 0;
}

int vx_mem_free(vx_buffer_h hbuffer) {
  if (nullptr == hbuffer) return 0;  // callbacks.inc:100-103
  vx_buffer* b = as_buf(hbuffer);
  cudaSetDevice(b->dev->cuda_dev);
  return destroy_buffer(b);  // stream-ordered: work already queued on the launch stream still sees the memory
}

int vx_mem_access(vx_buffer_h hbuffer, uint64_t offset, uint64_t size, int flags) {
  if (nullptr == hbuffer) return -1;
  vx_buffer* b = as_buf(hbuffer);
  if (offset + size > b->size) return -1;
  b->flags = flags;
  return 0;
}

int vx_mem_address(vx_buffer_h hbuffer, uint64_t* address) {
  if (nullptr == hbuffer || nullptr == address) return -1;
  *address = as_buf(hbuffer)->reported_addr;
  return 0;
}

int vx_mem_info(vx_device_h hdevice, uint64_t* mem_free, uint64_t* mem_used) {
  if (nullptr == hdevice) return -1;
  vx_device* dev = as_dev(hdevice);
  size_t f = 0, t = 0;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  VX_CUDA_OK(cudaMemGetInfo(&f, &t));
  if (mem_free) *mem_free = f;
  if (mem_used) *mem_used = dev->bytes_used;
  return 0;
}

// Staged uploads travel by a copy KERNEL that reads the page-locked ring through its device mapping, not by a DMA
// transfer: SM loads over PCIe leave the copy engines to the read-back that runs beside a render loop on the second
// stream (vx_cuda_copy_from_dev_async).  Measured on B200 the two are equal for the raster loop (1.84 against 1.86 ms
// per frame, profiles/r02d_draw3d_e2e_timeline*.txt); VX_CUDA_STAGE_KERNEL=0 selects cudaMemcpyAsync (A/B).
namespace {
__global__ void __launch_bounds__(256) stage_copy_kernel(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint64_t n) {
  const uint64_t tid = (uint64_t)blockIdx.x * 256u + threadIdx.x, nth = (uint64_t)gridDim.x * 256u;
  uint64_t done = 0;
  if ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {  // src is a 256-byte granule of the ring
    const uint64_t n16 = n >> 4;
    for (uint64_t i = tid; i < n16; i += nth) reinterpret_cast<uint4*>(dst)[i] = __ldcv(reinterpret_cast<const uint4*>(src) + i);
    done = n16 << 4;
  }
  for (uint64_t i = done + tid; i < n; i += nth) dst[i] = __ldcv(src + i);
}
}  // namespace

static bool stage_kernel_enabled() {
  static const bool on = env_int("VX_CUDA_STAGE_KERNEL", 1) != 0;
  return on;
}

static bool staged_uploads_enabled() {  // VX_CUDA_STAGED_UPLOADS=0: every upload blocks on the stream (A/B)
  static const bool on = env_int("VX_CUDA_STAGED_UPLOADS", 1) != 0;
  return on;
}

int vx_copy_to_dev(vx_buffer_h hbuffer, const void* host_ptr, uint64_t dst_offset, uint64_t size) {
  if (nullptr == hbuffer || nullptr == host_ptr) return -1;
  vx_buffer* b = as_buf(hbuffer);
  if (dst_offset + size > b->size) return -1;  // callbacks.inc:155-156
  vx_device* dev = b->dev;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  if (b->shadow) {
    // small host-written buffers (kernel arguments, kernel images): the bytes go to the shadow first (vx_start decodes
    // from it) and travel from there.  The shadow is pageable memory, which cudaMemcpyAsync stages before it returns,
    // so the caller's buffer and the shadow are both free on return WITHOUT waiting for the frames still queued on the
    // stream -- a render loop that uploads a new kernel_arg_t per frame keeps the device busy.
    memcpy(b->shadow + dst_offset, host_ptr, size);
    VX_CUDA_OK(cudaMemcpyAsync((char*)b->dptr + dst_offset, b->shadow + dst_offset, size, cudaMemcpyHostToDevice, dev->stream));
  } else if (size <= kStageSlotBytes && staged_uploads_enabled()) {
    // Medium uploads (a draw's tile / primitive buffers, textures, a packed frame): the bytes are copied
    // into a page-locked ring and travel from there, queued on the launch stream behind the kernels already in it.
    // The caller's buffer is free on return, as the reference's blocking copy promises
    // (runtime/simx/vortex.cpp:138-153), but the host does not wait for the device: it goes on binning the next draws
    // -- and, the ring being 32 MiB handed out by the byte, the next frames -- while this one renders.
    if (!dev->stage_ring) {
      VX_CUDA_OK(cudaHostAlloc(&dev->stage_ring, kStageRingBytes, cudaHostAllocMapped));
      void* mapped = nullptr;  // unified addressing: the device sees the ring at its host address
      if (cudaHostGetDevicePointer(&mapped, dev->stage_ring, 0) != cudaSuccess || mapped != dev->stage_ring) {
        cudaGetLastError();
        dev->stage_mapped = false;
      } else {
        dev->stage_mapped = true;
      }
      for (cudaEvent_t& e : dev->stage_done) VX_CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      dev->stage_head = dev->stage_count = 0;
      dev->stage_tail = 0;
    }
    const uint64_t need = size ? (size + 255) & ~255ull : 256;
    auto retire_oldest = [&]() {
      dev->stage_head = (dev->stage_head + 1) % kStageEntries;
      --dev->stage_count;
    };
    // transfers that have finished give their bytes back without any waiting
    while (dev->stage_count && cudaEventQuery(dev->stage_done[dev->stage_head]) == cudaSuccess) retire_oldest();
    if (dev->stage_count == 0) dev->stage_tail = 0;
    uint64_t off = dev->stage_tail;
    if (off + need > kStageRingBytes) off = 0;  // no transfer wraps around the end of the ring
    auto in_use = [&]() {
      for (uint32_t i = 0; i < dev->stage_count; ++i) {
        const auto& e = dev->stage_span[(dev->stage_head + i) % kStageEntries];
        if (off < e.end && e.begin < off + need) return true;
      }
      return false;
    };
    // the bytes are reused only after every transfer that reads them has completed (oldest first: stream order)
    while (dev->stage_count && (dev->stage_count == (uint32_t)kStageEntries || in_use())) {
      VX_CUDA_OK(cudaEventSynchronize(dev->stage_done[dev->stage_head]));
      retire_oldest();
    }
    uint8_t* stage = dev->stage_ring + off;
    memcpy(stage, host_ptr, size);
    if (dev->stage_mapped && stage_kernel_enabled() && size > 0) {
      const uint64_t units = (size + 15) / 16;
      const unsigned cap = (unsigned)dev->prop.multiProcessorCount * 8u;
      const unsigned grid = (unsigned)((units + 255) / 256 < cap ? (units + 255) / 256 : cap);
      stage_copy_kernel<<<grid, 256, 0, dev->stream>>>((uint8_t*)b->dptr + dst_offset, stage, size);
      VX_CUDA_OK(cudaGetLastError());
    } else {
      VX_CUDA_OK(cudaMemcpyAsync((char*)b->dptr + dst_offset, stage, size, cudaMemcpyHostToDevice, dev->stream));
    }
    const uint32_t slot = (dev->stage_head + dev->stage_count) % kStageEntries;
    VX_CUDA_OK(cudaEventRecord(dev->stage_done[slot], dev->stream));
    dev->stage_span[slot].begin = off;
    dev->stage_span[slot].end = off + need;
    ++dev->stage_count;
    dev->stage_tail = off + need;
  } else {
    VX_CUDA_OK(cudaMemcpyAsync((char*)b->dptr + dst_offset, host_ptr, size, cudaMemcpyHostToDevice, dev->stream));
    VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
  }
  b->generation++;
  return 0;
}

int vx_copy_from_dev(void* host_ptr, vx_buffer_h hbuffer, uint64_t src_offset, uint64_t size) {
  if (nullptr == hbuffer || nullptr == host_ptr) return -1;
  vx_buffer* b = as_buf(hbuffer);
  if (src_offset + size > b->size) return -1;  // callbacks.inc:166-167
  vx_device* dev = b->dev;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  VX_CUDA_OK(cudaMemcpyAsync(host_ptr, (const char*)b->dptr + src_offset, size, cudaMemcpyDeviceToHost, dev->stream));
  VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
  return srt::raster_take_fault(dev);  // what is read back may come from a draw that met a malformed tile buffer
}

// ------------------------------------------------------------------------------- execution
namespace {

// Wait for the last run, bring its device counters into dev->last_*, and -- if it ran with the performance
// monitor on -- add them to the free-running accumulators.  Idempotent per run.
int harvest(vx_device* dev) {
  if (dev->harvested) return 0;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
  dev->harvested = true;
  float ms = 0.0f;
  if (dev->run_recorded) VX_CUDA_OK(cudaEventElapsedTime(&ms, dev->ev_start, dev->ev_stop));
  if (dev->last_pipeline == SRT_PIPELINE_RAYTRACE) {
    if (srt::rt_read_counters(dev) != 0) return -1;
  } else if (dev->last_pipeline == SRT_PIPELINE_DRAW3D) {
    if (srt::raster_read_counters(dev) != 0) return -1;
    dev->last_counters_valid = true;
  }
  if (dev->run_mpm_class == 0) return 0;
  uint64_t* a = dev->acc;
  a[VX_PERF_CYCLES] += (uint64_t)((double)ms * 1e-3 * (double)dev->prop.clockRate * 1e3);
  a[VX_PERF_NS] += (uint64_t)((double)ms * 1e6);
  a[VX_PERF_RUNS] += 1;
  if (dev->last_pipeline == SRT_PIPELINE_RAYTRACE) {
    const uint64_t* c = dev->last_counters;
    a[VX_PERF_RT_PRIMARY] += dev->last_rays_primary;
    a[VX_PERF_RT_SHADOW] += c[CTR_RAYS_SHADOW];
    a[VX_PERF_RT_BOUNCE] += c[CTR_RAYS_BOUNCE];
    a[VX_PERF_RT_NODES] += c[CTR_NODES];
    a[VX_PERF_RT_TRIS] += c[CTR_TRIS];
    a[VX_PERF_RT_NODE_BYTES] += c[CTR_NODES] * dev->last_node_bytes;
    a[VX_PERF_RT_STORES] += dev->last_rt_stores;
  } else {
    a[VX_PERF_FRAGMENTS] += dev->last_raster_fragments;
    a[VX_PERF_TESTS] += dev->last_raster_tests;
    a[VX_PERF_TILES] += dev->last_raster_tiles;
    a[VX_PERF_TEX_READS] += dev->last_raster_cls[0];
    a[VX_PERF_OM_READS] += dev->last_raster_cls[1];
    a[VX_PERF_OM_WRITES] += dev->last_raster_cls[2];
    a[VX_PERF_RASTER_READS] += dev->last_raster_cls[3];
  }
  for (int i = 0; i < kMaxSms; ++i) dev->acc_instrs_sm[i] += dev->last_instrs_sm[i];
  return 0;
}

}  // namespace

int vx_start(vx_device_h hdevice, vx_buffer_h hkernel, vx_buffer_h harguments) {
  if (nullptr == hdevice || nullptr == hkernel || nullptr == harguments) return -1;
  vx_device* dev = as_dev(hdevice);
  vx_buffer* k = as_buf(hkernel);
  vx_buffer* args = as_buf(harguments);
  // the kernel image is {magic, pipeline, version} (include/skybox_rt.h); it is host-written and
  // small, so its shadow copy is authoritative
  struct { uint64_t magic; uint32_t pipeline, version; } img;
  if (!k->shadow || k->size < sizeof(img)) {
    fprintf(stderr, "[VXDRV] vx_start: kernel buffer is not a pipeline descriptor\n");
    return -1;
  }
  memcpy(&img, k->shadow, sizeof(img));
  if (img.magic != SRT_KERNEL_MAGIC) {
    fprintf(stderr, "[VXDRV] vx_start: bad kernel magic (RISC-V .vxbin images cannot run on the CUDA back-end)\n");
    return -1;
  }
  // decode the arguments before any run state is touched: a refused launch leaves the previous run's record intact
  kernel_arg_t rt_arg;
  draw3d_kernel_arg_t d3_arg;
  switch (img.pipeline) {
    case SRT_PIPELINE_RAYTRACE:
      if (!args->shadow || args->size < sizeof(rt_arg)) {
        fprintf(stderr, "[VXDRV] vx_start: arguments buffer too small for kernel_arg_t\n");
        return -1;
      }
      memcpy(&rt_arg, args->shadow, sizeof(rt_arg));
      break;
    case SRT_PIPELINE_DRAW3D:
      if (!args->shadow || args->size < sizeof(d3_arg)) {
        fprintf(stderr, "[VXDRV] vx_start: arguments buffer too small for the draw3d kernel_arg_t\n");
        return -1;
      }
      memcpy(&d3_arg, args->shadow, sizeof(d3_arg));
      break;
    default:
      fprintf(stderr, "[VXDRV] vx_start: unknown pipeline %u\n", img.pipeline);
      return -1;
  }
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  // VORTEX_PROFILING=<class> turns the performance monitor on for every run (stub/vortex.cpp:142-149)
  static const int profiling_mode = env_int("VORTEX_PROFILING", 0);
  if (profiling_mode != 0) dev->dcrs[VX_DCR_BASE_MPM_CLASS] = (uint32_t)profiling_mode;
  // a profiled run that nobody read yet goes into the accumulators before its counters are reused
  if (dev->run_mpm_class != 0 && harvest(dev) != 0) return -1;
  dev->run_mpm_class = (int)vx_dev_dcr(dev, VX_DCR_BASE_MPM_CLASS);
  dev->harvested = false;
  dev->run_recorded = false;
  dev->last_counters_valid = false;
  dev->last_pipeline = (int)img.pipeline;
  VX_CUDA_OK(cudaEventRecord(dev->ev_start, dev->stream));
  const int rc = img.pipeline == SRT_PIPELINE_RAYTRACE ? srt::rt_launch(dev, rt_arg) : srt::raster_launch(dev, d3_arg);
  if (rc != 0) {
    // nothing was enqueued (the launch functions validate before they launch): there is no run to harvest
    dev->harvested = true;
    dev->run_mpm_class = 0;
    return rc;
  }
  VX_CUDA_OK(cudaEventRecord(dev->ev_stop, dev->stream));
  dev->run_recorded = true;
  return 0;
}

int vx_ready_wait(vx_device_h hdevice, uint64_t timeout) {
  if (nullptr == hdevice) return -1;
  vx_device* dev = as_dev(hdevice);
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  struct timespec t0, now;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (uint64_t spins = 0;; ++spins) {
    cudaError_t e = cudaStreamQuery(dev->stream);
    if (e == cudaSuccess) return srt::raster_take_fault(dev);
    if (e != cudaErrorNotReady) {
      fprintf(stderr, "[VXDRV] device fault: %s\n", cudaGetErrorString(e));
      return -1;
    }
    clock_gettime(CLOCK_MONOTONIC, &now);
    const uint64_t ms = (uint64_t)(now.tv_sec - t0.tv_sec) * 1000ull + (uint64_t)((now.tv_nsec - t0.tv_nsec) / 1000000);
    if (ms >= timeout) return -1;  // simx/vortex.cpp:195-209
    if (spins > 2000) {
      struct timespec nap = {0, 20000};
      nanosleep(&nap, nullptr);
    }
  }
}

int vx_dcr_write(vx_device_h hdevice, uint32_t addr, uint32_t value) {
  if (nullptr == hdevice) return -1;
  if (!dcr_addr_ok(addr)) {
    fprintf(stderr, "[VXDRV] invalid DCR write address 0x%x\n", addr);
    return -1;
  }
  vx_device* dev = as_dev(hdevice);
  // The reference blocks here until the previous run completes (simx/vortex.cpp:211-218) because its
  // units read DCRs while they run.  This back-end latches every DCR it uses into the launch parameters
  // at vx_start, so a write cannot reach a run in flight and there is nothing to wait for: draws of one
  // frame queue back to back on the stream.
  dev->dcrs[addr] = value;
  return 0;
}

int vx_dcr_read(vx_device_h hdevice, uint32_t addr, uint32_t* value) {
  if (nullptr == hdevice || nullptr == value) return -1;
  if (!dcr_addr_ok(addr)) return -1;
  vx_device* dev = as_dev(hdevice);
  auto it = dev->dcrs.find(addr);
  *value = it == dev->dcrs.end() ? 0u : it->second;
  return 0;
}

int vx_mpm_query(vx_device_h hdevice, uint32_t addr, uint32_t core_id, uint64_t* value) {
  if (nullptr == hdevice || nullptr == value) return -1;
  vx_device* dev = as_dev(hdevice);
  if (core_id >= (uint32_t)dev->prop.multiProcessorCount) return -1;
  if (harvest(dev) != 0) return -1;
  float ms = 0.0f;
  if (dev->run_recorded) VX_CUDA_OK(cudaEventElapsedTime(&ms, dev->ev_start, dev->ev_stop));
  const uint32_t cls = vx_dev_dcr(dev, VX_DCR_BASE_MPM_CLASS);
  const bool draw3d = dev->last_pipeline == SRT_PIPELINE_DRAW3D;
  const uint64_t* c = dev->last_counters;
  const uint64_t* a = dev->acc;
  const uint64_t last_instrs =
      draw3d ? dev->last_raster_fragments : dev->last_rays_primary + c[CTR_RAYS_SHADOW] + c[CTR_RAYS_BOUNCE];
  // the two CSRs every class has (include/skybox_rt.h: accumulated while the monitor is on, else the last run)
  if (addr == VX_CSR_MCYCLE) {
    *value = cls ? a[VX_PERF_CYCLES] : (uint64_t)((double)ms * 1e-3 * (double)dev->prop.clockRate * 1e3);
    return 0;
  }
  if (addr == VX_CSR_MINSTRET) {
    *value = cls ? dev->acc_instrs_sm[core_id] : (core_id == 0 ? last_instrs : 0);
    return 0;
  }
  if (addr >= VX_CSR_MPM_USER && addr < VX_CSR_MPM_USER_END) {  // class CSRs
    const uint64_t reads = a[VX_PERF_RT_NODES] + a[VX_PERF_RT_TRIS] + a[VX_PERF_RASTER_READS] + a[VX_PERF_TEX_READS] + a[VX_PERF_OM_READS];
    const uint64_t writes = a[VX_PERF_RT_STORES] + a[VX_PERF_OM_WRITES];
    uint64_t v = 0;
    switch (cls) {
      case VX_DCR_MPM_CLASS_CORE:
        if (addr == VX_CSR_MPM_LOADS) v = reads;
        else if (addr == VX_CSR_MPM_STORES) v = writes;
        break;
      case VX_DCR_MPM_CLASS_MEM:
        if (addr == VX_CSR_MPM_MEM_READS) v = reads;
        else if (addr == VX_CSR_MPM_MEM_WRITES) v = writes;
        break;
      case VX_DCR_MPM_CLASS_TEX:
        if (addr == VX_CSR_MPM_TEX_READS || addr == VX_CSR_MPM_TCACHE_READS) v = a[VX_PERF_TEX_READS];
        break;
      case VX_DCR_MPM_CLASS_RASTER:
        if (addr == VX_CSR_MPM_RASTER_READS || addr == VX_CSR_MPM_RCACHE_READS) v = a[VX_PERF_RASTER_READS];
        break;
      case VX_DCR_MPM_CLASS_OM:
        if (addr == VX_CSR_MPM_OM_READS || addr == VX_CSR_MPM_OCACHE_READS) v = a[VX_PERF_OM_READS];
        else if (addr == VX_CSR_MPM_OM_WRITES || addr == VX_CSR_MPM_OCACHE_WRITES) v = a[VX_PERF_OM_WRITES];
        break;
      default: break;
    }
    *value = v;  // latency / stall / miss counters have no source inside a CUDA kernel: they read 0
    return 0;
  }
  // back-end CSRs: totals of the last vx_start
  switch (addr) {
    case VX_CSR_RT_ELAPSED_NS: *value = (uint64_t)((double)ms * 1e6); return 0;
    case VX_CSR_RT_LAUNCHES: *value = dev->last_launches; return 0;
    default: break;
  }
  if (draw3d) {
    switch (addr) {
      case VX_CSR_RASTER_FRAGMENTS: *value = dev->last_raster_fragments; return 0;
      case VX_CSR_RASTER_QUAD_TESTS: *value = dev->last_raster_tests; return 0;
      case VX_CSR_RASTER_TILES: *value = dev->last_raster_tiles; return 0;
      default: return -1;
    }
  }
  switch (addr) {
    case VX_CSR_RT_RAYS_PRIMARY: *value = dev->last_rays_primary; break;
    case VX_CSR_RT_RAYS_SHADOW: *value = c[CTR_RAYS_SHADOW]; break;
    case VX_CSR_RT_RAYS_BOUNCE: *value = c[CTR_RAYS_BOUNCE]; break;
    case VX_CSR_RT_NODES: *value = c[CTR_NODES]; break;
    case VX_CSR_RT_TRIS: *value = c[CTR_TRIS]; break;
    case VX_CSR_RT_NODE_BYTES: *value = dev->last_node_bytes; break;
    case VX_CSR_RT_WARP_NODE_STEPS: *value = c[CTR_WARP_NODE_STEPS]; break;
    case VX_CSR_RT_WARP_TRI_STEPS: *value = c[CTR_WARP_TRI_STEPS]; break;
    default: return -1;
  }
  return 0;
}

// Accumulated totals of the performance monitor for vx_dump_perf (vx_utils.cpp); slots = VX_PERF_* of vx_internal.h.
int vx_cuda_perf_totals(vx_device_h hdevice, uint64_t* out, uint32_t count) {
  if (nullptr == hdevice || nullptr == out) return -1;
  vx_device* dev = as_dev(hdevice);
  if (harvest(dev) != 0) return -1;
  for (uint32_t i = 0; i < count; ++i) out[i] = i < VX_PERF_COUNT ? dev->acc[i] : 0;
  return 0;
}

// ------------------------------------------------------------------------------- CUDA extras
int vx_cuda_stream(vx_device_h hdevice, void** stream) {
  if (nullptr == hdevice || nullptr == stream) return -1;
  *stream = (void*)as_dev(hdevice)->stream;
  return 0;
}

namespace {
__global__ void __launch_bounds__(256) fill32_kernel(uint4* __restrict__ dst, uint64_t n16, uint32_t* __restrict__ tail,
                                                     uint32_t ntail, uint32_t v) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint4 vv = make_uint4(v, v, v, v);
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) dst[i] = vv;
  if (blockIdx.x == 0 && threadIdx.x < ntail) tail[threadIdx.x] = v;
}
}  // namespace

int vx_cuda_fill32(vx_buffer_h hbuffer, uint64_t offset, uint64_t size, uint32_t value) {
  if (nullptr == hbuffer) return -1;
  vx_buffer* b = as_buf(hbuffer);
  if ((offset | size) & 3u || offset + size > b->size) return -1;
  if (size == 0) return 0;
  vx_device* dev = b->dev;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  char* p = (char*)b->dptr + offset;
  // words up to the first 16-byte boundary and after the last one go through the tail path
  uint64_t head = (16 - ((uintptr_t)p & 15)) & 15;
  if (head > size) head = size;
  if (head) fill32_kernel<<<1, 32, 0, dev->stream>>>(nullptr, 0, (uint32_t*)p, (uint32_t)(head / 4), value);
  const uint64_t n16 = (size - head) / 16, rest = (size - head) % 16;
  if (n16 || rest) {
    const uint64_t want = (n16 + 255) / 256;
    const unsigned grid = (unsigned)(want < 1 ? 1 : (want > 148ull * 16 ? 148ull * 16 : want));
    fill32_kernel<<<grid, 256, 0, dev->stream>>>((uint4*)(p + head), n16, (uint32_t*)(p + head + n16 * 16),
                                                 (uint32_t)(rest / 4), value);
  }
  VX_CUDA_OK(cudaGetLastError());
  b->generation++;
  if (b->shadow) {
    for (uint64_t i = 0; i < size / 4; ++i) memcpy(b->shadow + offset + 4 * i, &value, 4);
  }
  return 0;
}

namespace {
__global__ void __launch_bounds__(256) probe_read_kernel(const uint4* __restrict__ src, uint64_t n16, uint32_t* sink) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint32_t acc = 0;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n16; i += 4 * stride) {  // four independent 16-byte loads in flight per thread
    const uint4 a = __ldg(src + i), b = __ldg(src + i + stride), c = __ldg(src + i + 2 * stride), d = __ldg(src + i + 3 * stride);
    acc ^= a.x ^ a.y ^ a.z ^ a.w ^ b.x ^ b.y ^ b.z ^ b.w ^ c.x ^ c.y ^ c.z ^ c.w ^ d.x ^ d.y ^ d.z ^ d.w;
  }
  for (; i < n16; i += stride) { const uint4 a = __ldg(src + i); acc ^= a.x ^ a.y ^ a.z ^ a.w; }
  if (acc == 0x9e3779b9u) *sink = acc;  // keeps the loads alive; the buffer is zero-filled, so never taken
}
}  // namespace

int vx_cuda_probe_read_bandwidth(vx_device_h hdevice, uint64_t bytes, uint32_t repeats, float* ms) {
  if (nullptr == hdevice || nullptr == ms || bytes < 16 || (bytes & 15) || repeats == 0) return -1;
  vx_device* dev = as_dev(hdevice);
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  uint4* buf = nullptr;
  uint32_t* sink = nullptr;
  cudaEvent_t e0, e1;
  VX_CUDA_OK(cudaMalloc(&buf, bytes + 16));
  sink = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(buf) + bytes);
  int rc = -1;
  if (cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess) {
    const unsigned grid = (unsigned)dev->prop.multiProcessorCount * 8u;
    cudaMemsetAsync(buf, 0, bytes + 16, dev->stream);
    probe_read_kernel<<<grid, 256, 0, dev->stream>>>(buf, bytes / 16, sink);
    cudaEventRecord(e0, dev->stream);
    for (uint32_t r = 0; r < repeats; ++r) probe_read_kernel<<<grid, 256, 0, dev->stream>>>(buf, bytes / 16, sink);
    cudaEventRecord(e1, dev->stream);
    if (cudaStreamSynchronize(dev->stream) == cudaSuccess && cudaEventElapsedTime(ms, e0, e1) == cudaSuccess) rc = 0;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
  }
  cudaFree(buf);
  cudaGetLastError();
  return rc;
}

int vx_cuda_host_register(vx_device_h hdevice, void* ptr, uint64_t size) {
  if (nullptr == hdevice || nullptr == ptr || 0 == size) return -1;
  cudaSetDevice(as_dev(hdevice)->cuda_dev);
  cudaError_t e = cudaHostRegister(ptr, size, cudaHostRegisterDefault);
  if (e != cudaSuccess) {
    cudaGetLastError();  // not sticky; the caller simply keeps a pageable buffer
    return -1;
  }
  return 0;
}

int vx_cuda_host_unregister(vx_device_h hdevice, void* ptr) {
  if (nullptr == hdevice || nullptr == ptr) return -1;
  cudaSetDevice(as_dev(hdevice)->cuda_dev);
  if (cudaHostUnregister(ptr) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return 0;
}

// Asynchronous read-back: ordered after everything queued on the launch stream so far, executed on a second stream so
// that the frames queued afterwards overlap the transfer.
int vx_cuda_copy_from_dev_async(void* host_ptr, vx_buffer_h hbuffer, uint64_t src_offset, uint64_t size, uint64_t* ticket) {
  if (nullptr == hbuffer || nullptr == host_ptr || nullptr == ticket) return -1;
  vx_buffer* b = as_buf(hbuffer);
  if (src_offset + size > b->size) return -1;
  vx_device* dev = b->dev;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  if (!dev->copy_stream) {
    VX_CUDA_OK(cudaStreamCreateWithFlags(&dev->copy_stream, cudaStreamNonBlocking));
    VX_CUDA_OK(cudaEventCreateWithFlags(&dev->copy_fork, cudaEventDisableTiming));
    for (cudaEvent_t& e : dev->copy_done) VX_CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  const uint64_t t = dev->copy_next_ticket;
  if (t >= 16) VX_CUDA_OK(cudaEventSynchronize(dev->copy_done[t % 16]));  // the ring slot's previous copy
  VX_CUDA_OK(cudaEventRecord(dev->copy_fork, dev->stream));
  VX_CUDA_OK(cudaStreamWaitEvent(dev->copy_stream, dev->copy_fork, 0));
  VX_CUDA_OK(cudaMemcpyAsync(host_ptr, (const char*)b->dptr + src_offset, size, cudaMemcpyDeviceToHost, dev->copy_stream));
  VX_CUDA_OK(cudaEventRecord(dev->copy_done[t % 16], dev->copy_stream));
  dev->copy_next_ticket = t + 1;
  *ticket = t;
  return 0;
}

int vx_cuda_copy_wait(vx_device_h hdevice, uint64_t ticket) {
  if (nullptr == hdevice) return -1;
  vx_device* dev = as_dev(hdevice);
  if (!dev->copy_stream || ticket >= dev->copy_next_ticket) return -1;
  if (ticket + 16 < dev->copy_next_ticket) return 0;  // its ring slot was reused, which waited for it
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  VX_CUDA_OK(cudaEventSynchronize(dev->copy_done[ticket % 16]));
  return 0;
}

int vx_cuda_device(vx_device_h hdevice, int* ordinal) {
  if (nullptr == hdevice || nullptr == ordinal) return -1;
  *ordinal = as_dev(hdevice)->cuda_dev;
  return 0;
}

}  // extern "C"
