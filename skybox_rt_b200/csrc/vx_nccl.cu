// vx_nccl.cu -- multi-GPU frame assembly behind the C ABI (include/vortex_cuda.h: vx_cuda_comm_*,
// vx_cuda_frame_gather, vx_cuda_gather_bytes).
//
// No counterpart in the reference (single device; SURVEY.md section 2.1 "NCCL ... none").  The path shards by
// pixels: image rows are dealt to the ranks in blocks of rows_per_block (kernel_arg_t.rank_id / num_ranks /
// rows_per_block), every rank renders into a compact strip buffer, and ONE exchange per frame brings the strips to
// the root: N-1 ncclSend / ncclRecv pairs in one NCCL group on the launch stream (over NVLink / NVSwitch), followed
// by one de-interleave kernel on the root that writes every row of the frame once, 16 bytes per thread.  All of it
// is enqueued behind the kernels vx_start queued; nothing blocks the host and no other library sits in between.
//
// NCCL is bound at first use with dlopen, not at link time: a process that already carries a libnccl.so.2
// (one that imported torch, say) keeps using that copy, and single-GPU users of libvortex.so need no NCCL at all.
#include <dlfcn.h>
#include <nccl.h>  // types and enums only; every entry point goes through the table below
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "vx_internal.h"

namespace {

struct NcclApi {
  void* lib;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  const char* (*GetErrorString)(ncclResult_t);
};

NcclApi g_nccl;
std::once_flag g_nccl_once;
bool g_nccl_ok = false;

void nccl_bind() {
  void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the copy this process already uses, if any
  if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) {
    fprintf(stderr, "[VXDRV] multi-GPU: cannot load libnccl.so.2: %s\n", dlerror());
    return;
  }
  g_nccl.lib = lib;
#define SRT_BIND(field, name)                                                         \
  *reinterpret_cast<void**>(&g_nccl.field) = dlsym(lib, name);                        \
  if (!g_nccl.field) { fprintf(stderr, "[VXDRV] multi-GPU: libnccl has no %s\n", name); return; }
  SRT_BIND(GetUniqueId, "ncclGetUniqueId")
  SRT_BIND(CommInitRank, "ncclCommInitRank")
  SRT_BIND(CommDestroy, "ncclCommDestroy")
  SRT_BIND(GroupStart, "ncclGroupStart")
  SRT_BIND(GroupEnd, "ncclGroupEnd")
  SRT_BIND(Send, "ncclSend")
  SRT_BIND(Recv, "ncclRecv")
  SRT_BIND(GetErrorString, "ncclGetErrorString")
#undef SRT_BIND
  g_nccl_ok = true;
}

bool nccl_ready() {
  std::call_once(g_nccl_once, nccl_bind);
  return g_nccl_ok;
}

#define VX_NCCL_OK(expr)                                                                                   \
  do {                                                                                                     \
    ncclResult_t r_ = (expr);                                                                              \
    if (r_ != ncclSuccess) {                                                                               \
      fprintf(stderr, "[VXDRV] NCCL error at %s:%d: %s\n", __FILE__, __LINE__, g_nccl.GetErrorString(r_)); \
      return -1;                                                                                           \
    }                                                                                                      \
  } while (0)

struct vx_comm {
  vx_device* dev;
  ncclComm_t comm;
  uint32_t rank, num_ranks;
  char* d_staging;  // root: the other ranks' strips, back to back in rank order
  uint64_t staging_bytes;
  char* d_pack;     // in-place gather, non-root: this rank's row blocks packed into one strip
  uint64_t pack_bytes;
};

// One thread per 16 bytes of the frame: row y comes from rank (y / rpb) % n, local row (y / rpb / n) * rpb + y % rpb.
// `strip` is the root's own strip buffer, staging holds the strips of the other ranks at staging_off[rank] bytes.
struct GatherLayout {
  uint64_t off[64];  // byte offset of rank r's strip inside the staging area (unused for the root)
};
// strip == nullptr: the root's own rows are already in place in `frame` (in-place gather) and are left alone.
template <class T>
__global__ void __launch_bounds__(256)
frame_deinterleave_kernel(T* __restrict__ frame, const T* __restrict__ strip, const char* __restrict__ staging,
                          const GatherLayout lay, uint32_t row_elems, uint32_t height, uint32_t rpb, uint32_t n,
                          uint32_t root) {
  const uint32_t y = blockIdx.y;
  const uint32_t blk = y / rpb, owner = blk % n, ly = (blk / n) * rpb + y % rpb;
  if (owner == root && strip == nullptr) return;
  const T* src = owner == root ? strip : reinterpret_cast<const T*>(staging + lay.off[owner]);
  src += (size_t)ly * row_elems;
  T* dst = frame + (size_t)y * row_elems;
  for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < row_elems; x += gridDim.x * blockDim.x) dst[x] = __ldcs(src + x);
}

// The inverse on a sending rank of the in-place gather: its row blocks of the full-size frame -> one compact strip.
template <class T>
__global__ void __launch_bounds__(256)
frame_pack_kernel(T* __restrict__ strip, const T* __restrict__ frame, uint32_t row_elems, uint32_t rpb, uint32_t n,
                  uint32_t rank) {
  const uint32_t ly = blockIdx.y;  // local row of the strip
  const uint32_t y = ((ly / rpb) * n + rank) * rpb + ly % rpb;
  const T* src = frame + (size_t)y * row_elems;
  T* dst = strip + (size_t)ly * row_elems;
  for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < row_elems; x += gridDim.x * blockDim.x) dst[x] = __ldcs(src + x);
}

}  // namespace

extern "C" {

int vx_cuda_comm_unique_id(void* id) {
  if (nullptr == id || !nccl_ready()) return -1;
  static_assert(sizeof(ncclUniqueId) == VX_CUDA_COMM_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId u;
  VX_NCCL_OK(g_nccl.GetUniqueId(&u));
  memcpy(id, &u, sizeof(u));
  return 0;
}

int vx_cuda_comm_init(vx_device_h hdevice, const void* id, uint32_t rank, uint32_t num_ranks, vx_comm_h* hcomm) {
  if (nullptr == hdevice || nullptr == id || nullptr == hcomm || num_ranks == 0 || num_ranks > 64 || rank >= num_ranks)
    return -1;
  if (!nccl_ready()) return -1;
  vx_device* dev = static_cast<vx_device*>(hdevice);
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  ncclUniqueId u;
  memcpy(&u, id, sizeof(u));
  vx_comm* c = new vx_comm();
  memset(c, 0, sizeof(*c));
  c->dev = dev;
  c->rank = rank;
  c->num_ranks = num_ranks;
  ncclResult_t r = g_nccl.CommInitRank(&c->comm, (int)num_ranks, u, (int)rank);
  if (r != ncclSuccess) {
    fprintf(stderr, "[VXDRV] ncclCommInitRank(rank %u of %u): %s\n", rank, num_ranks, g_nccl.GetErrorString(r));
    delete c;
    return -1;
  }
  *hcomm = c;
  return 0;
}

int vx_cuda_comm_destroy(vx_comm_h hcomm) {
  if (nullptr == hcomm) return 0;
  vx_comm* c = static_cast<vx_comm*>(hcomm);
  cudaSetDevice(c->dev->cuda_dev);
  cudaStreamSynchronize(c->dev->stream);
  if (c->d_staging) cudaFree(c->d_staging);
  if (c->d_pack) cudaFree(c->d_pack);
  g_nccl.CommDestroy(c->comm);
  delete c;
  return 0;
}

int vx_cuda_frame_gather(vx_device_h hdevice, vx_comm_h hcomm, vx_buffer_h hstrip, vx_buffer_h hframe, uint32_t width,
                         uint32_t height, uint32_t rows_per_block, uint32_t root) {
  if (nullptr == hdevice || nullptr == hcomm || nullptr == hstrip || width == 0 || height == 0 || rows_per_block == 0)
    return -1;
  vx_device* dev = static_cast<vx_device*>(hdevice);
  vx_comm* c = static_cast<vx_comm*>(hcomm);
  if (c->dev != dev || root >= c->num_ranks) return -1;
  const uint32_t n = c->num_ranks;
  vx_buffer* strip = static_cast<vx_buffer*>(hstrip);
  const uint64_t row_bytes = (uint64_t)width * 4u;
  const uint64_t my_bytes = (uint64_t)srt_rank_rows(height, rows_per_block, n, c->rank) * row_bytes;
  if (strip->size < my_bytes) return -1;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  if (c->rank != root) {
    if (my_bytes) VX_NCCL_OK(g_nccl.Send(strip->dptr, my_bytes, ncclUint8, (int)root, c->comm, dev->stream));
    return 0;
  }
  if (nullptr == hframe) return -1;
  vx_buffer* frame = static_cast<vx_buffer*>(hframe);
  if (frame->size < (uint64_t)height * row_bytes) return -1;
  GatherLayout lay;
  memset(&lay, 0, sizeof(lay));
  uint64_t total = 0;
  for (uint32_t r = 0; r < n; ++r) {
    if (r == root) continue;
    lay.off[r] = total;
    total += ((uint64_t)srt_rank_rows(height, rows_per_block, n, r) * row_bytes + 255u) & ~(uint64_t)255u;
  }
  if (total > c->staging_bytes) {
    if (c->d_staging) {
      VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
      cudaFree(c->d_staging);
      c->d_staging = nullptr;
      c->staging_bytes = 0;
    }
    VX_CUDA_OK(cudaMalloc(&c->d_staging, total));
    c->staging_bytes = total;
  }
  if (n > 1) {
    VX_NCCL_OK(g_nccl.GroupStart());
    for (uint32_t r = 0; r < n; ++r) {
      if (r == root) continue;
      const uint64_t bytes = (uint64_t)srt_rank_rows(height, rows_per_block, n, r) * row_bytes;
      if (bytes) VX_NCCL_OK(g_nccl.Recv(c->d_staging + lay.off[r], bytes, ncclUint8, (int)r, c->comm, dev->stream));
    }
    VX_NCCL_OK(g_nccl.GroupEnd());
  }
  // de-interleave: every row of the frame written once; 16-byte elements when the rows allow it
  const bool vec = (row_bytes % 16u) == 0 && ((uintptr_t)frame->dptr % 16u) == 0 && ((uintptr_t)strip->dptr % 16u) == 0;
  const uint32_t elems = (uint32_t)(vec ? row_bytes / 16u : row_bytes / 4u);
  const dim3 grid((elems + 255u) / 256u > 8u ? 8u : (elems + 255u) / 256u, height);
  if (vec)
    frame_deinterleave_kernel<uint4><<<grid, 256, 0, dev->stream>>>(
        static_cast<uint4*>(frame->dptr), static_cast<const uint4*>(strip->dptr), c->d_staging, lay, elems, height,
        rows_per_block, n, root);
  else
    frame_deinterleave_kernel<uint32_t><<<grid, 256, 0, dev->stream>>>(
        static_cast<uint32_t*>(frame->dptr), static_cast<const uint32_t*>(strip->dptr), c->d_staging, lay, elems, height,
        rows_per_block, n, root);
  VX_CUDA_OK(cudaGetLastError());
  return 0;
}

// Frames whose ranks render into FULL-SIZE buffers at the rows' final positions (the raster path: tile rows dealt
// round-robin, every rank's colour buffer has the frame's geometry).  A sending rank packs its row blocks into one
// strip (one kernel) and sends it with ONE ncclSend; the root receives the strips into its staging area and
// scatters them to their rows of its own buffer (one kernel); its own rows never move.  (A first version sent
// every 32-row block with its own ncclSend / ncclRecv straight into place: 128 messages of 512 KiB reached 72 GB/s
// between two GPUs, one 64 MiB message 400+.)
int vx_cuda_frame_gather_inplace(vx_device_h hdevice, vx_comm_h hcomm, vx_buffer_h hframe, uint32_t width, uint32_t height,
                                 uint32_t rows_per_block, uint32_t root) {
  if (nullptr == hdevice || nullptr == hcomm || nullptr == hframe || width == 0 || height == 0 || rows_per_block == 0) return -1;
  vx_device* dev = static_cast<vx_device*>(hdevice);
  vx_comm* c = static_cast<vx_comm*>(hcomm);
  if (c->dev != dev || root >= c->num_ranks) return -1;
  vx_buffer* frame = static_cast<vx_buffer*>(hframe);
  const uint64_t row_bytes = (uint64_t)width * 4u;
  if (frame->size < (uint64_t)height * row_bytes) return -1;
  const uint32_t n = c->num_ranks;
  if (n == 1) return 0;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  const bool vec = (row_bytes % 16u) == 0 && ((uintptr_t)frame->dptr % 16u) == 0;
  const uint32_t elems = (uint32_t)(vec ? row_bytes / 16u : row_bytes / 4u);
  const uint32_t gx = (elems + 255u) / 256u > 8u ? 8u : (elems + 255u) / 256u;
  if (c->rank != root) {
    const uint32_t rows = srt_rank_rows(height, rows_per_block, n, c->rank);
    if (rows == 0) return 0;
    const uint64_t bytes = (uint64_t)rows * row_bytes;
    if (bytes > c->pack_bytes) {
      if (c->d_pack) {
        VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
        cudaFree(c->d_pack);
        c->d_pack = nullptr;
        c->pack_bytes = 0;
      }
      VX_CUDA_OK(cudaMalloc(&c->d_pack, bytes));
      c->pack_bytes = bytes;
    }
    if (vec)
      frame_pack_kernel<uint4><<<dim3(gx, rows), 256, 0, dev->stream>>>(reinterpret_cast<uint4*>(c->d_pack),
                                                                        static_cast<const uint4*>(frame->dptr), elems,
                                                                        rows_per_block, n, c->rank);
    else
      frame_pack_kernel<uint32_t><<<dim3(gx, rows), 256, 0, dev->stream>>>(reinterpret_cast<uint32_t*>(c->d_pack),
                                                                           static_cast<const uint32_t*>(frame->dptr), elems,
                                                                           rows_per_block, n, c->rank);
    VX_CUDA_OK(cudaGetLastError());
    VX_NCCL_OK(g_nccl.Send(c->d_pack, bytes, ncclUint8, (int)root, c->comm, dev->stream));
    return 0;
  }
  GatherLayout lay;
  memset(&lay, 0, sizeof(lay));
  uint64_t total = 0;
  for (uint32_t r = 0; r < n; ++r) {
    if (r == root) continue;
    lay.off[r] = total;
    total += ((uint64_t)srt_rank_rows(height, rows_per_block, n, r) * row_bytes + 255u) & ~(uint64_t)255u;
  }
  if (total > c->staging_bytes) {
    if (c->d_staging) {
      VX_CUDA_OK(cudaStreamSynchronize(dev->stream));
      cudaFree(c->d_staging);
      c->d_staging = nullptr;
      c->staging_bytes = 0;
    }
    VX_CUDA_OK(cudaMalloc(&c->d_staging, total));
    c->staging_bytes = total;
  }
  VX_NCCL_OK(g_nccl.GroupStart());
  for (uint32_t r = 0; r < n; ++r) {
    if (r == root) continue;
    const uint64_t bytes = (uint64_t)srt_rank_rows(height, rows_per_block, n, r) * row_bytes;
    if (bytes) VX_NCCL_OK(g_nccl.Recv(c->d_staging + lay.off[r], bytes, ncclUint8, (int)r, c->comm, dev->stream));
  }
  VX_NCCL_OK(g_nccl.GroupEnd());
  if (vec)
    frame_deinterleave_kernel<uint4><<<dim3(gx, height), 256, 0, dev->stream>>>(
        static_cast<uint4*>(frame->dptr), nullptr, c->d_staging, lay, elems, height, rows_per_block, n, root);
  else
    frame_deinterleave_kernel<uint32_t><<<dim3(gx, height), 256, 0, dev->stream>>>(
        static_cast<uint32_t*>(frame->dptr), nullptr, c->d_staging, lay, elems, height, rows_per_block, n, root);
  VX_CUDA_OK(cudaGetLastError());
  return 0;
}

int vx_cuda_gather_bytes(vx_device_h hdevice, vx_comm_h hcomm, vx_buffer_h hsrc, uint64_t bytes, vx_buffer_h hdst,
                         const uint64_t* offsets, const uint64_t* sizes, uint32_t root) {
  if (nullptr == hdevice || nullptr == hcomm) return -1;
  vx_device* dev = static_cast<vx_device*>(hdevice);
  vx_comm* c = static_cast<vx_comm*>(hcomm);
  if (c->dev != dev || root >= c->num_ranks) return -1;
  vx_buffer* src = static_cast<vx_buffer*>(hsrc);
  if (bytes && (nullptr == src || src->size < bytes)) return -1;
  VX_CUDA_OK(cudaSetDevice(dev->cuda_dev));
  if (c->rank != root) {
    if (bytes) VX_NCCL_OK(g_nccl.Send(src->dptr, bytes, ncclUint8, (int)root, c->comm, dev->stream));
    return 0;
  }
  if (nullptr == hdst || nullptr == offsets || nullptr == sizes) return -1;
  vx_buffer* dst = static_cast<vx_buffer*>(hdst);
  for (uint32_t r = 0; r < c->num_ranks; ++r)
    if (offsets[r] + sizes[r] > dst->size) return -1;
  if (sizes[root] != bytes) return -1;
  VX_NCCL_OK(g_nccl.GroupStart());
  for (uint32_t r = 0; r < c->num_ranks; ++r) {
    if (r == root || sizes[r] == 0) continue;
    VX_NCCL_OK(g_nccl.Recv(static_cast<char*>(dst->dptr) + offsets[r], sizes[r], ncclUint8, (int)r, c->comm, dev->stream));
  }
  VX_NCCL_OK(g_nccl.GroupEnd());
  if (bytes)
    VX_CUDA_OK(cudaMemcpyAsync(static_cast<char*>(dst->dptr) + offsets[root], src->dptr, bytes, cudaMemcpyDeviceToDevice,
                               dev->stream));
  return 0;
}

}  // extern "C"
