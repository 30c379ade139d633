// vgpu_spill_noalloc: the spill kernel for callers that must not allocate.
//
// Who cannot allocate here, and why.  The spill path exists to FREE device memory, so it
// is needed at exactly the moment a vGPU has none left: its HBM slice is full and the
// oversubscription manager wants to push chunks out to a tier.  A cudaMalloc for scratch at
// that point can fail outright, and when it is made from inside the preload shim's own
// cuMemAlloc hook it re-enters the quota accounting and is charged to the very vGPU that
// is being relieved.  Those callers get this entry point: same copy+checksum kernel as
// vgpu_spill(VGPU_XFER_LDST), but its per-chunk accumulators live in a fixed array that is
// part of the loaded module instead of memory obtained at run time.
//
// Consequences of the fixed array (read before use):
//   * capacity: at most VGPU_NOALLOC_MAX_CHUNKS chunks per call (4096 x 2 MiB = 8 GiB);
//     a larger batch is refused with VGPU_E_RANGE, split it into several calls.
//   * cost: 4096 x 32 B = 128 KiB of device memory on every GPU that loads the library,
//     held until the process exits.
//   * one array per GPU, shared by every call: two calls whose kernels overlap on the same
//     GPU (different streams, different threads) race on it and produce garbage checksums.
//     Serialise them (one stream, or an event between them).
//   * the array keeps whatever the previous call left.  The part a call uses is cleared on
//     the caller's stream before the kernel starts, so an earlier call that was aborted
//     half-way cannot poison this one.
//   * the first call on each GPU looks up the array's device address (the GPU must be the
//     current device).  Make that first call during start-up, outside any stream capture;
//     later calls do no set-up work.
// It calls no cudaMalloc / cudaFree / cudaHostAlloc and nothing that allocates on its
// behalf; inputs, outputs and the checksum array are the caller's.
#include <cuda_runtime.h>

#include <algorithm>
#include <mutex>

#include "vgpu_internal.h"
#include "vgpu_kernels.cuh"

namespace {

using vgpu::ChunkAcc;

// The scratch itself.  In host code this name is only a handle for the runtime — it is not
// a device address and must never be passed to a kernel; see noalloc_scratch() below.
static __device__ __align__(256) ChunkAcc g_noalloc_acc[VGPU_NOALLOC_MAX_CHUNKS];

constexpr int kMaxDev = 16;
std::mutex g_mu;
ChunkAcc *g_acc_addr[kMaxDev] = {nullptr};
int g_sms[kMaxDev] = {0};

// Device address of g_noalloc_acc on the CURRENT device, resolved once per device.
int noalloc_scratch(ChunkAcc **out, int *sms) {
    int dev = -1;
    VGPU_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDev) return vgpu::fail(VGPU_E_INVAL, "device %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_acc_addr[dev]) {
        ChunkAcc *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_noalloc_acc);
        if (e != cudaSuccess || !p)
            return vgpu::fail(VGPU_E_CUDA, "cudaGetSymbolAddress(g_noalloc_acc) on device %d: %s", dev,
                              cudaGetErrorString(e));
        VGPU_CUDA(cudaDeviceGetAttribute(&g_sms[dev], cudaDevAttrMultiProcessorCount, dev));
        g_acc_addr[dev] = p;
    }
    *out = g_acc_addr[dev];
    *sms = g_sms[dev];
    return VGPU_OK;
}

}  // namespace

extern "C" int vgpu_spill_noalloc(const void *src, void *dst, const uint32_t *src_slot,
                                  const uint32_t *dst_slot, size_t nchunks, size_t chunk_bytes,
                                  vgpu_cksum_t *cksum, void *stream) {
    using namespace vgpu;
    if (nchunks == 0) return VGPU_OK;
    if (nchunks > VGPU_NOALLOC_MAX_CHUNKS)
        return fail(VGPU_E_RANGE,
                    "vgpu_spill_noalloc: %zu chunks exceed the fixed scratch of %d chunks; split the batch",
                    nchunks, VGPU_NOALLOC_MAX_CHUNKS);
    if (!src || !dst || !cksum) return fail(VGPU_E_INVAL, "null pointer");
    if ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 31u)
        return fail(VGPU_E_INVAL, "slab base must be 32-byte aligned");
    if (chunk_bytes == 0 || chunk_bytes % 32 != 0 || chunk_bytes > (1ull << 30))
        return fail(VGPU_E_INVAL, "chunk_bytes=%zu must be a multiple of 32 in (0, 1 GiB]", chunk_bytes);

    ChunkAcc *acc = nullptr;
    int sms = 0;
    if (int rc = noalloc_scratch(&acc, &sms)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the accumulators of this batch must start at zero, whatever the last call left behind
    VGPU_CUDA(cudaMemsetAsync(acc, 0, nchunks * sizeof(ChunkAcc), st));

    constexpr int U = 4;
    constexpr uint32_t tile = LdstTile<U>::bytes;
    XferArgs a{};
    a.src = static_cast<const unsigned char *>(src);
    a.dst = static_cast<unsigned char *>(dst);
    a.src_slot = src_slot;
    a.dst_slot = dst_slot;
    a.chunk_bytes = chunk_bytes;
    a.nchunks = (uint32_t)nchunks;
    a.tiles_per_chunk = (uint32_t)((chunk_bytes + tile - 1) / tile);
    a.acc = acc;  // the resolved device address, not the symbol
    a.cksum = cksum;
    const uint64_t total = (uint64_t)nchunks * a.tiles_per_chunk;
    // Where the destination lives decides the grid, as in vgpu_spill: pinned host memory is PCIe-bound and 32 CTAs
    // saturate it — more would only park on the SMs of the tenants this eviction is meant to relieve.  (A query, not
    // an allocation.)
    uint64_t cap = (uint64_t)sms * 32;
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, dst) == cudaSuccess) {
        if (at.type == cudaMemoryTypeHost) cap = 32;
    } else {
        cudaGetLastError();
    }
    const unsigned grid = (unsigned)std::min<uint64_t>(total, cap);
    k_xfer_ldst<U, true><<<grid, kXfThreads, 0, st>>>(a);
    VGPU_CUDA(cudaGetLastError());
    return VGPU_OK;
}
