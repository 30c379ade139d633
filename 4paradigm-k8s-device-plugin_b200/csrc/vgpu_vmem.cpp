// Virtual device memory manager (include/vgpu.h, vgpu_vmem_*): one reserved VA range per vGPU
// allocation arena, 2 MiB chunks backed on demand through the CUDA VMM driver API, spilled to
// and restored from a tier slab by the fused copy+checksum kernels (vgpu_spill / vgpu_restore).
//
// Replaces the reference's reliance on UVM (cuMemoryAllocate@0x32146 -> cuMemAllocManaged) and
// completes its dead VMM branch (allocmode 1, @0x3225f-0x322a0).  Unlike UVM there are no page
// faults: residency changes only at explicit ensure/evict calls (allocation time, suspend /
// resume of a vGPU, or a monitor's decision), which is what makes every byte of traffic a
// bandwidth-bound batched kernel instead of fault-driven 4 KiB..2 MiB migrations.
//
// Physical handles are pooled: evicting a chunk unmaps it and parks its handle, fetching takes a
// parked handle first, so steady-state paging costs cuMemMap/cuMemUnmap/cuMemSetAccess only.
#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <mutex>
#include <vector>

#include "vgpu_internal.h"

using namespace vgpu;

namespace {

struct VmmApi {
    bool ok = false;
    CUresult (*AddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*AddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*Create)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*GetGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*GetErrorString)(CUresult, const char **) = nullptr;
};
VmmApi g_api;
std::once_flag g_once;

void load_api() {
    void *lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return;
    auto sym = [&](auto &fn, const char *name) {
        fn = reinterpret_cast<std::remove_reference_t<decltype(fn)>>(dlsym(lib, name));
        return fn != nullptr;
    };
    g_api.ok = sym(g_api.AddressReserve, "cuMemAddressReserve") && sym(g_api.AddressFree, "cuMemAddressFree") &&
               sym(g_api.Create, "cuMemCreate") && sym(g_api.Release, "cuMemRelease") && sym(g_api.Map, "cuMemMap") &&
               sym(g_api.Unmap, "cuMemUnmap") && sym(g_api.SetAccess, "cuMemSetAccess") &&
               sym(g_api.GetGranularity, "cuMemGetAllocationGranularity") && sym(g_api.GetErrorString, "cuGetErrorString");
}

int drv_fail(const char *what, CUresult r) {
    const char *s = nullptr;
    if (g_api.GetErrorString) g_api.GetErrorString(r, &s);
    return fail(r == CUDA_ERROR_OUT_OF_MEMORY ? VGPU_E_NOMEM : VGPU_E_CUDA, "%s -> %d (%s)", what, (int)r, s ? s : "?");
}

enum ChunkState : uint8_t { kUntouched = 0, kResident = 1, kSpilled = 2 };

struct Chunk {
    CUmemGenericAllocationHandle handle = 0;
    uint64_t last_use = 0;
    uint32_t tier_slot = 0;
    ChunkState state = kUntouched;
    vgpu_cksum_t cksum{0, 0};
};

constexpr size_t kCB = VGPU_CHUNK_BYTES;
constexpr size_t kBatch = 4096;  // chunks per kernel launch (8 GiB)

struct Vmem {
    int device = 0, tier = 0, peer = -1;
    CUdeviceptr base = 0;
    size_t nchunks = 0, resident_limit = 0, tier_slots = 0;
    void *tier_slab = nullptr;
    std::vector<Chunk> chunks;
    std::vector<CUmemGenericAllocationHandle> pool;  // parked physical chunks
    std::vector<uint32_t> free_slots;
    size_t resident = 0, spilled = 0;
    uint64_t clock = 0;
    // pinned + mapped staging (host writes, kernels read / write over PCIe; a few KiB per batch)
    uint32_t *src_slot = nullptr, *dst_slot = nullptr;
    vgpu_cksum_t *ck = nullptr;
    vgpu_restore_result_t *res = nullptr;
    CUmemAllocationProp prop{};
    CUmemAccessDesc access{};
    vgpu_vmem_stats_t st{};
    std::mutex mu;
};

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int map_chunk(Vmem *v, size_t c) {
    Chunk &ch = v->chunks[c];
    CUmemGenericAllocationHandle h;
    if (!v->pool.empty()) {
        h = v->pool.back();
        v->pool.pop_back();
    } else {
        CUresult r = g_api.Create(&h, kCB, &v->prop, 0);
        if (r != CUDA_SUCCESS) return drv_fail("cuMemCreate", r);
    }
    CUresult r = g_api.Map(v->base + c * kCB, kCB, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        v->pool.push_back(h);
        return drv_fail("cuMemMap", r);
    }
    ch.handle = h;
    return VGPU_OK;
}

// cuMemSetAccess over maximal runs of consecutive chunk indices (sorted input)
int set_access_runs(Vmem *v, const std::vector<size_t> &idx) {
    size_t i = 0;
    while (i < idx.size()) {
        size_t j = i + 1;
        while (j < idx.size() && idx[j] == idx[j - 1] + 1) ++j;
        CUresult r = g_api.SetAccess(v->base + idx[i] * kCB, (j - i) * kCB, &v->access, 1);
        if (r != CUDA_SUCCESS) return drv_fail("cuMemSetAccess", r);
        i = j;
    }
    return VGPU_OK;
}

// Spill `victims` (resident chunk indices) to the tier, then unmap them and park or release their memory.
int evict_list(Vmem *v, const std::vector<size_t> &victims, bool release) {
    if (victims.empty()) return VGPU_OK;
    if (!v->tier_slab && v->tier_slots)
        if (int rc = vgpu_slab_alloc(v->tier, v->device, v->peer, v->tier_slots * kCB, &v->tier_slab)) return rc;
    if (victims.size() > v->free_slots.size()) return fail(VGPU_E_NOMEM, "spill tier full: %zu victims, %zu free slots", victims.size(), v->free_slots.size());
    const double t0 = now_ms();
    for (size_t off = 0; off < victims.size(); off += kBatch) {
        const size_t n = std::min(kBatch, victims.size() - off);
        for (size_t k = 0; k < n; ++k) {
            v->src_slot[k] = (uint32_t)victims[off + k];
            v->dst_slot[k] = v->free_slots[v->free_slots.size() - 1 - k];
        }
        int rc = vgpu_spill(VGPU_XFER_AUTO, v->tier, (const void *)v->base, v->tier_slab, v->src_slot, v->dst_slot, n, kCB, v->ck, nullptr);
        if (rc) return rc;
        VGPU_CUDA(cudaStreamSynchronize(nullptr));
        for (size_t k = 0; k < n; ++k) {
            Chunk &ch = v->chunks[victims[off + k]];
            ch.cksum = v->ck[k];
            ch.tier_slot = v->dst_slot[k];
            CUresult r = g_api.Unmap(v->base + victims[off + k] * kCB, kCB);
            if (r != CUDA_SUCCESS) return drv_fail("cuMemUnmap", r);
            if (release) g_api.Release(ch.handle);  // explicit evict: the HBM goes back to the device
            else v->pool.push_back(ch.handle);      // making room inside ensure(): reused within microseconds
            ch.handle = 0;
            ch.state = kSpilled;
        }
        v->free_slots.resize(v->free_slots.size() - n);
        v->resident -= n;
        v->spilled += n;
        v->st.evictions += n;
    }
    v->st.evict_ms += now_ms() - t0;
    return VGPU_OK;
}

}  // namespace

extern "C" {

int vgpu_vmem_create(int device, uint64_t virtual_bytes, uint64_t resident_bytes, int tier, int peer_device,
                     uint64_t tier_bytes, void **handle) {
    if (!handle || !virtual_bytes) return fail(VGPU_E_INVAL, "bad vmem request");
    if (tier != VGPU_TIER_PEER && tier != VGPU_TIER_HOST) return fail(VGPU_E_INVAL, "tier must be VGPU_TIER_PEER or VGPU_TIER_HOST");
    std::call_once(g_once, load_api);
    if (!g_api.ok) return fail(VGPU_E_UNSUPPORTED, "CUDA VMM driver API not available");
    VGPU_CUDA(cudaSetDevice(device));
    VGPU_CUDA(cudaFree(0));
    Vmem *v = new Vmem();
    v->device = device;
    v->tier = tier;
    v->peer = peer_device;
    v->nchunks = (virtual_bytes + kCB - 1) / kCB;
    v->resident_limit = std::max<uint64_t>(1, resident_bytes / kCB);
    v->tier_slots = tier_bytes / kCB;
    v->prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    v->prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    v->prop.location.id = device;
    v->access.location = v->prop.location;
    v->access.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    size_t gran = 0;
    CUresult r = g_api.GetGranularity(&gran, &v->prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM);
    if (r != CUDA_SUCCESS || gran == 0 || kCB % gran != 0) {
        delete v;
        return r != CUDA_SUCCESS ? drv_fail("cuMemGetAllocationGranularity", r)
                                 : fail(VGPU_E_UNSUPPORTED, "allocation granularity %zu does not divide the 2 MiB chunk", gran);
    }
    r = g_api.AddressReserve(&v->base, v->nchunks * kCB, kCB, 0, 0);
    if (r != CUDA_SUCCESS) {
        delete v;
        return drv_fail("cuMemAddressReserve", r);
    }
    int rc = VGPU_OK;  // the tier slab is allocated at the first eviction, not here
    if (!rc) {
        cudaError_t e = cudaHostAlloc(&v->src_slot, kBatch * sizeof(uint32_t), cudaHostAllocMapped | cudaHostAllocPortable);
        if (e == cudaSuccess) e = cudaHostAlloc(&v->dst_slot, kBatch * sizeof(uint32_t), cudaHostAllocMapped | cudaHostAllocPortable);
        if (e == cudaSuccess) e = cudaHostAlloc(&v->ck, kBatch * sizeof(vgpu_cksum_t), cudaHostAllocMapped | cudaHostAllocPortable);
        if (e == cudaSuccess) e = cudaHostAlloc(&v->res, sizeof(vgpu_restore_result_t), cudaHostAllocMapped | cudaHostAllocPortable);
        if (e != cudaSuccess) rc = fail(VGPU_E_NOMEM, "pinned staging: %s", cudaGetErrorString(e));
    }
    if (rc) {
        vgpu_vmem_destroy(v);
        return rc;
    }
    v->chunks.resize(v->nchunks);
    v->free_slots.resize(v->tier_slots);
    for (size_t i = 0; i < v->tier_slots; ++i) v->free_slots[i] = (uint32_t)(v->tier_slots - 1 - i);
    *handle = v;
    return VGPU_OK;
}

int vgpu_vmem_base(void *handle, void **base) {
    if (!handle || !base) return fail(VGPU_E_INVAL, "null argument");
    *base = (void *)static_cast<Vmem *>(handle)->base;
    return VGPU_OK;
}

int vgpu_vmem_evict(void *handle, uint64_t first, uint64_t n) {
    Vmem *v = static_cast<Vmem *>(handle);
    if (!v || first + n > v->nchunks) return fail(VGPU_E_INVAL, "chunk range out of bounds");
    std::lock_guard<std::mutex> lk(v->mu);
    VGPU_CUDA(cudaSetDevice(v->device));
    std::vector<size_t> victims;
    for (size_t c = first; c < first + n; ++c)
        if (v->chunks[c].state == kResident) victims.push_back(c);
    int rc = evict_list(v, victims, true);
    for (auto h : v->pool) g_api.Release(h);
    v->pool.clear();
    return rc;
}

int vgpu_vmem_ensure(void *handle, uint64_t first, uint64_t n) {
    Vmem *v = static_cast<Vmem *>(handle);
    if (!v || first + n > v->nchunks) return fail(VGPU_E_INVAL, "chunk range out of bounds");
    if (n > v->resident_limit) return fail(VGPU_E_NOMEM, "%llu chunks requested, resident limit is %zu", (unsigned long long)n, v->resident_limit);
    std::lock_guard<std::mutex> lk(v->mu);
    VGPU_CUDA(cudaSetDevice(v->device));
    const uint64_t stamp = ++v->clock;
    std::vector<size_t> fresh, back;
    for (size_t c = first; c < first + n; ++c) {
        Chunk &ch = v->chunks[c];
        ch.last_use = stamp;
        if (ch.state == kUntouched) fresh.push_back(c);
        else if (ch.state == kSpilled) back.push_back(c);
    }
    const size_t need = fresh.size() + back.size();
    if (v->resident + need > v->resident_limit) {
        // least-recently-used resident chunks outside the requested range make room
        size_t want = v->resident + need - v->resident_limit;
        std::vector<size_t> cand;
        for (size_t c = 0; c < v->nchunks; ++c)
            if (v->chunks[c].state == kResident && v->chunks[c].last_use != stamp) cand.push_back(c);
        if (cand.size() < want) return fail(VGPU_E_NOMEM, "cannot make room: %zu needed, %zu evictable", want, cand.size());
        std::nth_element(cand.begin(), cand.begin() + (want - 1), cand.end(),
                         [&](size_t a, size_t b) { return v->chunks[a].last_use < v->chunks[b].last_use || (v->chunks[a].last_use == v->chunks[b].last_use && a < b); });
        cand.resize(want);
        std::sort(cand.begin(), cand.end());
        if (int rc = evict_list(v, cand, false)) return rc;
    }
    const double t0 = now_ms();
    std::vector<size_t> all(fresh);
    all.insert(all.end(), back.begin(), back.end());
    std::sort(all.begin(), all.end());
    for (size_t c : all)
        if (int rc = map_chunk(v, c)) return rc;
    if (int rc = set_access_runs(v, all)) return rc;
    for (size_t c : fresh) v->chunks[c].state = kResident;
    v->resident += fresh.size();
    int result = VGPU_OK;
    for (size_t off = 0; off < back.size(); off += kBatch) {
        const size_t m = std::min(kBatch, back.size() - off);
        for (size_t k = 0; k < m; ++k) {
            v->src_slot[k] = v->chunks[back[off + k]].tier_slot;
            v->dst_slot[k] = (uint32_t)back[off + k];
            v->ck[k] = v->chunks[back[off + k]].cksum;
        }
        v->res->bad_chunks = 0;
        v->res->first_bad_inv = 0;
        int rc = vgpu_restore(VGPU_XFER_AUTO, v->tier, v->tier_slab, (void *)v->base, v->src_slot, v->dst_slot, m, kCB, v->ck, nullptr, v->res, nullptr);
        if (rc) return rc;
        VGPU_CUDA(cudaStreamSynchronize(nullptr));
        if (v->res->bad_chunks) {
            v->st.checksum_failures += v->res->bad_chunks;
            result = fail(VGPU_E_MISMATCH, "%llu restored chunk(s) failed their checksum (first: chunk %u)",
                          (unsigned long long)v->res->bad_chunks, v->dst_slot[(size_t)~v->res->first_bad_inv]);
        }
        for (size_t k = 0; k < m; ++k) {
            Chunk &ch = v->chunks[back[off + k]];
            v->free_slots.push_back(ch.tier_slot);
            ch.state = kResident;
        }
        v->resident += m;
        v->spilled -= m;
        v->st.fetches += m;
    }
    v->st.fetch_ms += now_ms() - t0;
    return result;
}

int vgpu_vmem_stats(void *handle, vgpu_vmem_stats_t *out) {
    Vmem *v = static_cast<Vmem *>(handle);
    if (!v || !out) return fail(VGPU_E_INVAL, "null argument");
    std::lock_guard<std::mutex> lk(v->mu);
    *out = v->st;
    out->chunks_total = v->nchunks;
    out->chunks_resident = v->resident;
    out->chunks_spilled = v->spilled;
    out->resident_limit = v->resident_limit;
    out->tier_slots = v->tier_slots;
    return VGPU_OK;
}

int vgpu_vmem_destroy(void *handle) {
    Vmem *v = static_cast<Vmem *>(handle);
    if (!v) return VGPU_OK;
    cudaSetDevice(v->device);
    cudaDeviceSynchronize();
    for (size_t c = 0; c < v->chunks.size(); ++c)
        if (v->chunks[c].state == kResident && v->chunks[c].handle) {
            g_api.Unmap(v->base + c * kCB, kCB);
            g_api.Release(v->chunks[c].handle);
        }
    for (auto h : v->pool) g_api.Release(h);
    if (v->base) g_api.AddressFree(v->base, v->nchunks * kCB);
    if (v->tier_slab) vgpu_slab_free(v->tier, v->device, v->peer, v->tier_slab);
    if (v->src_slot) cudaFreeHost(v->src_slot);
    if (v->dst_slot) cudaFreeHost(v->dst_slot);
    if (v->ck) cudaFreeHost(v->ck);
    if (v->res) cudaFreeHost(v->res);
    delete v;
    return VGPU_OK;
}

}  // extern "C"
