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
#include <cstdlib>
#include <cstring>
#include <map>
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
    CUresult (*CtxGetCurrent)(CUcontext *) = nullptr;
    CUresult (*CtxSetCurrent)(CUcontext) = nullptr;
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
               sym(g_api.GetGranularity, "cuMemGetAllocationGranularity") && sym(g_api.GetErrorString, "cuGetErrorString") &&
               sym(g_api.CtxGetCurrent, "cuCtxGetCurrent") && sym(g_api.CtxSetCurrent, "cuCtxSetCurrent");
}

// The manager works in the device's primary context (cudaSetDevice), but it is called on the APPLICATION's thread —
// from inside the shim's cuMemAlloc / cuMemFree / cuMemPrefetchAsync hooks — where the current context may be one
// made with cuCtxCreate, the stand-in context of a second vGPU of the GPU, or a green context.  Whatever was current
// when an entry point is entered is current again when it returns.
struct CtxScope {
    CUcontext saved = nullptr;
    bool have = false;
    CtxScope() { have = g_api.ok && g_api.CtxGetCurrent(&saved) == CUDA_SUCCESS; }
    ~CtxScope() {
        if (have) g_api.CtxSetCurrent(saved);
    }
};

int drv_fail(const char *what, CUresult r) {
    const char *s = nullptr;
    if (g_api.GetErrorString) g_api.GetErrorString(r, &s);
    return fail(r == CUDA_ERROR_OUT_OF_MEMORY ? VGPU_E_NOMEM : VGPU_E_CUDA, "%s -> %d (%s)", what, (int)r, s ? s : "?");
}

// Residency unit = BLOCK of consecutive chunks (up to 512 x 2 MiB = 1 GiB, see vgpu_vmem_create), transfer and
// checksum unit = chunk.  Every VMM driver call costs 80-350 us whatever its size (measured on
// B200: cuMemCreate 100, cuMemSetAccess 215, cuMemUnmap 90 us per call; profiles/r01_vmm_costs.md),
// so mapping per 2 MiB chunk caps paging at ~5 GB/s; per 256 MiB block the calls cost ~0.4 ms
// against 0.37 ms of NVLink transfer and are overlapped with the kernels of the neighbouring
// blocks (launches are asynchronous; the host maps block k+1 while the GPU restores block k).
enum BlockState : uint8_t { kUntouched = 0, kResident = 1, kSpilled = 2 };

struct Block {
    CUmemGenericAllocationHandle handle = 0;
    uint64_t last_use = 0;
    uint32_t tier_slot = 0;  // block-sized slot in the tier slab
    uint8_t tier_idx = 0;    // which of the manager's tiers holds it
    BlockState state = kUntouched;
};

constexpr size_t kCB = VGPU_CHUNK_BYTES;

struct Vmem {
    int device = 0;
    CUdeviceptr base = 0;
    size_t nchunks = 0;         // as requested
    size_t bc = 0;              // chunks per block
    size_t nblocks = 0, resident_limit = 0 /* blocks */;
    // Spill tiers in order of preference (north_star: "first idle peer-GPU HBM over NVLink via P2P,
    // then pinned host RAM"): a victim goes to the first tier that still has a free slot.
    struct Tier {
        int kind = 0, peer = -1;
        void *slab = nullptr;       // peer tier: one allocation on the peer GPU, made at the first eviction that needs it
        std::vector<void *> piece;  // host tier: one pinned piece per slot IN USE, from the process-wide pool (host_piece_get)
        size_t slots = 0;           // blocks
        std::vector<uint32_t> free_slots;
        uint64_t evictions = 0;     // chunks
    };
    std::vector<Tier> tiers;
    std::vector<Block> blocks;
    std::vector<CUmemGenericAllocationHandle> pool;  // parked physical blocks
    size_t resident = 0, spilled = 0;  // blocks
    uint64_t clock = 0;
    vgpu_cksum_t *ck = nullptr;          // pinned + mapped: (X,S) of every chunk, written by the spill kernel,
                                         // read as `expect` by the restore kernel
    vgpu_restore_result_t *res = nullptr;  // pinned + mapped, one per block: which block failed is known exactly
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[2] = {nullptr, nullptr};
    CUmemAllocationProp prop{};
    CUmemAccessDesc access[2]{};  // this GPU, and the peer GPU when there is a peer tier (it may pull the spill)
    size_t access_n = 1;
    vgpu_vmem_stats_t st{};
    double t_sync = 0, t_launch = 0, t_wait = 0, t_unmap = 0, t_map = 0;  // where evict / ensure spend their wall time (VGPU_VMEM_TRACE)
    std::mutex mu;
    size_t block_bytes() const { return bc * kCB; }
};

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// Pinned host memory for the host tier, by the piece (one block).  Pinning costs ~40 ms per GiB (mmap + first touch on
// 16 threads + cudaHostRegister: 24 GB/s, against 55 GB/s for the transfer itself), so a piece is never given back
// to the OS while someone may need one again: a restored block's piece goes to a process-wide pool and the next block
// that leaves — of any range — takes it.  A process that pages in steady state pins nothing after the first round,
// and what is pinned is the high-water mark of what was spilled to the host at one time, not the sum of all ranges
// that ever left.  Unpinning is as slow as pinning (cudaHostUnregister + munmap: ~1 s for 16 GiB), so a restore never
// does it; only a range that is destroyed unpins its pieces, and only those the pool does not want
// (VGPU_VMEM_HOST_POOL_MB, default 8192, or as much as is spilled to the host at that moment if that is more).
struct HostPool {
    std::mutex mu;
    std::map<size_t, std::vector<void *>> free_;  // by piece size
    size_t bytes = 0;                             // idle in the pool
    size_t in_use = 0;                            // handed out: holding a spilled block
} g_host_pool;

size_t host_pool_cap() {
    const char *e = getenv("VGPU_VMEM_HOST_POOL_MB");
    return (e && *e ? strtoull(e, nullptr, 0) : 8192ull) << 20;
}
void *host_piece_get(Vmem *v, size_t bytes) {
    {
        std::lock_guard<std::mutex> lk(g_host_pool.mu);
        auto it = g_host_pool.free_.find(bytes);
        if (it != g_host_pool.free_.end() && !it->second.empty()) {
            void *p = it->second.back();
            it->second.pop_back();
            g_host_pool.bytes -= bytes;
            g_host_pool.in_use += bytes;
            return p;
        }
    }
    void *p = nullptr;
    const double ts = now_ms();
    const int rc = vgpu_slab_alloc(VGPU_TIER_HOST, v->device, -1, bytes, &p);
    v->st.tier_setup_ms += now_ms() - ts;
    if (rc != VGPU_OK) return nullptr;
    std::lock_guard<std::mutex> lk(g_host_pool.mu);
    g_host_pool.in_use += bytes;
    return p;
}
void host_piece_put(Vmem *v, void *p, size_t bytes, bool may_unpin) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(g_host_pool.mu);
        g_host_pool.in_use -= std::min(g_host_pool.in_use, bytes);
        if (!may_unpin || g_host_pool.bytes + bytes <= std::max(host_pool_cap(), g_host_pool.in_use)) {
            g_host_pool.free_[bytes].push_back(p);
            g_host_pool.bytes += bytes;
            return;
        }
    }
    vgpu_slab_free(VGPU_TIER_HOST, v->device, -1, p);
}
// where slot `slot` of tier t lives
char *slot_ptr(Vmem *v, Vmem::Tier &t, uint32_t slot) {
    return t.kind == VGPU_TIER_HOST ? (char *)t.piece[slot] : (char *)t.slab + (size_t)slot * v->block_bytes();
}
// a slot whose block is resident again (or gone) is free; its pinned piece goes back to the pool
void free_slot(Vmem *v, Vmem::Tier &t, uint32_t slot) {
    t.free_slots.push_back(slot);
    if (t.kind == VGPU_TIER_HOST && t.piece[slot]) {
        host_piece_put(v, t.piece[slot], v->block_bytes(), false);
        t.piece[slot] = nullptr;
    }
}

// Back the blocks `list` (ascending) with physical memory.  cuMemMap costs under a microsecond; cuMemSetAccess costs
// 0.2 ms per CALL whatever its size and — like cuMemUnmap — waits for the kernels that are running, so it is issued
// once per run of consecutive blocks and BEFORE any restore kernel is launched, never between two of them
// (interleaved per block, config 4 restored at 310-510 GB/s; this way the restore kernels run back to back).
// All or nothing: on failure whatever this call mapped is unmapped again and parked.
int map_blocks(Vmem *v, const std::vector<size_t> &list) {
    const double t0 = now_ms();
    size_t done = 0;
    int rc = VGPU_OK;
    for (; done < list.size(); ++done) {
        const size_t b = list[done];
        CUmemGenericAllocationHandle h;
        if (!v->pool.empty()) {
            h = v->pool.back();
            v->pool.pop_back();
        } else {
            CUresult r = g_api.Create(&h, v->block_bytes(), &v->prop, 0);
            if (r != CUDA_SUCCESS) {
                rc = drv_fail("cuMemCreate", r);
                break;
            }
        }
        CUresult r = g_api.Map(v->base + b * v->block_bytes(), v->block_bytes(), 0, h, 0);
        if (r != CUDA_SUCCESS) {
            v->pool.push_back(h);
            rc = drv_fail("cuMemMap", r);
            break;
        }
        v->blocks[b].handle = h;
    }
    for (size_t i = 0; i < done && !rc;) {  // one cuMemSetAccess per run of consecutive blocks
        size_t j = i + 1;
        while (j < done && list[j] == list[j - 1] + 1) ++j;
        CUresult r = g_api.SetAccess(v->base + list[i] * v->block_bytes(), (j - i) * v->block_bytes(), v->access, v->access_n);
        if (r != CUDA_SUCCESS) rc = drv_fail("cuMemSetAccess", r);
        i = j;
    }
    if (rc)
        for (size_t i = 0; i < done; ++i) {
            Block &bl = v->blocks[list[i]];
            g_api.Unmap(v->base + list[i] * v->block_bytes(), v->block_bytes());
            v->pool.push_back(bl.handle);
            bl.handle = 0;
        }
    v->t_map += now_ms() - t0;
    return rc;
}

double g_t_unmap_call = 0, g_t_release_call = 0, g_t_unmap_max = 0, g_t_release_max = 0;  // VGPU_VMEM_TRACE: process-wide
int unmap_block(Vmem *v, size_t b, bool release) {
    Block &bl = v->blocks[b];
    const double t0 = now_ms();
    CUresult r = g_api.Unmap(v->base + b * v->block_bytes(), v->block_bytes());
    const double t1 = now_ms();
    g_t_unmap_call += t1 - t0;
    g_t_unmap_max = std::max(g_t_unmap_max, t1 - t0);
    if (r != CUDA_SUCCESS) return drv_fail("cuMemUnmap", r);
    if (release) g_api.Release(bl.handle);  // explicit evict: the HBM goes back to the device
    else v->pool.push_back(bl.handle);      // making room inside ensure(): reused within microseconds
    const double t2 = now_ms();
    g_t_release_call += t2 - t1;
    g_t_release_max = std::max(g_t_release_max, t2 - t1);
    bl.handle = 0;
    return VGPU_OK;
}

int begin_spill(Vmem *v, size_t b);
void give_slot_back(Vmem *v, Block &bl) {
    Vmem::Tier &t = v->tiers[bl.tier_idx];
    t.free_slots.push_back(bl.tier_slot);
    t.evictions -= v->bc;
}

// Spill resident blocks `victims` to the tier, then unmap them and park or release their memory.  All spill kernels
// are queued first and run back to back; the unmaps follow when the stream is idle.  (Round 1 unmapped block i-1 while
// block i was being spilled, to overlap the two.  cuMemUnmap waits for the running kernel — 1.5 ms per 1 GiB block
// instead of 0.1 ms on an idle GPU — so nothing overlapped and the link idled meanwhile: 525 GB/s; now the NVLink rate.)
// State and counters move per block, once its bytes are safely in the tier AND its HBM is unmapped; whatever fails,
// every block is afterwards either fully resident (slot returned) or fully spilled, and the counters say so.
int evict_blocks(Vmem *v, const std::vector<size_t> &victims, bool release) {
    if (victims.empty()) return VGPU_OK;
    size_t free_total = 0;
    for (auto &t : v->tiers) free_total += t.free_slots.size();
    if (victims.size() > free_total)
        return fail(VGPU_E_NOMEM, "spill tier full: %zu blocks to evict, %zu free slots", victims.size(), free_total);
    const double t0 = now_ms();
    VGPU_CUDA(cudaDeviceSynchronize());  // whatever the application queued on the victims must be done
    v->t_sync += now_ms() - t0;
    size_t launched = 0;
    int rc = VGPU_OK;
    for (size_t i = 0; i < victims.size() && !rc; ++i) {
        const double tl = now_ms();
        rc = begin_spill(v, victims[i]);  // a failed one took nothing: the block stays resident
        v->t_launch += now_ms() - tl;
        if (!rc) launched = i + 1;
    }
    // every block whose spill was queued is in the tier once the stream is idle
    const double tw = now_ms();
    const cudaError_t se = launched ? cudaStreamSynchronize(v->stream) : cudaSuccess;
    v->t_wait += now_ms() - tw;
    if (se != cudaSuccess) {
        if (!rc) rc = fail(VGPU_E_CUDA, "cudaStreamSynchronize (evict): %s", cudaGetErrorString(se));
        for (size_t j = 0; j < launched; ++j) give_slot_back(v, v->blocks[victims[j]]);  // cannot trust the copies
        launched = 0;
    }
    const double tu = now_ms();
    for (size_t j = 0; j < launched; ++j) {
        Block &bl = v->blocks[victims[j]];
        if (int urc = unmap_block(v, victims[j], release)) {
            give_slot_back(v, bl);  // still mapped, still resident: the copy in the tier is simply dropped
            if (!rc) rc = urc;
            continue;
        }
        bl.state = kSpilled;
        v->resident -= 1;
        v->spilled += 1;
        v->st.evictions += v->bc;
    }
    v->t_unmap += now_ms() - tu;
    v->st.evict_ms += now_ms() - t0;
    return rc;
}


// ---- one block at a time: the steps evict_blocks / vgpu_vmem_ensure do in bulk, for vgpu_vmem_swap's lock step

// Pick the tier slot for block b (first tier with room whose slab can be had) and queue its spill.  -> VGPU_OK with the
// slot taken and the kernel on v->stream, or an error with nothing taken.
int begin_spill(Vmem *v, size_t b) {
    Block &bl = v->blocks[b];
    size_t ti = 0;
    uint32_t slot = 0;
    for (;; ++ti) {
        if (ti == v->tiers.size()) return fail(VGPU_E_NOMEM, "spill tier full: no tier can take block %zu", b);
        Vmem::Tier &c = v->tiers[ti];
        if (c.free_slots.empty()) continue;
        slot = c.free_slots.back();
        if (c.kind == VGPU_TIER_HOST) {  // pinned by the piece, from the pool if it has one
            if (c.piece[slot] || (c.piece[slot] = host_piece_get(v, v->block_bytes())) != nullptr) break;
        } else {
            if (c.slab) break;
            const double ts = now_ms();
            const int src = vgpu_slab_alloc(c.kind, v->device, c.peer, c.slots * v->block_bytes(), &c.slab);
            v->st.tier_setup_ms += now_ms() - ts;  // one-time: a context on the peer and its slab
            if (src == VGPU_OK) break;
        }
        // e.g. the peer GPU is not idle after all, or no more host memory can be pinned: give this tier up
        c.free_slots.clear();
        c.slots = 0;
    }
    Vmem::Tier &t = v->tiers[ti];
    bl.tier_idx = (uint8_t)ti;
    bl.tier_slot = slot;
    t.free_slots.pop_back();
    t.evictions += v->bc;
    // a range the peer was not given access to (see create) cannot be pulled, whatever the policy says by now
    const int variant = (t.kind == VGPU_TIER_PEER && v->access_n < 2 && vgpu_policy_get(VGPU_TIER_PEER, v->bc) == VGPU_XFER_PULL)
                            ? VGPU_XFER_TMA : VGPU_XFER_AUTO;
    const int rc = vgpu_spill(variant, t.kind, (const void *)(v->base + b * v->block_bytes()), slot_ptr(v, t, slot), nullptr, nullptr,
                              v->bc, kCB, v->ck + b * v->bc, v->stream);
    if (rc) give_slot_back(v, bl);
    return rc;
}
// The spill of block b has run (ok) or its stream failed (!ok): unmap it and release its memory (or park the handle in
// v->pool when the caller has a use for it), or give the slot back.
int finish_spill(Vmem *v, size_t b, bool ok, bool release) {
    Block &bl = v->blocks[b];
    if (!ok) {
        give_slot_back(v, bl);
        return fail(VGPU_E_CUDA, "spill of block %zu did not complete", b);
    }
    if (int urc = unmap_block(v, b, release)) {
        give_slot_back(v, bl);
        return urc;
    }
    bl.state = kSpilled;
    v->resident -= 1;
    v->spilled += 1;
    v->st.evictions += v->bc;
    return VGPU_OK;
}
int begin_restore(Vmem *v, size_t b, int variant) {
    Block &bl = v->blocks[b];
    Vmem::Tier &t = v->tiers[bl.tier_idx];
    v->res[b].bad_chunks = 0;
    v->res[b].first_bad_inv = 0;
    return vgpu_restore(variant, t.kind, slot_ptr(v, t, bl.tier_slot), (void *)(v->base + b * v->block_bytes()), nullptr, nullptr, v->bc,
                        kCB, v->ck + b * v->bc, nullptr, &v->res[b], v->stream);
}
// The restore of block b was queued and its stream is idle (drained) or broken: verdict, one retry through the copy
// engine, then either resident (slot freed) or unmapped again and still spilled with its tier copy intact.
int finish_restore(Vmem *v, size_t b, bool drained) {
    Block &bl = v->blocks[b];
    bool good = drained && v->res[b].bad_chunks == 0;
    if (drained && !good) {
        v->st.restore_retries += 1;
        if (begin_restore(v, b, VGPU_XFER_CE) == VGPU_OK && cudaStreamSynchronize(v->stream) == cudaSuccess) good = v->res[b].bad_chunks == 0;
    }
    if (good) {
        free_slot(v, v->tiers[bl.tier_idx], bl.tier_slot);
        bl.state = kResident;
        v->resident += 1;
        v->spilled -= 1;
        v->st.fetches += v->bc;
        return VGPU_OK;
    }
    if (drained) v->st.checksum_failures += v->res[b].bad_chunks;
    unmap_block(v, b, false);
    return fail(VGPU_E_MISMATCH, "block %zu: %llu restored chunk(s) failed their checksum (first: chunk %llu); the block stays in the tier", b,
                (unsigned long long)v->res[b].bad_chunks, (unsigned long long)(b * v->bc + (~v->res[b].first_bad_inv)));
}

}  // namespace

extern "C" {

int vgpu_vmem_create(int device, uint64_t virtual_bytes, uint64_t resident_bytes, int tier, int peer_device,
                     uint64_t tier_bytes, void **handle) {
    if (tier != VGPU_TIER_PEER && tier != VGPU_TIER_HOST) return fail(VGPU_E_INVAL, "tier must be VGPU_TIER_PEER or VGPU_TIER_HOST");
    return vgpu_vmem_create_tiered(device, virtual_bytes, resident_bytes, tier == VGPU_TIER_PEER ? peer_device : -1,
                                   tier == VGPU_TIER_PEER ? tier_bytes : 0, tier == VGPU_TIER_HOST ? tier_bytes : 0, handle);
}

int vgpu_vmem_create_tiered(int device, uint64_t virtual_bytes, uint64_t resident_bytes, int peer_device,
                            uint64_t peer_bytes, uint64_t host_bytes, void **handle) {
    if (!handle || !virtual_bytes) return fail(VGPU_E_INVAL, "bad vmem request");
    if (peer_bytes && (peer_device < 0 || peer_device == device)) return fail(VGPU_E_INVAL, "peer tier needs another GPU");
    std::call_once(g_once, load_api);
    if (!g_api.ok) return fail(VGPU_E_UNSUPPORTED, "CUDA VMM driver API not available");
    CtxScope keep;
    VGPU_CUDA(cudaSetDevice(device));
    VGPU_CUDA(cudaFree(0));
    Vmem *v = new Vmem();
    v->device = device;
    v->nchunks = (virtual_bytes + kCB - 1) / kCB;
    // block size: up to 512 chunks (1 GiB: config 4 pages at 616 / 608 GB/s incl. VMM calls, vs 509 / 285
    // with 256 MiB blocks), but keep >= 16 blocks in a range so that LRU paging has something to choose from
    size_t bc = 1;
    while (bc < 512 && bc * 2 * 16 <= v->nchunks) bc *= 2;
    if (const char *e = getenv("VGPU_VMEM_BLOCK_CHUNKS")) bc = std::max<long>(1, atol(e));
    v->bc = std::min(bc, v->nchunks);
    v->nblocks = (v->nchunks + v->bc - 1) / v->bc;
    v->resident_limit = std::max<uint64_t>(1, (resident_bytes + v->block_bytes() - 1) / v->block_bytes());  // whole blocks
    for (int k = 0; k < 2; ++k) {  // peer first, then host
        const uint64_t bytes = k == 0 ? peer_bytes : host_bytes;
        if (!bytes) continue;
        Vmem::Tier t;
        t.kind = k == 0 ? VGPU_TIER_PEER : VGPU_TIER_HOST;
        t.peer = k == 0 ? peer_device : -1;
        t.slots = (bytes + v->block_bytes() - 1) / v->block_bytes();
        t.free_slots.resize(t.slots);
        if (t.kind == VGPU_TIER_HOST) t.piece.assign(t.slots, nullptr);
        for (size_t i = 0; i < t.slots; ++i) t.free_slots[i] = (uint32_t)(t.slots - 1 - i);
        v->tiers.push_back(std::move(t));
    }
    v->prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    v->prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    v->prop.location.id = device;
    v->access[0].location = v->prop.location;
    v->access[0].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    // The peer GPU needs access to OUR range only for the PULL form of the spill (its kernel reads our HBM), i.e. when
    // the peer tier's policy can select it (after vgpu_calibrate on an idle peer).  Otherwise every map / unmap would
    // also have to update the peer's page tables for nothing.
    if (peer_bytes && vgpu_policy_get(VGPU_TIER_PEER, v->bc) == VGPU_XFER_PULL) {
        v->access[1].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        v->access[1].location.id = peer_device;
        v->access[1].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        v->access_n = 2;
    }
    size_t gran = 0;
    CUresult r = g_api.GetGranularity(&gran, &v->prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM);
    if (r != CUDA_SUCCESS || gran == 0 || kCB % gran != 0) {
        delete v;
        return r != CUDA_SUCCESS ? drv_fail("cuMemGetAllocationGranularity", r)
                                 : fail(VGPU_E_UNSUPPORTED, "allocation granularity %zu does not divide the 2 MiB chunk", gran);
    }
    r = g_api.AddressReserve(&v->base, v->nblocks * v->block_bytes(), kCB, 0, 0);
    if (r != CUDA_SUCCESS) {
        delete v;
        return drv_fail("cuMemAddressReserve", r);
    }
    int rc = VGPU_OK;  // the tier slab is allocated at the first eviction, not here
    cudaError_t e = cudaHostAlloc(&v->ck, v->nblocks * v->bc * sizeof(vgpu_cksum_t), cudaHostAllocMapped | cudaHostAllocPortable);
    if (e == cudaSuccess) e = cudaHostAlloc(&v->res, v->nblocks * sizeof(vgpu_restore_result_t), cudaHostAllocMapped | cudaHostAllocPortable);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&v->ev[0], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&v->ev[1], cudaEventDisableTiming);
    if (e != cudaSuccess) rc = fail(VGPU_E_NOMEM, "vmem staging: %s", cudaGetErrorString(e));
    if (rc) {
        vgpu_vmem_destroy(v);
        return rc;
    }
    v->blocks.resize(v->nblocks);
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
    if (n == 0) return VGPU_OK;
    std::lock_guard<std::mutex> lk(v->mu);
    CtxScope keep;
    VGPU_CUDA(cudaSetDevice(v->device));
    std::vector<size_t> victims;
    for (size_t b = first / v->bc; b <= (first + n - 1) / v->bc; ++b)
        if (v->blocks[b].state == kResident) victims.push_back(b);
    int rc = evict_blocks(v, victims, true);
    for (auto h : v->pool) g_api.Release(h);
    v->pool.clear();
    return rc;
}

int vgpu_vmem_ensure(void *handle, uint64_t first, uint64_t n) {
    Vmem *v = static_cast<Vmem *>(handle);
    if (!v || first + n > v->nchunks) return fail(VGPU_E_INVAL, "chunk range out of bounds");
    if (n == 0) return VGPU_OK;
    const size_t b0 = first / v->bc, b1 = (first + n - 1) / v->bc;
    if (b1 - b0 + 1 > v->resident_limit)
        return fail(VGPU_E_NOMEM, "%llu chunks requested, resident limit is %zu", (unsigned long long)n, v->resident_limit * v->bc);
    std::lock_guard<std::mutex> lk(v->mu);
    CtxScope keep;
    VGPU_CUDA(cudaSetDevice(v->device));
    const uint64_t stamp = ++v->clock;
    std::vector<size_t> fresh, back;
    for (size_t b = b0; b <= b1; ++b) {
        Block &bl = v->blocks[b];
        bl.last_use = stamp;
        if (bl.state == kUntouched) fresh.push_back(b);
        else if (bl.state == kSpilled) back.push_back(b);
    }
    const size_t need = fresh.size() + back.size();
    if (v->resident + need > v->resident_limit) {
        // least-recently-used resident blocks outside the requested range make room
        size_t want = v->resident + need - v->resident_limit;
        std::vector<size_t> cand;
        for (size_t b = 0; b < v->nblocks; ++b)
            if (v->blocks[b].state == kResident && v->blocks[b].last_use != stamp) cand.push_back(b);
        if (cand.size() < want) return fail(VGPU_E_NOMEM, "cannot make room: %zu blocks needed, %zu evictable", want, cand.size());
        std::nth_element(cand.begin(), cand.begin() + (want - 1), cand.end(), [&](size_t a, size_t b) {
            return v->blocks[a].last_use < v->blocks[b].last_use || (v->blocks[a].last_use == v->blocks[b].last_use && a < b);
        });
        cand.resize(want);
        std::sort(cand.begin(), cand.end());
        if (int rc = evict_blocks(v, cand, false)) return rc;
    }
    const double t0 = now_ms();
    int result = VGPU_OK;
    // never-touched blocks and spilled blocks are backed in one pass (map_blocks: all or nothing) ...
    std::vector<size_t> to_map(fresh);
    to_map.insert(to_map.end(), back.begin(), back.end());
    std::sort(to_map.begin(), to_map.end());
    result = map_blocks(v, to_map);
    if (result == VGPU_OK)
        for (size_t b : fresh) {
            v->blocks[b].state = kResident;
            v->resident += 1;
        }
    // ... then the restore kernels of the spilled ones run back to back.  Each launch reports into its own result
    // slot, so a checksum failure names its block.
    std::vector<size_t> launched;
    for (size_t i = 0; i < back.size() && !result; ++i) {
        if ((result = begin_restore(v, back[i], VGPU_XFER_AUTO)) != VGPU_OK) {
            for (size_t k = i; k < back.size(); ++k) unmap_block(v, back[k], false);  // nothing was launched into these: they stay spilled, slots taken
            break;
        }
        launched.push_back(back[i]);
    }
    if (!launched.empty()) {
        const bool drained = cudaStreamSynchronize(v->stream) == cudaSuccess;
        if (!drained && !result) result = fail(VGPU_E_CUDA, "cudaStreamSynchronize (ensure)");
        // What arrived must match what was spilled.  A block that does not is tried once more through the other mover
        // (copy engine + checksum kernel); if it still fails it is unmapped again and stays spilled — the tier holds
        // the only good copy, slot and checksums untouched, so a later ensure() can succeed (finish_restore).
        for (size_t b : launched) {
            const int f = finish_restore(v, b, drained);
            if (f != VGPU_OK && !result) result = f;
        }
    }
    v->st.fetch_ms += now_ms() - t0;
    return result;
}

// One range leaves while another arrives.  A caller whose HBM is full (the shim's residency engine, a monitor) would
// otherwise evict the victim — the link busy in one direction — and only then ensure the range it needs — the link busy
// in the other.  PCIe and NVLink are full duplex: here the two transfers run in lock step, one block each way per step,
// on the two ranges' own streams.  The VMM calls of a step (unmap + release of the block that just left, create + map +
// set-access of the block that arrives next) are issued while the GPU is idle, because cuMemUnmap / cuMemSetAccess wait
// for running kernels (see map_blocks); the arriving block only ever takes memory that a leaving block freed in THIS
// call — where both ranges have the same block size it takes that very physical block, handed over without a release /
// create — so the call never holds more HBM than before it.  All resident blocks of `out` leave; of [first, first+n) of
// `in` as many missing blocks arrive as that memory allows — vgpu_vmem_missing tells what is left, vgpu_vmem_ensure
// or another swap brings it.  Per-block state, checksum verification and the failed-restore handling are ensure()'s.
int vgpu_vmem_swap(void *out_handle, void *in_handle, uint64_t first, uint64_t n) {
    Vmem *o = static_cast<Vmem *>(out_handle), *v = static_cast<Vmem *>(in_handle);
    if (!o || !v || o == v) return fail(VGPU_E_INVAL, "swap needs two different ranges");
    if (first + n > v->nchunks) return fail(VGPU_E_INVAL, "chunk range out of bounds");
    if (o->device != v->device) return fail(VGPU_E_INVAL, "both ranges must live on one device");
    std::unique_lock<std::mutex> lo(o->mu, std::defer_lock), lv(v->mu, std::defer_lock);
    std::lock(lo, lv);
    CtxScope keep;
    VGPU_CUDA(cudaSetDevice(v->device));
    std::vector<size_t> out, in;
    for (size_t b = 0; b < o->nblocks; ++b)
        if (o->blocks[b].state == kResident) out.push_back(b);
    if (n) {
        const uint64_t stamp = ++v->clock;
        for (size_t b = first / v->bc; b <= (first + n - 1) / v->bc; ++b) {
            v->blocks[b].last_use = stamp;
            if (v->blocks[b].state != kResident) in.push_back(b);
        }
    }
    if (v->resident + in.size() > v->resident_limit)
        return fail(VGPU_E_NOMEM, "%zu blocks to bring in, resident limit is %zu", in.size(), v->resident_limit);
    size_t free_total = 0;
    for (auto &t : o->tiers) free_total += t.free_slots.size();
    if (out.size() > free_total) return fail(VGPU_E_NOMEM, "spill tier full: %zu blocks to evict, %zu free slots", out.size(), free_total);
    const double t0 = now_ms();
    VGPU_CUDA(cudaDeviceSynchronize());  // whatever the application queued on the leaving range must be done
    size_t oi = 0, ii = 0;
    uint64_t credit = 0;  // bytes freed by blocks that left, not yet taken by blocks that arrived
    int rc = VGPU_OK;
    while (!rc && (oi < out.size() || (ii < in.size() && credit >= v->block_bytes()))) {
        // -- GPU idle: the VMM calls of this step
        bool arriving = false;
        if (ii < in.size() && credit >= v->block_bytes()) {
            if ((rc = map_blocks(v, std::vector<size_t>{in[ii]})) != VGPU_OK) break;
            credit -= v->block_bytes();
            if (v->blocks[in[ii]].state == kUntouched) {  // never written: nothing to copy
                v->blocks[in[ii]].state = kResident;
                v->resident += 1;
                ++ii;
            } else {
                arriving = true;
            }
        }
        // -- both directions queued back to back, then both awaited
        bool leaving = false;
        const double tl = now_ms();
        if (oi < out.size()) {
            if ((rc = begin_spill(o, out[oi])) == VGPU_OK) leaving = true;
        }
        if (arriving) {
            const int r = begin_restore(v, in[ii], VGPU_XFER_AUTO);
            if (r != VGPU_OK) {
                unmap_block(v, in[ii], false);  // nothing was launched into it: it stays spilled, slot taken
                arriving = false;
                if (!rc) rc = r;
            }
        }
        const double tw = now_ms();
        o->t_launch += tw - tl;
        const bool so = !leaving || cudaStreamSynchronize(o->stream) == cudaSuccess;
        const bool sv = !arriving || cudaStreamSynchronize(v->stream) == cudaSuccess;
        const double tu = now_ms();
        o->t_wait += tu - tw;
        if (leaving) {
            // Same block size on both sides (the usual case: 1 GiB): the physical block that just left is handed to
            // the range that arrives — unmap there, map here — instead of being released and created anew
            // (cuMemRelease of 1 GiB takes up to 60 ms, cuMemCreate several).
            const bool hand_over = o->block_bytes() == v->block_bytes() && ii < in.size();
            const int f = finish_spill(o, out[oi], so, !hand_over);
            if (f == VGPU_OK) {
                credit += o->block_bytes();
                if (hand_over && !o->pool.empty()) {
                    v->pool.push_back(o->pool.back());
                    o->pool.pop_back();
                }
            } else if (!rc) {
                rc = f;
            }
            ++oi;
        }
        if (arriving) {
            const int f = finish_restore(v, in[ii], sv);
            if (f != VGPU_OK && !rc) rc = f;
            ++ii;
        }
        o->t_unmap += now_ms() - tu;  // unmap + release of the block that left, slot and piece bookkeeping of the one that arrived
    }
    for (Vmem *m : {o, v}) {
        for (auto h : m->pool) g_api.Release(h);
        m->pool.clear();
    }
    const double dt = now_ms() - t0;
    o->st.evict_ms += dt;
    v->st.fetch_ms += dt;
    return rc;
}

// Chunks of [first, first+n) that a vgpu_vmem_ensure of the same range would have to bring in (their block is not
// resident): what a caller needs to know to make room BEFORE it asks.
int vgpu_vmem_missing(void *handle, uint64_t first, uint64_t n, uint64_t *chunks) {
    Vmem *v = static_cast<Vmem *>(handle);
    if (!v || !chunks || first + n > v->nchunks) return fail(VGPU_E_INVAL, "chunk range out of bounds");
    *chunks = 0;
    if (n == 0) return VGPU_OK;
    std::lock_guard<std::mutex> lk(v->mu);
    for (size_t b = first / v->bc; b <= (first + n - 1) / v->bc; ++b)
        if (v->blocks[b].state != kResident) *chunks += v->bc;
    return VGPU_OK;
}

// Where a chunk lives right now: *tier = VGPU_TIER_LOCAL and *ptr = its address in the range while its block is
// resident; the tier kind and the chunk's address inside that tier's slab while it is spilled; *ptr = NULL for a
// block that was never touched.
int vgpu_vmem_locate(void *handle, uint64_t chunk, int *tier, void **ptr) {
    Vmem *v = static_cast<Vmem *>(handle);
    if (!v || !tier || !ptr || chunk >= v->nchunks) return fail(VGPU_E_INVAL, "bad locate request");
    std::lock_guard<std::mutex> lk(v->mu);
    const Block &bl = v->blocks[chunk / v->bc];
    *tier = VGPU_TIER_LOCAL;
    *ptr = nullptr;
    if (bl.state == kResident) {
        *ptr = (void *)(v->base + chunk * kCB);
    } else if (bl.state == kSpilled) {
        const Vmem::Tier &t = v->tiers[bl.tier_idx];
        *tier = t.kind;
        *ptr = slot_ptr(v, const_cast<Vmem::Tier &>(t), bl.tier_slot) + (chunk % v->bc) * kCB;
    }
    return VGPU_OK;
}

int vgpu_vmem_stats(void *handle, vgpu_vmem_stats_t *out) {
    Vmem *v = static_cast<Vmem *>(handle);
    if (!v || !out) return fail(VGPU_E_INVAL, "null argument");
    std::lock_guard<std::mutex> lk(v->mu);
    *out = v->st;
    out->chunks_total = v->nchunks;
    out->chunks_resident = v->resident * v->bc;
    out->chunks_spilled = v->spilled * v->bc;
    out->resident_limit = v->resident_limit * v->bc;
    out->tier_slots = 0;
    for (auto &t : v->tiers) out->tier_slots += t.slots * v->bc;
    out->evictions_peer = out->evictions_host = 0;
    for (auto &t : v->tiers) (t.kind == VGPU_TIER_PEER ? out->evictions_peer : out->evictions_host) += t.evictions;
    return VGPU_OK;
}

int vgpu_vmem_destroy(void *handle) {
    Vmem *v = static_cast<Vmem *>(handle);
    if (!v) return VGPU_OK;
    CtxScope keep;
    if (getenv("VGPU_VMEM_TRACE"))
        fprintf(stderr, "[vgpu_vmem] evict %.1f ms = device sync %.1f + spill launches %.1f + waits %.1f + unmap %.1f + tier set-up %.1f; fetch %.1f ms of which map %.1f\n",
                v->st.evict_ms, v->t_sync, v->t_launch, v->t_wait, v->t_unmap, v->st.tier_setup_ms, v->st.fetch_ms, v->t_map);
    if (getenv("VGPU_VMEM_TRACE"))
        fprintf(stderr, "[vgpu_vmem] process so far: cuMemUnmap %.1f ms (longest %.1f), cuMemRelease %.1f ms (longest %.1f)\n", g_t_unmap_call, g_t_unmap_max,
                g_t_release_call, g_t_release_max);
    cudaSetDevice(v->device);
    cudaDeviceSynchronize();
    for (size_t b = 0; b < v->blocks.size(); ++b)
        if (v->blocks[b].state == kResident && v->blocks[b].handle) {
            g_api.Unmap(v->base + b * v->block_bytes(), v->block_bytes());
            g_api.Release(v->blocks[b].handle);
        }
    for (auto h : v->pool) g_api.Release(h);
    if (v->base) g_api.AddressFree(v->base, v->nblocks * v->block_bytes());
    for (auto &t : v->tiers) {
        if (t.slab) vgpu_slab_free(t.kind, v->device, t.peer, t.slab);
        for (void *p : t.piece) host_piece_put(v, p, v->block_bytes(), true);  // to the pool for the ranges that remain, up to its cap
    }
    if (v->ck) cudaFreeHost(v->ck);
    if (v->res) cudaFreeHost(v->res);
    if (v->stream) cudaStreamDestroy(v->stream);
    for (auto &e : v->ev)
        if (e) cudaEventDestroy(e);
    delete v;
    return VGPU_OK;
}

}  // extern "C"
