// Host side of the data-path C-ABI declared in include/vgpu.h: launch logic,
// per-(device,stream) scratch, the AUTO policy table, tier slabs, memtest
// drivers.  No CPU fallback: every entry point either launches the sm_100a
// kernels in vgpu_kernels.cuh (or, for VGPU_XFER_CE, the copy engine plus the
// checksum kernel) or fails.
#include <cuda_runtime.h>
#include <sys/mman.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

#include "vgpu_internal.h"
#include "vgpu_kernels.cuh"

namespace vgpu {

char *err_buf() {
    static thread_local char buf[kErrLen] = {0};
    return buf;
}

namespace {

constexpr int kMaxDev = 16;  // reference cap: "Max Gpus Per Node can't excced 16" (libvgpu.so@0x27f0d)

struct Scratch {
    ChunkAcc *acc = nullptr;  // self-cleaning per-chunk accumulators
    size_t acc_cap = 0;
    vgpu_memtest_dev_result_t *mt = nullptr;
    vgpu_memtest_dev_result_t *mt_host = nullptr;  // pinned + mapped: written by k_mt_result_publish (a pageable
                                                   // D2H copy would hold the context lock until the stream
                                                   // drains and serialise the other vGPUs' host threads)
};

struct DevState {
    int sm_count = 0;
    std::map<void *, Scratch> scratch;  // keyed by stream
};

std::mutex g_mu;
DevState g_dev[kMaxDev];

int cur_device(int *dev) {
    VGPU_CUDA(cudaGetDevice(dev));
    if (*dev < 0 || *dev >= kMaxDev) return fail(VGPU_E_INVAL, "device %d out of range", *dev);
    return VGPU_OK;
}

int sm_count(int dev, int *out) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_dev[dev].sm_count)
        VGPU_CUDA(cudaDeviceGetAttribute(&g_dev[dev].sm_count, cudaDevAttrMultiProcessorCount, dev));
    *out = g_dev[dev].sm_count;
    return VGPU_OK;
}

// Scratch is allocated lazily, zeroed once, and kept until process exit.
int get_acc(int dev, void *stream, size_t nchunks, ChunkAcc **out) {
    std::lock_guard<std::mutex> lk(g_mu);
    Scratch &s = g_dev[dev].scratch[stream];
    if (s.acc_cap < nchunks) {
        size_t cap = std::max<size_t>(4096, nchunks);
        ChunkAcc *p = nullptr;
        VGPU_CUDA(cudaMalloc(&p, cap * sizeof(ChunkAcc)));
        // zero it ON THE CALLER'S STREAM: a memset on the legacy stream is not ordered before a
        // kernel on a non-blocking stream (torch streams, green-context streams)
        VGPU_CUDA(cudaMemsetAsync(p, 0, cap * sizeof(ChunkAcc), static_cast<cudaStream_t>(stream)));
        if (s.acc) {
            // the old block may still be in use by launches queued on this stream
            VGPU_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
            VGPU_CUDA(cudaFree(s.acc));
        }
        s.acc = p;
        s.acc_cap = cap;
    }
    *out = s.acc;
    return VGPU_OK;
}

int get_mt(int dev, void *stream, vgpu_memtest_dev_result_t **out, vgpu_memtest_dev_result_t **host) {
    std::lock_guard<std::mutex> lk(g_mu);
    Scratch &s = g_dev[dev].scratch[stream];
    if (!s.mt) VGPU_CUDA(cudaMalloc(&s.mt, sizeof(*s.mt)));
    if (!s.mt_host) VGPU_CUDA(cudaHostAlloc(&s.mt_host, sizeof(*s.mt_host), cudaHostAllocMapped | cudaHostAllocPortable));
    *out = s.mt;
    *host = s.mt_host;
    return VGPU_OK;
}

// ---- AUTO policy ----------------------------------------------------------
// policy[tier] is a list of (min_chunks, variant), highest matching min wins.
// Defaults come from profiles/ (first GPU measurements); vgpu_calibrate
// overwrites them with what it measures on the box at hand.
struct Rule {
    size_t min_chunks;
    int variant;
};
std::vector<Rule> g_policy[3] = {
    {{0, VGPU_XFER_TMA}, {8, VGPU_XFER_LDST}},    // local HBM: bulk-copy ring below 8 chunks (+5-10 %), LD/ST from 8 up (+3-9 %; profiles/r02_small_batches.md)
    {{0, VGPU_XFER_TMA}},                         // peer HBM: bulk-copy ring >= LD/ST at every size (891 vs 851 GB/s at 8 chunks)
    {{0, VGPU_XFER_LDST}, {8, VGPU_XFER_CE}},    // pinned host: copy engine +5 % from 8 chunks up (55.1 vs 52.7 GB/s)
};

int policy_lookup(int tier, size_t nchunks) {
    std::lock_guard<std::mutex> lk(g_mu);
    int v = VGPU_XFER_LDST;
    size_t best = 0;
    for (const Rule &r : g_policy[tier])
        if (r.min_chunks <= nchunks && r.min_chunks >= best) {
            best = r.min_chunks;
            v = r.variant;
        }
    return v;
}

// ---- transfer launch --------------------------------------------------------
constexpr int kTmaTile = 32768, kTmaStages = 2, kTmaLag = 1, kTmaConsumerWarps = 4;  // 64 KiB ring, 3 CTAs/SM: best of the shapes swept
constexpr int kTmaGridMult = 16;  // CTAs per launch = resident CTAs x 16 (grid-stride over the rest)
// link tiers: CTAs per launch (see grid_cap).
// Measured on B200 beside ten busy memtest partitions (profiles/r02_concurrency.md): the host tier is at the PCIe rate
// from 16-32 CTAs on (51-52 GB/s) and from 64 CTAs on it takes the SMs from the tenants (memtest 7.0 -> 3.2 TB/s);
// the peer tier needs the wide grid for the last 2 % of NVLink (592 CTAs: 696 / 736 GB/s, 4736: 714 / 781).
constexpr uint64_t kGridHostLdst = 32, kGridHostTma = 16, kGridPeerLdst = 4736, kGridPeerTma = 7104;

// CTAs per launch.  Local tier (HBM -> HBM): a grid-stride loop over 32 KiB tiles with 32 CTAs per SM (4 waves
// of the 8 resident).  Unlike the write-only / read-only memtest kernels, the mixed read+write copy is FASTER with
// some persistence: 3276 GB/s payload at 32/SM vs 2520 GB/s with one tile per CTA (4096 chunks,
// profiles/r01_memtest_sweeps.md).  Link tiers (peer HBM over NVLink, pinned host over PCIe) are bound by the link,
// not by the SMs: their grid only has to cover the link's bandwidth-delay product, and every CTA beyond that just
// sits on an SM waiting for the link — on SMs that belong to the tenants' partitions when the transfer runs in
// the primary context (a whole-GPU grid of host-tier CTAs halved a concurrent memtest, profiles/r02_concurrency.md).
// vgpu_policy_set_grid overrides per tier; VGPU_XFER_CTAS_PER_SM overrides the local tier for experiments.
unsigned g_grid_override[3] = {0, 0, 0};  // CTAs per launch per tier, 0 = default (under g_mu)

uint64_t grid_cap(int tier, int variant, int sms) {
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (g_grid_override[tier]) return g_grid_override[tier];
    }
    static const long cap = [] {
        const char *e = getenv("VGPU_XFER_CTAS_PER_SM");
        return e ? atol(e) : 0L;
    }();
    const uint64_t tma_local = (uint64_t)sms * 3 * kTmaGridMult;  // 3 resident CTAs per SM (shared-memory bound)
    switch (tier) {
        case VGPU_TIER_HOST: return variant == VGPU_XFER_TMA ? kGridHostTma : kGridHostLdst;
        case VGPU_TIER_PEER: return variant == VGPU_XFER_TMA ? kGridPeerTma : kGridPeerLdst;
        default: return variant == VGPU_XFER_TMA ? tma_local : (uint64_t)sms * (uint64_t)(cap > 0 ? cap : 32);
    }
}

bool aligned32(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 31u) == 0; }

int host_and_device_readable(const void *p, bool *ok) {
    cudaPointerAttributes at{};
    cudaError_t e = cudaPointerGetAttributes(&at, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *ok = false;
        return VGPU_OK;
    }
    // pinned or managed: the host walks it to issue the copies AND the checksum kernel reads it
    *ok = at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
    return VGPU_OK;
}

int launch_xfer(int variant, int tier, bool copy, const void *src, void *dst,
                const uint32_t *src_slot, const uint32_t *dst_slot, size_t nchunks,
                size_t chunk_bytes, const vgpu_cksum_t *expect, vgpu_cksum_t *cksum,
                vgpu_restore_result_t *result, void *stream, bool restore) {
    if (nchunks == 0) return VGPU_OK;
    if (!src || (copy && !dst)) return fail(VGPU_E_INVAL, "null slab pointer");
    if (!aligned32(src) || (copy && !aligned32(dst)))
        return fail(VGPU_E_INVAL, "slab base must be 32-byte aligned");
    if (chunk_bytes == 0 || chunk_bytes % 32 != 0 || chunk_bytes > (1ull << 30))
        return fail(VGPU_E_INVAL, "chunk_bytes=%zu must be a multiple of 32 in (0, 1 GiB]", chunk_bytes);
    if (nchunks > (1u << 24)) return fail(VGPU_E_INVAL, "nchunks=%zu too large", nchunks);
    if (expect && !result) return fail(VGPU_E_INVAL, "expect given without a result block");
    if (tier < 0 || tier > 2) return fail(VGPU_E_INVAL, "tier=%d", tier);
    if (variant == VGPU_XFER_AUTO) variant = policy_lookup(tier, nchunks);
    if (variant == VGPU_XFER_PULL) variant = VGPU_XFER_TMA;  // a restore already runs on the reading side
    if (variant < VGPU_XFER_LDST || variant > VGPU_XFER_CE) return fail(VGPU_E_INVAL, "variant=%d", variant);

    int dev, sms;
    if (int rc = cur_device(&dev)) return rc;
    if (int rc = sm_count(dev, &sms)) return rc;
    XferArgs a{};
    a.src = static_cast<const unsigned char *>(src);
    a.dst = static_cast<unsigned char *>(dst);
    a.src_slot = src_slot;
    a.dst_slot = dst_slot;
    a.chunk_bytes = chunk_bytes;
    a.nchunks = (uint32_t)nchunks;
    a.cksum = cksum;
    a.expect = expect;
    a.result = result;
    if (int rc = get_acc(dev, stream, nchunks, &a.acc)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    if (variant == VGPU_XFER_CE && copy) {
        // Copy engine moves the bytes; the checksum kernel then reads the
        // local side (src for a spill, dst for a restore) at HBM speed.
        if (!src_slot && !dst_slot) {
            VGPU_CUDA(cudaMemcpyAsync(dst, src, nchunks * chunk_bytes, cudaMemcpyDefault, st));
        } else {
            bool ok1 = true, ok2 = true;
            if (src_slot) host_and_device_readable(src_slot, &ok1);
            if (dst_slot) host_and_device_readable(dst_slot, &ok2);
            if (!ok1 || !ok2)
                return fail(VGPU_E_UNSUPPORTED, "VGPU_XFER_CE needs slot tables in pinned or managed memory (read by host and GPU)");
            // coalesce runs of consecutive slots into one copy each
            size_t c = 0;
            while (c < nchunks) {
                size_t s0 = src_slot ? src_slot[c] : c, d0 = dst_slot ? dst_slot[c] : c, n = 1;
                while (c + n < nchunks && (src_slot ? src_slot[c + n] : c + n) == s0 + n &&
                       (dst_slot ? dst_slot[c + n] : c + n) == d0 + n)
                    ++n;
                VGPU_CUDA(cudaMemcpyAsync(a.dst + d0 * chunk_bytes, a.src + s0 * chunk_bytes,
                                          n * chunk_bytes, cudaMemcpyDefault, st));
                c += n;
            }
        }
        if (restore) {  // checksum what arrived
            a.src = a.dst;
            a.src_slot = a.dst_slot;
        }
        copy = false;
        variant = VGPU_XFER_LDST;
    }

    if (variant == VGPU_XFER_LDST || !copy) {
        // 32 KiB tiles (4 x 32 B per thread in flight) at every batch size: smaller tiles were tried for small batches
        // and only add folds (profiles/r02_small_batches.md)
        constexpr int kU = 4;
        constexpr uint32_t tile = LdstTile<kU>::bytes;
        a.tiles_per_chunk = (uint32_t)((chunk_bytes + tile - 1) / tile);
        const uint64_t total = (uint64_t)nchunks * a.tiles_per_chunk;
        const unsigned grid = (unsigned)std::min<uint64_t>(total, grid_cap(copy ? tier : VGPU_TIER_LOCAL, VGPU_XFER_LDST, sms));
        if (copy) k_xfer_ldst<kU, true><<<grid, kXfThreads, 0, st>>>(a);
        else k_xfer_ldst<kU, false><<<grid, kXfThreads, 0, st>>>(a);
    } else {  // VGPU_XFER_TMA
        a.tiles_per_chunk = (uint32_t)((chunk_bytes + kTmaTile - 1) / kTmaTile);
        const uint64_t total = (uint64_t)nchunks * a.tiles_per_chunk;
        constexpr size_t smem = (size_t)kTmaTile * kTmaStages;
        constexpr int threads = 32 * (kTmaConsumerWarps + 1);
        auto kfn = k_xfer_tma<kTmaTile, kTmaStages, kTmaLag, kTmaConsumerWarps, true>;
        {
            static bool attr_done[kMaxDev] = {false};  // per device, under g_mu: callers come in from several host threads
            std::lock_guard<std::mutex> lk(g_mu);
            if (!attr_done[dev]) {
                VGPU_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                attr_done[dev] = true;
            }
        }
        static_assert((220 * 1024) / (smem + 1024) == 3, "grid_cap assumes 3 resident CTAs per SM");
        const unsigned grid = (unsigned)std::min<uint64_t>(total, grid_cap(tier, VGPU_XFER_TMA, sms));
        kfn<<<grid, threads, smem, st>>>(a);
    }
    VGPU_CUDA(cudaGetLastError());
    return VGPU_OK;
}

}  // namespace
}  // namespace vgpu

using namespace vgpu;

extern "C" {

int vgpu_abi_version(void) { return VGPU_ABI_VERSION; }
const char *vgpu_last_error(void) { return err_buf(); }

int vgpu_checksum(const void *src, const uint32_t *src_slot, size_t nchunks, size_t chunk_bytes,
                  vgpu_cksum_t *cksum, void *stream) {
    if (!cksum && nchunks) return fail(VGPU_E_INVAL, "cksum is NULL");
    return launch_xfer(VGPU_XFER_LDST, VGPU_TIER_LOCAL, false, src, nullptr, src_slot, nullptr,
                       nchunks, chunk_bytes, nullptr, cksum, nullptr, stream, false);
}

// Peer-tier spill as a PULL: the bulk-copy kernel runs on the (idle) GPU that owns the spill slab,
// reads the victim chunks out of this GPU's HBM over NVLink and stores them locally.  Ordered after
// the caller's stream through an event, and the caller's stream waits for its completion, so the
// call keeps its "asynchronous on `stream`" contract.
namespace {
struct PullCtx {
    cudaStream_t stream = nullptr;  // on the peer device
    cudaEvent_t ready = nullptr;    // recorded on the caller's stream (this device)
    cudaEvent_t done = nullptr;     // recorded on the peer stream
};
std::map<std::pair<int, int>, PullCtx> g_pull;  // (this device, peer device)
std::mutex g_pull_mu;

int pull_spill(const void *src, void *dst, const uint32_t *src_slot, const uint32_t *dst_slot, size_t nchunks,
               size_t chunk_bytes, vgpu_cksum_t *cksum, void *stream) {
    int dev;
    if (int rc = cur_device(&dev)) return rc;
    cudaPointerAttributes at{};
    VGPU_CUDA(cudaPointerGetAttributes(&at, dst));
    const int peer = at.device;
    if (at.type != cudaMemoryTypeDevice || peer == dev || peer < 0 || peer >= kMaxDev)
        return fail(VGPU_E_INVAL, "VGPU_XFER_PULL needs a destination slab in another GPU's memory");
    PullCtx *pc;
    {
        // one-time set-up per (device, peer) pair, under the lock: the entry may be reached from several host threads
        std::lock_guard<std::mutex> lk(g_pull_mu);
        pc = &g_pull[{dev, peer}];
        if (!pc->stream) {
            PullCtx fresh;
            VGPU_CUDA(cudaEventCreateWithFlags(&fresh.ready, cudaEventDisableTiming));
            cudaError_t e = cudaSetDevice(peer);
            if (e == cudaSuccess) {
                e = cudaDeviceEnablePeerAccess(dev, 0);  // the peer must be able to read our HBM
                if (e == cudaErrorPeerAccessAlreadyEnabled) {
                    cudaGetLastError();  // clear the sticky "already enabled"
                    e = cudaSuccess;
                }
            }
            if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&fresh.stream, cudaStreamNonBlocking);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&fresh.done, cudaEventDisableTiming);
            cudaSetDevice(dev);
            if (e != cudaSuccess) {
                if (fresh.stream) cudaStreamDestroy(fresh.stream);
                if (fresh.done) cudaEventDestroy(fresh.done);
                cudaEventDestroy(fresh.ready);
                return fail(VGPU_E_CUDA, "pull set-up %d <- %d: %s", peer, dev, cudaGetErrorString(e));
            }
            *pc = fresh;
        }
    }
    VGPU_CUDA(cudaEventRecord(pc->ready, static_cast<cudaStream_t>(stream)));
    VGPU_CUDA(cudaSetDevice(peer));
    cudaError_t e = cudaStreamWaitEvent(pc->stream, pc->ready, 0);
    int rc = e == cudaSuccess ? launch_xfer(VGPU_XFER_TMA, VGPU_TIER_PEER, true, src, dst, src_slot, dst_slot, nchunks, chunk_bytes,
                                            nullptr, cksum, nullptr, pc->stream, false)
                              : fail(VGPU_E_CUDA, "cudaStreamWaitEvent: %s", cudaGetErrorString(e));
    if (!rc && cudaEventRecord(pc->done, pc->stream) != cudaSuccess) rc = fail(VGPU_E_CUDA, "cudaEventRecord(done)");
    cudaSetDevice(dev);
    if (rc) return rc;
    VGPU_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), pc->done, 0));
    return VGPU_OK;
}
}  // namespace

int vgpu_spill(int variant, int tier, const void *src, void *dst, const uint32_t *src_slot,
               const uint32_t *dst_slot, size_t nchunks, size_t chunk_bytes, vgpu_cksum_t *cksum,
               void *stream) {
    if (!cksum && nchunks) return fail(VGPU_E_INVAL, "cksum is NULL");
    if (variant == VGPU_XFER_AUTO && tier >= 0 && tier <= 2) variant = policy_lookup(tier, nchunks);
    if (variant == VGPU_XFER_PULL) {
        if (nchunks == 0) return VGPU_OK;
        if (tier != VGPU_TIER_PEER) return fail(VGPU_E_INVAL, "VGPU_XFER_PULL is a peer-tier spill variant");
        return pull_spill(src, dst, src_slot, dst_slot, nchunks, chunk_bytes, cksum, stream);
    }
    return launch_xfer(variant, tier, true, src, dst, src_slot, dst_slot, nchunks, chunk_bytes,
                       nullptr, cksum, nullptr, stream, false);
}

int vgpu_restore(int variant, int tier, const void *src, void *dst, const uint32_t *src_slot,
                 const uint32_t *dst_slot, size_t nchunks, size_t chunk_bytes,
                 const vgpu_cksum_t *expect, vgpu_cksum_t *cksum, vgpu_restore_result_t *result,
                 void *stream) {
    return launch_xfer(variant, tier, true, src, dst, src_slot, dst_slot, nchunks, chunk_bytes,
                       expect, cksum, result, stream, true);
}

// ---- host-buffer forms ------------------------------------------------------
static int pin_if_needed(const void *host, size_t bytes, bool *registered) {
    *registered = false;
    cudaPointerAttributes at{};
    cudaError_t e = cudaPointerGetAttributes(&at, host);
    if (e != cudaSuccess) cudaGetLastError();
    if (e == cudaSuccess && at.type == cudaMemoryTypeHost) return VGPU_OK;
    VGPU_CUDA(cudaHostRegister(const_cast<void *>(host), bytes, cudaHostRegisterPortable | cudaHostRegisterMapped));
    *registered = true;
    return VGPU_OK;
}

// The host-buffer calls run on a stream of their own (non-blocking, one per device) and stage checksums and the
// result block through pinned memory: on the legacy stream, with a pageable read-back, the call would hold the
// context lock for the length of the transfer and stall every other host thread's launches (measured: a memtest
// running beside it in four threads lost a third of its rate).  Calls on one device are serialised.
namespace {
struct HostIo {
    std::mutex mu;
    cudaStream_t st = nullptr;
    vgpu_cksum_t *d_ck = nullptr, *h_ck = nullptr;
    size_t cap = 0;
    vgpu_restore_result_t *d_rr = nullptr, *h_rr = nullptr;
};
HostIo g_hio[kMaxDev];

int hio_prepare(HostIo &h, size_t nchunks) {
    if (!h.st) VGPU_CUDA(cudaStreamCreateWithFlags(&h.st, cudaStreamNonBlocking));
    if (!h.d_rr) VGPU_CUDA(cudaMalloc(&h.d_rr, sizeof(*h.d_rr)));
    if (!h.h_rr) VGPU_CUDA(cudaHostAlloc(&h.h_rr, sizeof(*h.h_rr), cudaHostAllocDefault));
    if (h.cap < nchunks) {
        VGPU_CUDA(cudaStreamSynchronize(h.st));
        if (h.d_ck) cudaFree(h.d_ck);
        if (h.h_ck) cudaFreeHost(h.h_ck);
        h.d_ck = h.h_ck = nullptr;
        h.cap = 0;
        const size_t cap = std::max<size_t>(4096, nchunks);
        VGPU_CUDA(cudaMalloc(&h.d_ck, cap * sizeof(vgpu_cksum_t)));
        VGPU_CUDA(cudaHostAlloc(&h.h_ck, cap * sizeof(vgpu_cksum_t), cudaHostAllocDefault));
        h.cap = cap;
    }
    return VGPU_OK;
}
}  // namespace

int vgpu_spill_to_host(int variant, int device, const void *d_src, void *host, size_t nchunks,
                       size_t chunk_bytes, vgpu_cksum_t *h_cksum) {
    if (!host || !h_cksum) return fail(VGPU_E_INVAL, "null host pointer");
    if (device < 0 || device >= kMaxDev) return fail(VGPU_E_INVAL, "device %d out of range", device);
    if (nchunks == 0) return VGPU_OK;
    VGPU_CUDA(cudaSetDevice(device));
    HostIo &h = g_hio[device];
    std::lock_guard<std::mutex> lk(h.mu);
    if (int rc = hio_prepare(h, nchunks)) return rc;
    bool reg = false;
    if (int rc = pin_if_needed(host, nchunks * chunk_bytes, &reg)) return rc;
    void *h_dev = nullptr;
    int rc = VGPU_OK;
    cudaError_t e = cudaHostGetDevicePointer(&h_dev, host, 0);
    if (e != cudaSuccess) rc = fail(VGPU_E_CUDA, "cudaHostGetDevicePointer: %s", cudaGetErrorString(e));
    if (!rc) rc = vgpu_spill(variant, VGPU_TIER_HOST, d_src, h_dev, nullptr, nullptr, nchunks, chunk_bytes, h.d_ck, h.st);
    if (!rc) {
        e = cudaMemcpyAsync(h.h_ck, h.d_ck, nchunks * sizeof(vgpu_cksum_t), cudaMemcpyDeviceToHost, h.st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h.st);
        if (e != cudaSuccess) rc = fail(VGPU_E_CUDA, "spill to host: %s", cudaGetErrorString(e));
        else memcpy(h_cksum, h.h_ck, nchunks * sizeof(vgpu_cksum_t));
    } else {
        cudaStreamSynchronize(h.st);
    }
    if (reg) cudaHostUnregister(host);
    return rc;
}

int vgpu_restore_from_host(int variant, int device, const void *host, void *d_dst, size_t nchunks,
                           size_t chunk_bytes, const vgpu_cksum_t *h_expect, uint64_t *bad_chunks) {
    if (!host || !h_expect || !bad_chunks) return fail(VGPU_E_INVAL, "null host pointer");
    if (device < 0 || device >= kMaxDev) return fail(VGPU_E_INVAL, "device %d out of range", device);
    *bad_chunks = 0;
    if (nchunks == 0) return VGPU_OK;
    VGPU_CUDA(cudaSetDevice(device));
    HostIo &h = g_hio[device];
    std::lock_guard<std::mutex> lk(h.mu);
    if (int rc = hio_prepare(h, nchunks)) return rc;
    bool reg = false;
    if (int rc = pin_if_needed(host, nchunks * chunk_bytes, &reg)) return rc;
    void *h_dev = nullptr;
    int rc = VGPU_OK;
    cudaError_t e = cudaHostGetDevicePointer(&h_dev, const_cast<void *>(host), 0);
    if (e != cudaSuccess) rc = fail(VGPU_E_CUDA, "cudaHostGetDevicePointer: %s", cudaGetErrorString(e));
    if (!rc) {
        memcpy(h.h_ck, h_expect, nchunks * sizeof(vgpu_cksum_t));
        e = cudaMemcpyAsync(h.d_ck, h.h_ck, nchunks * sizeof(vgpu_cksum_t), cudaMemcpyHostToDevice, h.st);
        if (e == cudaSuccess) e = cudaMemsetAsync(h.d_rr, 0, sizeof(*h.d_rr), h.st);
        if (e != cudaSuccess) rc = fail(VGPU_E_CUDA, "stage expect: %s", cudaGetErrorString(e));
    }
    if (!rc) rc = vgpu_restore(variant, VGPU_TIER_HOST, h_dev, d_dst, nullptr, nullptr, nchunks, chunk_bytes, h.d_ck, nullptr, h.d_rr, h.st);
    if (!rc) {
        e = cudaMemcpyAsync(h.h_rr, h.d_rr, sizeof(*h.h_rr), cudaMemcpyDeviceToHost, h.st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h.st);
        if (e != cudaSuccess) rc = fail(VGPU_E_CUDA, "restore from host: %s", cudaGetErrorString(e));
    } else {
        cudaStreamSynchronize(h.st);
    }
    const vgpu_restore_result_t hres = rc ? vgpu_restore_result_t{} : *h.h_rr;
    if (reg) cudaHostUnregister(const_cast<void *>(host));
    if (rc) return rc;
    *bad_chunks = hres.bad_chunks;
    if (hres.bad_chunks)
        return fail(VGPU_E_MISMATCH, "%llu chunk(s) failed checksum, first %llu",
                    (unsigned long long)hres.bad_chunks, (unsigned long long)~hres.first_bad_inv);
    return VGPU_OK;
}

// ---- slabs --------------------------------------------------------------------
// Pinned host slabs: cudaHostAlloc pins at ~2.5 GiB/s on this box (3.2 s for 8 GiB; the first
// suspend of an 80 GiB vGPU spent 30 s there).  Transparent-huge-page backed anonymous memory,
// first-touched by several threads and then registered, is pinned at 20+ GiB/s (cudaHostRegister
// itself: 0.15 s for 8 GiB).  Falls back to cudaHostAlloc whenever a step is not available.
namespace {
std::mutex g_slab_mu;
std::map<void *, size_t> g_registered_slabs;  // host slabs that came from mmap + cudaHostRegister

void *host_slab_fast(int device, size_t bytes) {
    int can = 0;
    if (cudaDeviceGetAttribute(&can, cudaDevAttrCanUseHostPointerForRegisteredMem, device) != cudaSuccess || !can) return nullptr;
    const size_t huge = 2u << 20;
    const size_t len = (bytes + huge - 1) / huge * huge;
    void *m = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (m == MAP_FAILED) return nullptr;
    madvise(m, len, MADV_HUGEPAGE);
    unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    if (len < (64u << 20)) nt = 1;
    std::vector<std::thread> th;
    const size_t per = (len / huge + nt - 1) / nt * huge;
    for (unsigned t = 0; t < nt; ++t)
        th.emplace_back([=] {
            const size_t a = std::min(len, per * t), b = std::min(len, per * (t + 1));
            for (size_t o = a; o < b; o += 4096) static_cast<volatile char *>(m)[o] = 0;  // first touch
        });
    for (auto &x : th) x.join();
    if (cudaHostRegister(m, len, cudaHostRegisterPortable | cudaHostRegisterMapped) != cudaSuccess) {
        cudaGetLastError();
        munmap(m, len);
        return nullptr;
    }
    std::lock_guard<std::mutex> lk(g_slab_mu);
    g_registered_slabs[m] = len;
    return m;
}
}  // namespace

int vgpu_slab_alloc(int tier, int device, int peer_device, size_t bytes, void **out) {
    if (!out || !bytes) return fail(VGPU_E_INVAL, "bad slab request");
    *out = nullptr;
    switch (tier) {
        case VGPU_TIER_LOCAL:
            VGPU_CUDA(cudaSetDevice(device));
            VGPU_CUDA(cudaMalloc(out, bytes));
            return VGPU_OK;
        case VGPU_TIER_PEER: {
            if (peer_device == device) return fail(VGPU_E_INVAL, "peer == device");
            int can = 0;
            VGPU_CUDA(cudaDeviceCanAccessPeer(&can, device, peer_device));
            if (!can) return fail(VGPU_E_UNSUPPORTED, "GPU %d cannot access GPU %d", device, peer_device);
            VGPU_CUDA(cudaSetDevice(peer_device));
            VGPU_CUDA(cudaMalloc(out, bytes));
            VGPU_CUDA(cudaSetDevice(device));
            cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess)
                return fail(VGPU_E_CUDA, "cudaDeviceEnablePeerAccess(%d): %s", peer_device, cudaGetErrorString(e));
            return VGPU_OK;
        }
        case VGPU_TIER_HOST:
            VGPU_CUDA(cudaSetDevice(device));
            if ((*out = host_slab_fast(device, bytes)) != nullptr) return VGPU_OK;
            VGPU_CUDA(cudaHostAlloc(out, bytes, cudaHostAllocPortable | cudaHostAllocMapped));
            return VGPU_OK;
    }
    return fail(VGPU_E_INVAL, "tier=%d", tier);
}

int vgpu_slab_free(int tier, int device, int peer_device, void *p) {
    if (!p) return VGPU_OK;
    switch (tier) {
        case VGPU_TIER_LOCAL:
            VGPU_CUDA(cudaSetDevice(device));
            VGPU_CUDA(cudaFree(p));
            return VGPU_OK;
        case VGPU_TIER_PEER:
            VGPU_CUDA(cudaSetDevice(peer_device));
            VGPU_CUDA(cudaFree(p));
            VGPU_CUDA(cudaSetDevice(device));
            return VGPU_OK;
        case VGPU_TIER_HOST: {
            size_t len = 0;
            {
                std::lock_guard<std::mutex> lk(g_slab_mu);
                auto it = g_registered_slabs.find(p);
                if (it != g_registered_slabs.end()) {
                    len = it->second;
                    g_registered_slabs.erase(it);
                }
            }
            if (len) {
                VGPU_CUDA(cudaHostUnregister(p));
                munmap(p, len);
            } else {
                VGPU_CUDA(cudaFreeHost(p));
            }
            return VGPU_OK;
        }
    }
    return fail(VGPU_E_INVAL, "tier=%d", tier);
}

// ---- policy -------------------------------------------------------------------
int vgpu_policy_get(int tier, size_t nchunks) {
    if (tier < 0 || tier > 2) return fail(VGPU_E_INVAL, "tier=%d", tier);
    return policy_lookup(tier, nchunks);
}

int vgpu_policy_set(int tier, size_t min_chunks, int variant) {
    if (tier < 0 || tier > 2 || variant < VGPU_XFER_LDST || variant > VGPU_XFER_PULL || (variant == VGPU_XFER_PULL && tier != VGPU_TIER_PEER))
        return fail(VGPU_E_INVAL, "tier=%d variant=%d", tier, variant);
    std::lock_guard<std::mutex> lk(g_mu);
    for (Rule &r : g_policy[tier])
        if (r.min_chunks == min_chunks) {
            r.variant = variant;
            return VGPU_OK;
        }
    g_policy[tier].push_back({min_chunks, variant});
    return VGPU_OK;
}

int vgpu_policy_set_grid(int tier, unsigned ctas) {
    if (tier < 0 || tier > 2) return fail(VGPU_E_INVAL, "tier=%d", tier);
    std::lock_guard<std::mutex> lk(g_mu);
    g_grid_override[tier] = ctas;
    return VGPU_OK;
}

int vgpu_calibrate(int device, int tier, int peer_device, size_t max_chunks) {
    if (tier < 0 || tier > 2) return fail(VGPU_E_INVAL, "tier=%d", tier);
    if (max_chunks == 0) max_chunks = 512;
    const size_t chunk = VGPU_CHUNK_BYTES;
    void *near_slab = nullptr, *far_slab = nullptr;
    vgpu_cksum_t *ck = nullptr;
    int rc = vgpu_slab_alloc(VGPU_TIER_LOCAL, device, -1, max_chunks * chunk, &near_slab);
    if (!rc) rc = vgpu_slab_alloc(tier, device, peer_device, max_chunks * chunk, &far_slab);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (!rc && cudaMalloc(&ck, max_chunks * sizeof(*ck)) != cudaSuccess) rc = fail(VGPU_E_NOMEM, "cksum alloc");
    if (!rc && (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess))
        rc = fail(VGPU_E_CUDA, "event create");
    if (!rc) rc = vgpu_memtest_fill(near_slab, max_chunks * chunk, VGPU_PAT_RANDOM, 0, 0, nullptr);
    for (size_t n = 1; !rc && n <= max_chunks; n *= 8) {
        float best_ms = 0;
        int best = VGPU_XFER_LDST;
        for (int v = VGPU_XFER_LDST; v <= (tier == VGPU_TIER_PEER ? VGPU_XFER_PULL : VGPU_XFER_CE) && !rc; ++v) {
            float ms_min = 1e30f;
            for (int it = 0; it < 4 && !rc; ++it) {  // first iteration is warm-up
                cudaEventRecord(e0, 0);
                rc = vgpu_spill(v, tier, near_slab, far_slab, nullptr, nullptr, n, chunk, ck, nullptr);
                cudaEventRecord(e1, 0);
                if (cudaEventSynchronize(e1) != cudaSuccess) rc = fail(VGPU_E_CUDA, "calibrate sync");
                float ms = 0;
                cudaEventElapsedTime(&ms, e0, e1);
                if (it) ms_min = std::min(ms_min, ms);
            }
            if (v == VGPU_XFER_LDST || ms_min < best_ms) {
                best_ms = ms_min;
                best = v;
            }
        }
        if (!rc) vgpu_policy_set(tier, n == 1 ? 0 : n, best);
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (ck) cudaFree(ck);
    vgpu_slab_free(tier, device, peer_device, far_slab);
    vgpu_slab_free(VGPU_TIER_LOCAL, device, -1, near_slab);
    return rc;
}

// SM ids that kernels launched on `stream` in the CURRENT context run on (a green-context partition shows its own
// SMs only): mask[i] bit b = SM 64*i+b.  Synchronous; 4096 CTAs of 128 threads, a few microseconds each.
int vgpu_sm_probe(void *stream, uint64_t mask[4], unsigned *count) {
    if (!mask) return fail(VGPU_E_INVAL, "null mask");
    unsigned long long *d = nullptr;
    VGPU_CUDA(cudaMalloc(&d, 32));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaError_t e = cudaMemsetAsync(d, 0, 32, st);
    if (e == cudaSuccess) {
        k_smid_mask<<<4096, 128, 0, st>>>(d);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(mask, d, 32, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d);
    if (e != cudaSuccess) return fail(VGPU_E_CUDA, "sm probe: %s", cudaGetErrorString(e));
    if (count) {
        *count = 0;
        for (int i = 0; i < 4; ++i) *count += (unsigned)__builtin_popcountll(mask[i]);
    }
    return VGPU_OK;
}

// Copy-engine rate of the link behind a tier: plain cudaMemcpyAsync of `bytes` from this GPU's HBM to a
// slab of the tier (out) and back (in), best of 3 after a warm-up, CUDA events.  This is the
// "measured practical peak" the kernel rates are reported against next to the nominal link rate.
int vgpu_link_rate(int device, int tier, int peer_device, size_t bytes, double *out_gbps, double *in_gbps) {
    if (tier < 0 || tier > 2 || !bytes || !out_gbps || !in_gbps) return fail(VGPU_E_INVAL, "bad link-rate request");
    void *near_slab = nullptr, *far_slab = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = vgpu_slab_alloc(VGPU_TIER_LOCAL, device, -1, bytes, &near_slab);
    if (!rc) rc = vgpu_slab_alloc(tier, device, peer_device, bytes, &far_slab);
    if (!rc && (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess)) rc = fail(VGPU_E_CUDA, "event create");
    double best[2] = {0, 0};
    for (int dir = 0; dir < 2 && !rc; ++dir)
        for (int it = 0; it < 4 && !rc; ++it) {  // first iteration is warm-up
            cudaEventRecord(e0, 0);
            cudaError_t e = dir == 0 ? cudaMemcpyAsync(far_slab, near_slab, bytes, cudaMemcpyDefault, 0)
                                     : cudaMemcpyAsync(near_slab, far_slab, bytes, cudaMemcpyDefault, 0);
            cudaEventRecord(e1, 0);
            if (e != cudaSuccess || cudaEventSynchronize(e1) != cudaSuccess) rc = fail(VGPU_E_CUDA, "link-rate copy: %s", cudaGetErrorString(e));
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            if (it && ms > 0) best[dir] = std::max(best[dir], (double)bytes / (ms * 1e6));
        }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    vgpu_slab_free(tier, device, peer_device, far_slab);
    vgpu_slab_free(VGPU_TIER_LOCAL, device, -1, near_slab);
    if (rc) return rc;
    *out_gbps = best[0];
    *in_gbps = best[1];
    return VGPU_OK;
}

// ---- memtest ------------------------------------------------------------------
static int mt_geom(const void *buf, uint64_t bytes, uint64_t *head, uint64_t *nvec, uint64_t *tail) {
    if (!buf) return fail(VGPU_E_INVAL, "null buffer");
    if ((reinterpret_cast<uintptr_t>(buf) & 7u) || (bytes & 7u))
        return fail(VGPU_E_INVAL, "buffer and size must be 8-byte aligned");
    const uint64_t words = bytes / 8;
    uint64_t h = ((32 - (reinterpret_cast<uintptr_t>(buf) & 31u)) & 31u) / 8;
    if (h > words) h = words;
    *head = h;
    *nvec = (words - h) / 4;
    *tail = (words - h) % 4;
    return VGPU_OK;
}

static uint64_t mt_key(int pattern, uint64_t pass, uint64_t seed) {
    return pattern == VGPU_PAT_WALKING_ONES ? pass : (kGolden ^ pass ^ seed);
}

static int mt_grid(uint64_t nvec, int threads, unsigned *grid) {
    const uint64_t per = (uint64_t)threads * kMtU;
    const uint64_t want = std::max<uint64_t>(1, (nvec + per - 1) / per);
    if (want > 0x7fffffffull) return fail(VGPU_E_INVAL, "buffer too large for one launch");
    *grid = (unsigned)want;
    return VGPU_OK;
}

int vgpu_memtest_fill(void *buf, uint64_t bytes, int pattern, uint64_t pass, uint64_t seed, void *stream) {
    if (bytes == 0) return VGPU_OK;
    uint64_t head = 0, nvec = 0, tail = 0;
    unsigned grid = 0;
    if (int rc = mt_geom(buf, bytes, &head, &nvec, &tail)) return rc;
    if (pattern != VGPU_PAT_WALKING_ONES && pattern != VGPU_PAT_RANDOM) return fail(VGPU_E_INVAL, "pattern=%d", pattern);
    if (int rc = mt_grid(nvec, pattern == VGPU_PAT_RANDOM ? kMtThreadsR : kMtThreadsW, &grid)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint64_t *p = static_cast<uint64_t *>(buf);
    const uint64_t key = mt_key(pattern, pass, seed);
    if (pattern == VGPU_PAT_WALKING_ONES)
        k_memtest_fill<VGPU_PAT_WALKING_ONES, kMtThreadsW><<<grid, kMtThreadsW, 0, st>>>(p, head, nvec, tail, key);
    else
        k_memtest_fill<VGPU_PAT_RANDOM, kMtThreadsR><<<grid, kMtThreadsR, 0, st>>>(p, head, nvec, tail, key);
    VGPU_CUDA(cudaGetLastError());
    return VGPU_OK;
}

int vgpu_memtest_verify(const void *buf, uint64_t bytes, int pattern, uint64_t pass, uint64_t seed,
                        vgpu_memtest_dev_result_t *d_result, void *stream) {
    if (!d_result) return fail(VGPU_E_INVAL, "null result");
    if (bytes == 0) return VGPU_OK;
    uint64_t head = 0, nvec = 0, tail = 0;
    unsigned grid = 0;
    if (int rc = mt_geom(buf, bytes, &head, &nvec, &tail)) return rc;
    if (pattern != VGPU_PAT_WALKING_ONES && pattern != VGPU_PAT_RANDOM) return fail(VGPU_E_INVAL, "pattern=%d", pattern);
    if (int rc = mt_grid(nvec, pattern == VGPU_PAT_RANDOM ? kMtThreadsR : kMtThreadsW, &grid)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint64_t *p = static_cast<const uint64_t *>(buf);
    const uint64_t key = mt_key(pattern, pass, seed);
    if (pattern == VGPU_PAT_WALKING_ONES)
        k_memtest_verify<VGPU_PAT_WALKING_ONES, kMtThreadsW><<<grid, kMtThreadsW, 0, st>>>(p, head, nvec, tail, key, d_result);
    else
        k_memtest_verify<VGPU_PAT_RANDOM, kMtThreadsR><<<grid, kMtThreadsR, 0, st>>>(p, head, nvec, tail, key, d_result);
    VGPU_CUDA(cudaGetLastError());
    return VGPU_OK;
}

int vgpu_memtest_run(void *buf, uint64_t bytes, int pattern, uint64_t first_pass, uint64_t npasses,
                     uint64_t seed, void *stream, vgpu_memtest_result_t *out) {
    if (!out) return fail(VGPU_E_INVAL, "null result");
    int dev;
    if (int rc = cur_device(&dev)) return rc;
    vgpu_memtest_dev_result_t *d_res, *h_res;
    if (int rc = get_mt(dev, stream, &d_res, &h_res)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
    for (auto &e : ev)
        if (cudaError_t ce = cudaEventCreate(&e); ce != cudaSuccess) {
            for (auto &d : ev)
                if (d) cudaEventDestroy(d);
            return fail(VGPU_E_CUDA, "cudaEventCreate: %s", cudaGetErrorString(ce));
        }
    vgpu_memtest_result_t r{};
    r.bytes = bytes;
    r.first_bad_word = ~0ull;
    r.first_bad_pattern = r.first_bad_pass = -1;
    int rc = VGPU_OK;
    for (uint64_t p = first_pass; p < first_pass + npasses && !rc; ++p) {
        k_mt_result_reset<<<1, 1, 0, st>>>(d_res);
        cudaEventRecord(ev[0], st);
        rc = vgpu_memtest_fill(buf, bytes, pattern, p, seed, stream);
        cudaEventRecord(ev[1], st);
        if (!rc) rc = vgpu_memtest_verify(buf, bytes, pattern, p, seed, d_res, stream);
        cudaEventRecord(ev[2], st);
        if (!rc) {
            k_mt_result_publish<<<1, 1, 0, st>>>(d_res, h_res);  // mapped pinned memory: no copy-engine job (see the kernel)
            cudaError_t e = cudaGetLastError();
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            if (e != cudaSuccess) rc = fail(VGPU_E_CUDA, "memtest pass %llu: %s", (unsigned long long)p, cudaGetErrorString(e));
        }
        if (rc) break;
        const vgpu_memtest_dev_result_t h = *h_res;
        float f = 0, v = 0;
        cudaEventElapsedTime(&f, ev[0], ev[1]);
        cudaEventElapsedTime(&v, ev[1], ev[2]);
        r.fill_ms += f;
        r.verify_ms += v;
        r.passes += 1;
        r.mismatches += h.mismatches;
        r.bad_bits |= h.bad_bits;
        if (h.mismatches && r.first_bad_pattern < 0) {
            r.first_bad_pattern = pattern;
            r.first_bad_pass = (int32_t)p;
            r.first_bad_word = ~h.first_bad_inv;
        }
    }
    for (auto &e : ev) cudaEventDestroy(e);
    if (rc) return rc;
    const double ms = r.fill_ms + r.verify_ms;
    r.gbps = ms > 0 ? 2.0 * (double)bytes * (double)r.passes / (ms * 1e6) : 0.0;
    *out = r;
    return VGPU_OK;
}

int vgpu_memtest(int device, uint64_t bytes, unsigned patterns, uint64_t random_passes, uint64_t seed,
                 vgpu_memtest_result_t *out) {
    if (!out) return fail(VGPU_E_INVAL, "null result");
    if (!(patterns & 3u)) return fail(VGPU_E_INVAL, "patterns mask selects nothing");
    VGPU_CUDA(cudaSetDevice(device));
    void *buf = nullptr;
    cudaError_t e = cudaMalloc(&buf, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(VGPU_E_NOMEM, "cudaMalloc(%llu): %s", (unsigned long long)bytes, cudaGetErrorString(e));
    }
    vgpu_memtest_result_t total{};
    total.bytes = bytes;
    total.first_bad_word = ~0ull;
    total.first_bad_pattern = total.first_bad_pass = -1;
    int rc = VGPU_OK;
    for (int pat = 0; pat < 2 && !rc; ++pat) {
        if (!(patterns & (1u << pat))) continue;
        vgpu_memtest_result_t r{};
        rc = vgpu_memtest_run(buf, bytes, pat, 0, pat == VGPU_PAT_WALKING_ONES ? 64 : random_passes, seed, nullptr, &r);
        if (rc) break;
        total.passes += r.passes;
        total.mismatches += r.mismatches;
        total.bad_bits |= r.bad_bits;
        total.fill_ms += r.fill_ms;
        total.verify_ms += r.verify_ms;
        if (r.mismatches && total.first_bad_pattern < 0) {
            total.first_bad_pattern = r.first_bad_pattern;
            total.first_bad_pass = r.first_bad_pass;
            total.first_bad_word = r.first_bad_word;
        }
    }
    cudaFree(buf);
    if (rc) return rc;
    const double ms = total.fill_ms + total.verify_ms;
    total.gbps = ms > 0 ? 2.0 * (double)bytes * (double)total.passes / (ms * 1e6) : 0.0;
    *out = total;
    return total.mismatches ? fail(VGPU_E_MISMATCH, "%llu mismatching words", (unsigned long long)total.mismatches) : VGPU_OK;
}

}  // extern "C"
