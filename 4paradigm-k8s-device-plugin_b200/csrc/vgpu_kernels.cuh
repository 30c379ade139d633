// Device code of the vGPU data path for sm_100a: memtest fill/verify and the
// fused chunk copy+checksum kernels (256-bit LD/ST variant and cp.async.bulk
// ring variant).  Everything here is HBM/NVLink/PCIe-bound integer work; no
// tensor cores (BASELINE.json north_star: "nothing on this path is a
// contraction").
//
// Reference anchors: chunk size = IPCSIZE (libvgpu.so .data@0x620c8, 2 MiB);
// the reference itself has no kernels — these replace its
// cuMemAllocManaged route (libvgpu.so@0x32146 cuMemoryAllocate) and give
// /usr/bin/vgpuvalidator (server.go:518-519) a real memtest.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "vgpu.h"

namespace vgpu {

constexpr uint64_t kGolden = 0x9E3779B97F4A7C15ull;

// ---------------------------------------------------------------------------
// small PTX helpers
// ---------------------------------------------------------------------------
struct alignas(32) U64x4 {
    uint64_t a, b, c, d;
};

// 256-bit streaming load/store (SASS: LDG.E.NA.ENL2.256 / STG.E.NA.ENL2.256).
// No .nc: the far side may be peer HBM or mapped host memory.
__device__ __forceinline__ U64x4 ld256(const void *p) {
    U64x4 v;
    asm volatile("ld.global.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(v.a), "=l"(v.b), "=l"(v.c), "=l"(v.d)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ void st256(void *p, const U64x4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(v.a),
                 "l"(v.b), "l"(v.c), "l"(v.d)
                 : "memory");
}

__device__ __forceinline__ uint64_t splitmix64_mix(uint64_t z) {  // the output function, z = state + golden
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) { return splitmix64_mix(x + kGolden); }

// Expected word `i` of a memtest pass.  `key` = pass for walking ones,
// (kGolden ^ pass ^ seed) for random.
template <int PAT>
__device__ __forceinline__ uint64_t pattern_word(uint64_t i, uint64_t key) {
    if (PAT == VGPU_PAT_WALKING_ONES) return 1ull << ((i + key) & 63);
    return splitmix64(key + i);
}

// ---------------------------------------------------------------------------
// memtest
// ---------------------------------------------------------------------------
// Layout handled by both kernels: `head` scalar words up to 32-byte alignment,
// `nvec` 32-byte vectors, `tail` scalar words.  Word indices are global from
// the start of the buffer so the CPU oracle agrees bit for bit.
//
// Shape: NOT persistent.  One CTA owns one contiguous tile of T*U vectors and
// exits; the hardware block scheduler hands out tiles in address order, which
// keeps the set of DRAM pages being written/read at any moment compact.
// Measured on B200 (profiles/r01_memtest_sweeps.md): 7.59 TB/s for fill and
// verify this way vs 6.35 / 7.2 TB/s with a 148x8 grid-stride loop.
constexpr int kMtU = 4;           // vectors per thread => 4 x 32 B in flight per thread
constexpr int kMtThreadsW = 256;  // walking ones: 7.59 TB/s fill and verify at 256 x 4
constexpr int kMtThreadsR = 128;  // random (splitmix64, ALU-heavy): 7.57 / 7.25 TB/s at 128 x 4, ~10 % less at 256

template <int PAT, int T>
__global__ void __launch_bounds__(T)
k_memtest_fill(uint64_t *__restrict__ buf, uint64_t head, uint64_t nvec, uint64_t tail,
               uint64_t key) {
    uint64_t *vbase = buf + head;
    const uint64_t base = (uint64_t)blockIdx.x * (T * kMtU) + threadIdx.x;
#pragma unroll
    for (int u = 0; u < kMtU; ++u) {
        const uint64_t v = base + (uint64_t)u * T;
        if (v < nvec) {
            const uint64_t i = head + 4 * v;
            U64x4 w;
            w.a = pattern_word<PAT>(i + 0, key);
            w.b = pattern_word<PAT>(i + 1, key);
            w.c = pattern_word<PAT>(i + 2, key);
            w.d = pattern_word<PAT>(i + 3, key);
            st256(vbase + 4 * v, w);
        }
    }
    if (blockIdx.x == 0) {  // ragged ends (at most 3 + 3 words)
        const uint64_t tid = threadIdx.x;
        if (tid < head) buf[tid] = pattern_word<PAT>(tid, key);
        if (tid < tail) {
            const uint64_t i = head + 4 * nvec + tid;
            buf[i] = pattern_word<PAT>(i, key);
        }
    }
}

struct MtBad {
    uint64_t count, bits, first;
};

__device__ __forceinline__ void mt_check(uint64_t got, uint64_t exp, uint64_t i, MtBad &b) {
    const uint64_t d = got ^ exp;
    if (d) {
        b.count += 1;
        b.bits |= d;
        b.first = i < b.first ? i : b.first;
    }
}

// Cold path: some word of this thread's tile mismatched.  Recount everything the
// thread owns with full bookkeeping.  Kept out of line so the hot loop stays at ~48
// registers (the ALU-bound random verify went 6.2 -> 7.25 TB/s from this alone).
template <int PAT>
__device__ __noinline__ void memtest_verify_slow(const uint64_t *buf, uint64_t head, uint64_t nvec,
                                                 uint64_t tail, uint64_t key, uint64_t base, int T,
                                                 vgpu_memtest_dev_result_t *res) {
    MtBad bad{0, 0, ~0ull};
    const uint64_t *vbase = buf + head;
    for (int u = 0; u < kMtU; ++u) {
        const uint64_t v = base + (uint64_t)u * T;
        if (v < nvec)
            for (int k = 0; k < 4; ++k) {
                const uint64_t i = head + 4 * v + k;
                mt_check(vbase[4 * v + k], pattern_word<PAT>(i, key), i, bad);
            }
    }
    if (blockIdx.x == 0) {
        const uint64_t tid = threadIdx.x;
        if (tid < head) mt_check(buf[tid], pattern_word<PAT>(tid, key), tid, bad);
        if (tid < tail) {
            const uint64_t i = head + 4 * nvec + tid;
            mt_check(buf[i], pattern_word<PAT>(i, key), i, bad);
        }
    }
    if (bad.count) {  // integer add / or / max: order-independent, bit-exact vs the oracle
        atomicAdd((unsigned long long *)&res->mismatches, (unsigned long long)bad.count);
        atomicOr((unsigned long long *)&res->bad_bits, (unsigned long long)bad.bits);
        atomicMax((unsigned long long *)&res->first_bad_inv, (unsigned long long)~bad.first);
    }
}

template <int PAT, int T>
__global__ void __launch_bounds__(T)
k_memtest_verify(const uint64_t *__restrict__ buf, uint64_t head, uint64_t nvec, uint64_t tail,
                 uint64_t key, vgpu_memtest_dev_result_t *__restrict__ res) {
    const uint64_t *vbase = buf + head;
    const uint64_t base = (uint64_t)blockIdx.x * (T * kMtU) + threadIdx.x;
    U64x4 w[kMtU];
#pragma unroll
    for (int u = 0; u < kMtU; ++u) {  // all loads first: kMtU x 32 B in flight per thread
        const uint64_t v = base + (uint64_t)u * T;
        if (v < nvec) w[u] = ld256(vbase + 4 * v);
    }
    uint64_t acc = 0;  // OR of got^expected over everything this thread owns
    // Random pattern: word i is mix(z), z = key + golden + i.  The 16 words of a thread lie within 4*(kMtU-1)*T + 3
    // indices of each other, so unless the low 32 bits of z wrap inside that span (one thread in ~2.8 million) they all
    // share the HIGH word of z — and with it half of the first xorshift and a third of the first multiply, which the
    // compiler then computes once per thread instead of once per word: 20 % fewer ALU-pipe instructions in a kernel that
    // is bound by that pipe (6.5 -> 7.5 TB/s on the whole GPU, profiles/r02_concurrency.md §6).  Same bits either way.
    const uint64_t zt = key + kGolden + head + 4 * base;
    if (PAT == VGPU_PAT_RANDOM && (uint32_t)zt <= 0xFFFFFFFFu - (4u * (kMtU - 1) * T + 3u)) {
        const uint32_t lo = (uint32_t)zt;
        const uint64_t hi = zt & 0xFFFFFFFF00000000ull;
#pragma unroll
        for (int u = 0; u < kMtU; ++u) {
            const uint64_t v = base + (uint64_t)u * T;
            if (v < nvec) {
                const uint32_t c = 4u * u * T;
                acc |= (w[u].a ^ splitmix64_mix(hi | (uint32_t)(lo + c + 0))) | (w[u].b ^ splitmix64_mix(hi | (uint32_t)(lo + c + 1))) |
                       (w[u].c ^ splitmix64_mix(hi | (uint32_t)(lo + c + 2))) | (w[u].d ^ splitmix64_mix(hi | (uint32_t)(lo + c + 3)));
            }
        }
    } else {
#pragma unroll
        for (int u = 0; u < kMtU; ++u) {
            const uint64_t v = base + (uint64_t)u * T;
            if (v < nvec) {
                const uint64_t i = head + 4 * v;
                acc |= (w[u].a ^ pattern_word<PAT>(i + 0, key)) | (w[u].b ^ pattern_word<PAT>(i + 1, key)) |
                       (w[u].c ^ pattern_word<PAT>(i + 2, key)) | (w[u].d ^ pattern_word<PAT>(i + 3, key));
            }
        }
    }
    if (blockIdx.x == 0) {
        const uint64_t tid = threadIdx.x;
        if (tid < head) acc |= buf[tid] ^ pattern_word<PAT>(tid, key);
        if (tid < tail) {
            const uint64_t i = head + 4 * nvec + tid;
            acc |= buf[i] ^ pattern_word<PAT>(i, key);
        }
    }
    if (acc) memtest_verify_slow<PAT>(buf, head, nvec, tail, key, base, T, res);  // rare
}

// Result block of a blocking memtest pass: zeroed before the pass and handed to the host after it by one-thread
// kernels (a store to mapped pinned memory), NOT by cudaMemsetAsync / cudaMemcpyAsync — a 24-byte copy-engine job
// queues behind whatever bulk spill another vGPU has on the same engine (40 ms for 2 GiB to the host), and every
// validator pass would wait for it.
static __global__ void k_mt_result_reset(vgpu_memtest_dev_result_t *res) { *res = vgpu_memtest_dev_result_t{0, 0, 0}; }
static __global__ void k_mt_result_publish(const vgpu_memtest_dev_result_t *res, volatile vgpu_memtest_dev_result_t *host) {
    host->mismatches = res->mismatches;
    host->first_bad_inv = res->first_bad_inv;
    host->bad_bits = res->bad_bits;
    __threadfence_system();
}

// Which SMs does a launch in the current context land on?  Every CTA ORs its %smid into a 256-bit mask and spins a
// little so that the grid really spreads over everything it may use.
static __global__ void k_smid_mask(unsigned long long *mask) {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (threadIdx.x == 0) atomicOr(&mask[(smid >> 6) & 3], 1ull << (smid & 63));
    const long long t0 = clock64();
    while (clock64() - t0 < 20000) {
    }
}

// ---------------------------------------------------------------------------
// chunk transfer + checksum
// ---------------------------------------------------------------------------
// Per-chunk accumulator in scratch memory.  Warps fold their partial (X,S) in
// with integer atomics (associative + commutative => bit-exact in any order)
// and take a ticket; the warp holding the last ticket publishes the result
// and resets the slot to zero, so the scratch is clean for the next launch
// without a memset.
struct alignas(32) ChunkAcc {
    unsigned long long x, s;
    unsigned int ticket, pad0;
    unsigned long long pad1;
};

struct XferArgs {
    const unsigned char *src;
    unsigned char *dst;
    const uint32_t *src_slot;
    const uint32_t *dst_slot;
    uint64_t chunk_bytes;
    uint32_t nchunks;
    uint32_t tiles_per_chunk;
    ChunkAcc *acc;
    vgpu_cksum_t *cksum;               // may be null
    const vgpu_cksum_t *expect;        // may be null
    vgpu_restore_result_t *result;     // required iff expect
};

__device__ __forceinline__ void cksum_word(uint64_t w, uint32_t i, uint64_t &x, uint64_t &s) {
    x ^= w;
    s += w * (uint64_t)(2u * i + 1u);
}

// Fold one tile's partial (X,S) into the chunk's accumulator and publish when the chunk is complete.
// Called by all threads of NW full warps (the whole CTA in the LD/ST kernel, the consumer warps in the bulk-copy
// kernel).  The warps first combine through shared memory so that ONE lane per tile touches the accumulator: all
// atomics of a chunk hit the same three addresses and the L2 serialises them — with one fold per warp a 2 MiB chunk
// cost 1536 same-address atomics (~5 us, the whole run time of a small batch; profiles/r02_small_batches.md), with
// one per tile it costs 192.  `stage` is NW x 2 u64 of shared memory that the caller alternates between consecutive
// tiles (so one barrier per tile is enough); `sync` is the barrier over exactly the participating warps.
template <int NW, typename Sync>
__device__ __forceinline__ void chunk_fold(const XferArgs &a, uint32_t c, uint64_t x, uint64_t s, uint32_t warp,
                                           uint64_t (*stage)[2], Sync sync, uint32_t tickets_per_chunk) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        x ^= __shfl_xor_sync(0xffffffffu, x, o);
        s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    const uint32_t lane = threadIdx.x & 31;
    if (lane == 0) {
        stage[warp][0] = x;
        stage[warp][1] = s;
    }
    sync();
    if (warp != 0) return;
    x = lane < NW ? stage[lane][0] : 0;
    s = lane < NW ? stage[lane][1] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {  // NW <= 32
        x ^= __shfl_xor_sync(0xffffffffu, x, o);
        s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    if (lane == 0) {
        ChunkAcc *acc = a.acc + c;
        atomicXor(&acc->x, (unsigned long long)x);  // fire-and-forget reductions ...
        atomicAdd(&acc->s, (unsigned long long)s);
        // ... ordered before the ticket by its release half; the acquire half lets whoever draws the last ticket read
        // every contribution.  One acq_rel atomic instead of fence + atomic + fence: the serial tail of a one-chunk
        // transfer is a chain of L2 round trips, and each fence was one more of them.
        unsigned int t;
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(&acc->ticket) : "memory");
        if (t == tickets_per_chunk - 1) {
            const uint64_t X = __ldcg(&acc->x);  // both loads in flight together (L2 is where the atomics landed)
            const uint64_t S = __ldcg(&acc->s);
            __stcg(&acc->x, 0ull);               // clean for the next launch: nobody else touches this slot any more
            __stcg(&acc->s, 0ull);
            __stcg(&acc->ticket, 0u);
            if (a.cksum) {
                a.cksum[c].x = X;
                a.cksum[c].s = S;
            }
            if (a.expect) {
                if (X != a.expect[c].x || S != a.expect[c].s) {
                    atomicAdd((unsigned long long *)&a.result->bad_chunks, 1ull);
                    atomicMax((unsigned long long *)&a.result->first_bad_inv,
                              ~(unsigned long long)c);
                }
            }
        }
    }
}

// ---- variant 1: 256-bit vectorised LD/ST --------------------------------
constexpr int kXfThreads = 256;
template <int U>
struct LdstTile {
    static constexpr uint32_t bytes = kXfThreads * 32u * U;
};

template <int U, bool COPY>
__global__ void __launch_bounds__(kXfThreads) k_xfer_ldst(const XferArgs a) {
    constexpr uint32_t kTile = LdstTile<U>::bytes;
    const uint64_t total = (uint64_t)a.nchunks * a.tiles_per_chunk;
    __shared__ uint64_t stage[2][kXfThreads / 32][2];
    uint32_t flip = 0;
    for (uint64_t tile = blockIdx.x; tile < total; tile += gridDim.x, flip ^= 1) {
        const uint32_t c = (uint32_t)(tile / a.tiles_per_chunk);
        const uint32_t t = (uint32_t)(tile % a.tiles_per_chunk);
        const uint64_t toff = (uint64_t)t * kTile;
        const unsigned char *s =
            a.src + (uint64_t)(a.src_slot ? a.src_slot[c] : c) * a.chunk_bytes + toff;
        unsigned char *d = COPY ? a.dst + (uint64_t)(a.dst_slot ? a.dst_slot[c] : c) * a.chunk_bytes + toff
                                : nullptr;
        const uint64_t left = a.chunk_bytes - toff;  // bytes of this chunk at/after the tile start
        U64x4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t off = (uint32_t)u * (kXfThreads * 32u) + threadIdx.x * 32u;
            if (off < left) v[u] = ld256(s + off);
            else v[u] = U64x4{0, 0, 0, 0};
        }
        uint64_t x = 0, sum = 0;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t off = (uint32_t)u * (kXfThreads * 32u) + threadIdx.x * 32u;
            if (off < left) {
                if (COPY) st256(d + off, v[u]);
                const uint32_t i = (uint32_t)((toff + off) >> 3);
                cksum_word(v[u].a, i + 0, x, sum);
                cksum_word(v[u].b, i + 1, x, sum);
                cksum_word(v[u].c, i + 2, x, sum);
                cksum_word(v[u].d, i + 3, x, sum);
            }
        }
        chunk_fold<kXfThreads / 32>(a, c, x, sum, threadIdx.x >> 5, stage[flip], [] { __syncthreads(); }, a.tiles_per_chunk);
    }
}

// ---- variant 2: cp.async.bulk ring (SASS: UBLKCP) -------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes,
                                         uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *gdst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
                 "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// One CTA runs an NS-stage ring of TILE-byte shared-memory buffers, warp-specialised.
// Warp 0 is the producer: one lane keeps D bulk loads in flight, issues the bulk store of each
// tile as soon as it has landed, and recycles a buffer only when (a) the consumer warps have
// signalled `done` for it and (b) wait_group.read says the store engine has finished reading it.
// Warps 1..NCW only checksum out of shared memory and arrive on `done`.  No CTA-wide barrier
// after set-up; the producer's waits never stall a consumer.  Shape chosen on B200
// (profiles/r01_memtest_sweeps.md §4b): 2 stages x 32 KiB, store lag 1, 4 consumer warps, 3 CTAs/SM:
// 3.26 TB/s payload HBM->HBM, level with the LD/ST kernel; deeper rings cost residency and lose.
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int TILE, int NS, int D, int NCW, bool COPY>
__global__ void __launch_bounds__(32 * (NCW + 1)) k_xfer_tma(const XferArgs a) {
    static_assert(D >= 1 && D < NS, "store lag must leave at least one spare buffer");
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ __align__(8) uint64_t full[NS];
    __shared__ __align__(8) uint64_t done[NS];
    const uint64_t total = (uint64_t)a.nchunks * a.tiles_per_chunk;
    if (blockIdx.x >= total) return;
    const uint64_t my_n = (total - blockIdx.x + gridDim.x - 1) / gridDim.x;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < NS; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&done[i], NCW);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto tile_geom = [&](uint64_t k, uint32_t &c, uint64_t &toff, uint32_t &len) {
        const uint64_t tile = blockIdx.x + k * gridDim.x;
        c = (uint32_t)(tile / a.tiles_per_chunk);
        toff = (uint64_t)(tile % a.tiles_per_chunk) * TILE;
        const uint64_t left = a.chunk_bytes - toff;
        len = left < (uint64_t)TILE ? (uint32_t)left : (uint32_t)TILE;
    };

    if (threadIdx.x < 32) {  // ---------------- producer warp
        if (threadIdx.x != 0) return;
        for (uint64_t k = 0; k < my_n + D; ++k) {
            if (k < my_n) {
                const int st = (int)(k % NS);
                if (k >= NS) {
                    mbar_wait(&done[st], (uint32_t)(((k / NS) - 1) & 1));  // consumers are off tile k-NS
                    if (COPY) bulk_wait_read<NS - 1 - D>();                // and so is the store engine
                }
                uint32_t c, len;
                uint64_t toff;
                tile_geom(k, c, toff, len);
                const unsigned char *s = a.src + (uint64_t)(a.src_slot ? a.src_slot[c] : c) * a.chunk_bytes + toff;
                mbar_expect_tx(&full[st], len);
                bulk_g2s(ring + (size_t)st * TILE, s, len, &full[st]);
            }
            if (COPY && k >= D) {  // tile j = k - D has had D iterations to land
                const uint64_t j = k - D;
                const int sj = (int)(j % NS);
                uint32_t c, len;
                uint64_t toff;
                tile_geom(j, c, toff, len);
                mbar_wait(&full[sj], (uint32_t)((j / NS) & 1));
                unsigned char *d = a.dst + (uint64_t)(a.dst_slot ? a.dst_slot[c] : c) * a.chunk_bytes + toff;
                bulk_s2g(d, ring + (size_t)sj * TILE, len);
                bulk_commit();
            }
        }
        if (COPY) bulk_wait_all();
        return;
    }
    // ---------------- consumer warps
    const uint32_t ct = threadIdx.x - 32;
    __shared__ uint64_t stage[2][NCW][2];
    auto consumer_sync = [] { asm volatile("bar.sync 1, %0;" ::"n"(NCW * 32) : "memory"); };  // the consumer warps only
    for (uint64_t k = 0; k < my_n; ++k) {
        uint32_t c, len;
        uint64_t toff;
        tile_geom(k, c, toff, len);
        const int st = (int)(k % NS);
        const unsigned char *buf = ring + (size_t)st * TILE;
        mbar_wait(&full[st], (uint32_t)((k / NS) & 1));
        uint64_t x = 0, sum = 0;
#pragma unroll 4
        for (uint32_t off = ct * 16u; off < len; off += NCW * 32u * 16u) {
            const ulonglong2 w = *reinterpret_cast<const ulonglong2 *>(buf + off);
            const uint32_t i = (uint32_t)((toff + off) >> 3);
            cksum_word(w.x, i + 0, x, sum);
            cksum_word(w.y, i + 1, x, sum);
        }
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&done[st]);  // this warp no longer reads the buffer
        chunk_fold<NCW>(a, c, x, sum, ct >> 5, stage[k & 1], consumer_sync, a.tiles_per_chunk);
    }
}

}  // namespace vgpu
