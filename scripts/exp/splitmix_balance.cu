// Experiment: the random-pattern memtest kernels are bound by the ALU pipe (SHF / LOP3 / IADD3, 64 lanes/clk/SM) while the
// FMA pipe (IMAD) has room.  Variants move work across: 64-bit add as one IMAD.WIDE.U32 with an opaque multiplier 1, and
// hi >> k as IMAD.HI.U32 with an opaque 2^(32-k).  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o smx splitmix_balance.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__constant__ uint32_t kOpq[4] = {1u, 4u, 32u, 2u};  // 1, 2^(32-30), 2^(32-27), 2^(32-31): values ptxas cannot fold

constexpr uint64_t kGolden = 0x9E3779B97F4A7C15ull;
struct alignas(32) U64x4 { uint64_t a, b, c, d; };
__device__ __forceinline__ U64x4 ld256(const void *p) {
    U64x4 v;
    asm volatile("ld.global.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(v.a), "=l"(v.b), "=l"(v.c), "=l"(v.d) : "l"(p));
    return v;
}
__device__ __forceinline__ void st256(void *p, const U64x4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(v.a), "l"(v.b), "l"(v.c), "l"(v.d) : "memory");
}
__device__ __forceinline__ uint64_t splitmix_plain(uint64_t x) {
    uint64_t z = x + kGolden;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ uint32_t mulhi(uint32_t x, uint32_t m) {
    uint32_t r;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(m));
    return r;
}
// z ^ (z >> k); FMA = take the high half's shift on the FMA pipe
template <int K, bool FMA>
__device__ __forceinline__ uint64_t xorshift(uint64_t z, uint32_t m) {
    if (!FMA) return z ^ (z >> K);
    uint32_t lo = (uint32_t)z, hi = (uint32_t)(z >> 32);
    uint32_t slo = __funnelshift_r(lo, hi, K);
    uint32_t shi = mulhi(hi, m);
    return ((uint64_t)(hi ^ shi) << 32) | (lo ^ slo);
}
// V bit0: wide add, bit1: shift30 on fma, bit2: shift27 on fma, bit3: shift31 on fma
template <int V>
__device__ __forceinline__ uint64_t word(uint64_t base /* per thread: key+golden+4*(block part + tid) */, uint32_t off /* compile-time after unrolling */) {
    uint64_t z;
    if (V & 1) asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(z) : "r"(kOpq[0]), "r"(off), "l"(base));
    else z = base + off;
    z = xorshift<30, (V & 2) != 0>(z, kOpq[1]) * 0xBF58476D1CE4E5B9ull;
    z = xorshift<27, (V & 4) != 0>(z, kOpq[2]) * 0x94D049BB133111EBull;
    return xorshift<31, (V & 8) != 0>(z, kOpq[3]);
}
constexpr int U = 4;
template <int V, int T>
__global__ void __launch_bounds__(T) k_verify(const uint64_t *__restrict__ buf, uint64_t nvec, uint64_t key, unsigned long long *bad) {
    const uint64_t blk = (uint64_t)blockIdx.x * (T * U);
    const uint64_t base = key + kGolden + 4 * (blk + threadIdx.x);
    U64x4 w[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const uint64_t v = blk + threadIdx.x + (uint64_t)u * T;
        if (v < nvec) w[u] = ld256(buf + 4 * v);
    }
    uint64_t acc = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const uint32_t o = 4 * (u * T);
        if (blk + threadIdx.x + (uint64_t)u * T < nvec)
            acc |= (w[u].a ^ word<V>(base, o)) | (w[u].b ^ word<V>(base, o + 1)) | (w[u].c ^ word<V>(base, o + 2)) | (w[u].d ^ word<V>(base, o + 3));
    }
    if (acc) atomicAdd(bad, 1ull);
}
template <int V, int T>
__global__ void __launch_bounds__(T) k_fill(uint64_t *__restrict__ buf, uint64_t nvec, uint64_t key) {
    const uint64_t blk = (uint64_t)blockIdx.x * (T * U);
    const uint64_t base = key + kGolden + 4 * (blk + threadIdx.x);
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const uint64_t v = blk + threadIdx.x + (uint64_t)u * T;
        const uint32_t o = 4 * (u * T);
        if (v < nvec) {
            U64x4 w{word<V>(base, o), word<V>(base, o + 1), word<V>(base, o + 2), word<V>(base, o + 3)};
            st256(buf + 4 * v, w);
        }
    }
}
__global__ void k_fill_plain(uint64_t *buf, uint64_t n, uint64_t key) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) buf[i] = splitmix_plain(key + i);
}
template <int V, int T>
static void run(uint64_t *buf, uint64_t nvec, uint64_t key, unsigned long long *bad, int grid_cap) {
    const unsigned grid = (unsigned)((nvec + T * U - 1) / (T * U));
    (void)grid_cap;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best_v = 1e9f, best_f = 1e9f;
    cudaMemset(bad, 0, 8);
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k_verify<V, T><<<grid, T>>>(buf, nvec, key, bad);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best_v) best_v = ms;
    }
    unsigned long long hb = 0; cudaMemcpy(&hb, bad, 8, cudaMemcpyDeviceToHost);
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k_fill<V, T><<<grid, T>>>(buf, nvec, key);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best_f) best_f = ms;
    }
    // the variant's fill must verify clean under the plain variant
    cudaMemset(bad, 0, 8);
    k_verify<0, 128><<<(unsigned)((nvec + 128 * U - 1) / (128 * U)), 128>>>(buf, nvec, key, bad);
    unsigned long long hb2 = 0; cudaMemcpy(&hb2, bad, 8, cudaMemcpyDeviceToHost);
    const double gb = nvec * 32.0 / 1e9;
    printf("V=%2d T=%3d  verify %7.1f GB/s  fill %7.1f GB/s  bad_vs_plain_fill=%llu  bad_of_own_fill=%llu  %s\n", V, T, gb / (best_v * 1e-3), gb / (best_f * 1e-3), hb, hb2,
           cudaGetErrorString(cudaGetLastError()));
}
int main(int argc, char **argv) {
    const uint64_t gib = argc > 1 ? strtoull(argv[1], 0, 0) : 8;
    const uint64_t n = gib << 27, nvec = n / 4, key = 0x1234567ull ^ kGolden;
    uint64_t *buf; unsigned long long *bad;
    cudaMalloc(&buf, n * 8); cudaMalloc(&bad, 8);
    k_fill_plain<<<148 * 8, 256>>>(buf, n, key);
    cudaDeviceSynchronize();
#define RUN(V) run<V, 128>(buf, nvec, key, bad, 0); k_fill_plain<<<148 * 8, 256>>>(buf, n, key); cudaDeviceSynchronize();
    RUN(0) RUN(2) RUN(4) RUN(8) RUN(6) RUN(10) RUN(12) RUN(14) RUN(0)
    run<6, 256>(buf, nvec, key, bad, 0); k_fill_plain<<<148 * 8, 256>>>(buf, n, key); cudaDeviceSynchronize();
    run<14, 256>(buf, nvec, key, bad, 0);
    return 0;
}
