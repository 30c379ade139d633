// Round 2 probe: does the tensor-map form of the bulk copy (cp.async.bulk.tensor, SASS
// UTMALDG / UTMASTG) move a linear HBM -> HBM copy any faster than the plain bulk copy the product uses
// (cp.async.bulk, SASS UBLKCP)?  Same skeleton for both: one producer lane per CTA, 2-stage 32 KiB ring, store lag 1,
// 3 CTAs/SM, grid-stride; copy only (no checksum warps), so the difference is the DMA form alone.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_tensor_probe tma_tensor_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s line %d: %s\n",#x,__LINE__,cudaGetErrorString(e)); exit(1);} }while(0)
#define CU(x) do{CUresult e=(x); if(e!=CUDA_SUCCESS){const char*s=0;cuGetErrorString(e,&s);printf("ERR %s line %d: %d %s\n",#x,__LINE__,(int)e,s?s:"?"); exit(1);} }while(0)
constexpr int TILE = 32768, NS = 2, ROWB = 1024, ROWS = TILE / ROWB;  // a tile = 32 rows of 1 KiB = 32 KiB contiguous
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t *b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t ph) {
    asm volatile("{\n.reg .pred P1;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra D;\nbra W;\nD:\n}" ::"r"(s32(b)), "r"(ph) : "memory");
}
template <bool TENSOR>
__global__ void __launch_bounds__(32) k_copy(const __grid_constant__ CUtensorMap tsrc, const __grid_constant__ CUtensorMap tdst,
                                            const unsigned char *src, unsigned char *dst, uint64_t ntiles) {
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ __align__(8) uint64_t full[NS];
    if (blockIdx.x >= ntiles) return;
    if (threadIdx.x == 0) {
        for (int i = 0; i < NS; ++i) mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (threadIdx.x != 0) return;
    const uint64_t my_n = (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
    for (uint64_t k = 0; k < my_n + 1; ++k) {
        if (k < my_n) {
            const int st = (int)(k % NS);
            if (k >= NS) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            const uint64_t tile = blockIdx.x + k * gridDim.x;
            mbar_expect(&full[st], TILE);
            if (TENSOR)
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s32(ring + st * TILE)),
                             "l"(&tsrc), "r"(0), "r"((int)(tile * ROWS)), "r"(s32(&full[st]))
                             : "memory");
            else
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(ring + st * TILE)),
                             "l"(src + tile * TILE), "r"(TILE), "r"(s32(&full[st]))
                             : "memory");
        }
        if (k >= 1) {
            const uint64_t j = k - 1, tile = blockIdx.x + j * gridDim.x;
            const int sj = (int)(j % NS);
            mbar_wait(&full[sj], (uint32_t)((j / NS) & 1));
            if (TENSOR)
                asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&tdst), "r"(0), "r"((int)(tile * ROWS)),
                             "r"(s32(ring + sj * TILE))
                             : "memory");
            else
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + tile * TILE), "r"(s32(ring + sj * TILE)), "r"(TILE) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
static CUtensorMap make_map(void *p, uint64_t bytes, CUtensorMapL2promotion l2) {
    CUtensorMap m;
    cuuint64_t gdim[2] = {ROWB / 4, bytes / ROWB};
    cuuint64_t gstr[1] = {ROWB};
    cuuint32_t box[2] = {ROWB / 4, ROWS};
    cuuint32_t est[2] = {1, 1};
    CU(cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, p, gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, l2,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE));
    return m;
}
int main(int argc, char **argv) {
    const double gib = argc > 1 ? atof(argv[1]) : 4.0;
    const uint64_t bytes = (uint64_t)(gib * (1ull << 30)) / TILE * TILE, ntiles = bytes / TILE;
    CK(cudaSetDevice(0));
    unsigned char *src, *dst;
    CK(cudaMalloc(&src, bytes));
    CK(cudaMalloc(&dst, bytes));
    CK(cudaMemset(src, 0x5a, bytes));
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const size_t smem = (size_t)TILE * NS;
    CK(cudaFuncSetAttribute(k_copy<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(k_copy<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    const CUtensorMapL2promotion promos[3] = {CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B};
    const char *pn[3] = {"none", "128B", "256B"};
    for (int mult : {1, 4, 16}) {
        const unsigned grid = (unsigned)(ntiles < (uint64_t)sms * 3 * mult ? ntiles : (uint64_t)sms * 3 * mult);
        for (int v = 0; v < 4; ++v) {
            CUtensorMap ts = make_map(src, bytes, promos[v ? v - 1 : 0]), td = make_map(dst, bytes, promos[v ? v - 1 : 0]);
            float best = 1e30f;
            for (int r = 0; r < 5; ++r) {
                CK(cudaMemsetAsync(dst, 0, 4096));
                CK(cudaEventRecord(a));
                if (v == 0) k_copy<false><<<grid, 32, smem>>>(ts, td, src, dst, ntiles);
                else k_copy<true><<<grid, 32, smem>>>(ts, td, src, dst, ntiles);
                CK(cudaEventRecord(b));
                CK(cudaEventSynchronize(b));
                float ms;
                CK(cudaEventElapsedTime(&ms, a, b));
                if (r && ms < best) best = ms;
            }
            CK(cudaGetLastError());
            unsigned char probe[2];
            CK(cudaMemcpy(probe, dst + bytes - 2, 2, cudaMemcpyDeviceToHost));
            printf("grid %6u (x%-2d) %-28s %7.1f GB/s payload  (last bytes %02x %02x)\n", grid, mult,
                   v == 0 ? "cp.async.bulk (UBLKCP)" : (v == 1 ? "tensor map, L2 promo none" : (v == 2 ? "tensor map, L2 promo 128B" : "tensor map, L2 promo 256B")),
                   bytes / (best * 1e6), probe[0], probe[1]);
            (void)pn;
        }
    }
    return 0;
}
