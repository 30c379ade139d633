// Spatial SM partitioning with Blackwell green contexts.
//
// Replaces the reference's time-averaged launch throttle — rate_limiter
// (libvgpu.so@0x463b1), the 120 ms utilization_watcher (@0x471a7) and the
// token arithmetic in delta (@0x46712) — with a hard spatial cap: a vGPU whose
// CUDA_DEVICE_SM_LIMIT is p gets floor(SMs*p/100) SMs (rounded by the driver
// to its granule) and its kernels can only ever be scheduled there.
//
// The driver API is resolved with dlopen/dlsym so that libvgpu_dp.so still
// loads (and its symbol table can be checked) on a machine without a GPU.
#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstring>
#include <mutex>

#include "vgpu_internal.h"

using namespace vgpu;

namespace {

struct Drv {
    void *lib = nullptr;
    CUresult (*DeviceGet)(CUdevice *, int) = nullptr;
    CUresult (*DeviceGetDevResource)(CUdevice, CUdevResource *, CUdevResourceType) = nullptr;
    CUresult (*DevSmResourceSplitByCount)(CUdevResource *, unsigned int *, const CUdevResource *, CUdevResource *,
                                          unsigned int, unsigned int) = nullptr;
    CUresult (*DevResourceGenerateDesc)(CUdevResourceDesc *, CUdevResource *, unsigned int) = nullptr;
    CUresult (*GreenCtxCreate)(CUgreenCtx *, CUdevResourceDesc, CUdevice, unsigned int) = nullptr;
    CUresult (*GreenCtxDestroy)(CUgreenCtx) = nullptr;
    CUresult (*CtxFromGreenCtx)(CUcontext *, CUgreenCtx) = nullptr;
    CUresult (*GreenCtxStreamCreate)(CUstream *, CUgreenCtx, unsigned int, int) = nullptr;
    CUresult (*GreenCtxGetDevResource)(CUgreenCtx, CUdevResource *, CUdevResourceType) = nullptr;
    CUresult (*StreamDestroy)(CUstream) = nullptr;
    CUresult (*GetErrorString)(CUresult, const char **) = nullptr;
};

Drv g_drv;
std::once_flag g_once;
bool g_ok = false;

template <typename F>
bool sym(F &fn, const char *name) {
    fn = reinterpret_cast<F>(dlsym(g_drv.lib, name));
    return fn != nullptr;
}

void load_driver() {
    g_drv.lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
    if (!g_drv.lib) return;
    g_ok = sym(g_drv.DeviceGet, "cuDeviceGet") && sym(g_drv.DeviceGetDevResource, "cuDeviceGetDevResource") &&
           sym(g_drv.DevSmResourceSplitByCount, "cuDevSmResourceSplitByCount") &&
           sym(g_drv.DevResourceGenerateDesc, "cuDevResourceGenerateDesc") &&
           sym(g_drv.GreenCtxCreate, "cuGreenCtxCreate") && sym(g_drv.GreenCtxDestroy, "cuGreenCtxDestroy") &&
           sym(g_drv.CtxFromGreenCtx, "cuCtxFromGreenCtx") &&
           sym(g_drv.GreenCtxStreamCreate, "cuGreenCtxStreamCreate") &&
           sym(g_drv.GreenCtxGetDevResource, "cuGreenCtxGetDevResource") &&
           sym(g_drv.StreamDestroy, "cuStreamDestroy_v2") && sym(g_drv.GetErrorString, "cuGetErrorString");
}

int drv_fail(const char *what, CUresult r) {
    const char *s = nullptr;
    if (g_drv.GetErrorString) g_drv.GetErrorString(r, &s);
    return fail(VGPU_E_CUDA, "%s -> %d (%s)", what, (int)r, s ? s : "?");
}

}  // namespace

extern "C" {

int vgpu_sm_partition_create(int device, unsigned percent, unsigned n, vgpu_sm_partition_t *out) {
    return vgpu_sm_partition_create_ex(device, percent, n, 0, out);
}

int vgpu_sm_partition_create_ex(int device, unsigned percent, unsigned n, unsigned flags, vgpu_sm_partition_t *out) {
    const unsigned split_flags = (flags & VGPU_SM_FINE) ? CU_DEV_SM_RESOURCE_SPLIT_IGNORE_SM_COSCHEDULING : 0;
    if (!out || n == 0 || n > 64) return fail(VGPU_E_INVAL, "bad partition request");
    if (percent == 0 || percent > 100) return fail(VGPU_E_INVAL, "percent=%u must be in 1..100", percent);
    std::call_once(g_once, load_driver);
    if (!g_ok) return fail(VGPU_E_UNSUPPORTED, "libcuda.so.1 with green-context support not available");
    VGPU_CUDA(cudaSetDevice(device));
    VGPU_CUDA(cudaFree(0));  // make sure the primary context exists
    CUdevice dev;
    CUresult r = g_drv.DeviceGet(&dev, device);
    if (r != CUDA_SUCCESS) return drv_fail("cuDeviceGet", r);
    CUdevResource all;
    r = g_drv.DeviceGetDevResource(dev, &all, CU_DEV_RESOURCE_TYPE_SM);
    if (r != CUDA_SUCCESS) return drv_fail("cuDeviceGetDevResource", r);
    const unsigned total = all.sm.smCount;
    unsigned want = total * percent / 100;
    if (want < 1) want = 1;
    // The driver rounds minCount up to its granule (8 SMs on sm_90+), which can leave fewer
    // than n groups (148 SMs, 25 %: 37 -> 40 -> only 3 groups).  A vGPU must never get MORE
    // than its share, so step the ask down to the largest count that still yields n
    // disjoint groups (148 SMs, 25 %, n=4: 4 x 32).  Co-scheduling (clusters, cta_group::2)
    // is kept unless the caller passes VGPU_SM_FINE.
    CUdevResource groups[64], rest;
    unsigned nb = 0;
    for (;; --want) {
        nb = n;
        r = g_drv.DevSmResourceSplitByCount(groups, &nb, &all, &rest, split_flags, want);
        const bool formed = r == CUDA_SUCCESS && nb >= n;
        if (formed && (groups[0].sm.smCount <= total * percent / 100 || want <= 1)) break;
        if (want <= 1) {
            if (r != CUDA_SUCCESS) return drv_fail("cuDevSmResourceSplitByCount", r);
            return fail(VGPU_E_NOMEM, "driver formed %u partitions, %u requested", nb, n);
        }
    }
    memset(out, 0, sizeof(*out) * n);
    for (unsigned i = 0; i < n; ++i) {
        CUdevResourceDesc desc;
        CUgreenCtx g;
        CUcontext c;
        CUstream s;
        r = g_drv.DevResourceGenerateDesc(&desc, &groups[i], 1);
        if (r == CUDA_SUCCESS) r = g_drv.GreenCtxCreate(&g, desc, dev, CU_GREEN_CTX_DEFAULT_STREAM);
        if (r != CUDA_SUCCESS) {
            vgpu_sm_partition_destroy(out, i);
            return drv_fail("cuGreenCtxCreate", r);
        }
        out[i].green_ctx = g;
        r = g_drv.CtxFromGreenCtx(&c, g);
        if (r == CUDA_SUCCESS) r = g_drv.GreenCtxStreamCreate(&s, g, CU_STREAM_NON_BLOCKING, 0);
        if (r != CUDA_SUCCESS) {
            vgpu_sm_partition_destroy(out, i + 1);
            return drv_fail("cuGreenCtxStreamCreate", r);
        }
        CUdevResource got;
        r = g_drv.GreenCtxGetDevResource(g, &got, CU_DEV_RESOURCE_TYPE_SM);
        out[i].ctx = c;
        out[i].stream = s;
        out[i].sm_count = r == CUDA_SUCCESS ? got.sm.smCount : groups[i].sm.smCount;
        out[i].sm_total = total;
    }
    return VGPU_OK;
}

int vgpu_sm_partition_destroy(vgpu_sm_partition_t *parts, unsigned n) {
    if (!parts) return VGPU_OK;
    if (!g_ok) return fail(VGPU_E_UNSUPPORTED, "driver not loaded");
    for (unsigned i = 0; i < n; ++i) {
        if (parts[i].stream) g_drv.StreamDestroy((CUstream)parts[i].stream);
        if (parts[i].green_ctx) g_drv.GreenCtxDestroy((CUgreenCtx)parts[i].green_ctx);
        memset(&parts[i], 0, sizeof(parts[i]));
    }
    return VGPU_OK;
}

}  // extern "C"
