/*
 * mock_cuda.c — a fake libcuda.so.1 (no GPU) under which both the reference
 * vgpu/libvgpu.so and the new shim can be executed and compared.
 * TEST INFRASTRUCTURE ONLY.  Not derived from any NVIDIA source:
 * every entry point is a few lines of bookkeeping over a fake address space.
 *
 * Behaviour knobs (environment):
 *   MOCK_GPU_COUNT   number of devices (default 1, max 16)
 *   MOCK_TOTAL_MIB   physical memory per device in MiB (default 16384)
 *   MOCK_CUDA_LOG    append one line per "real" driver call to this file
 *   MOCK_CTX_MIB     what a live context costs per device, as the mock NVML's process list reports it
 *                    (default 0: no process is ever listed, the reference's host-pid probe finds nothing)
 *   MOCK_FAIL_ALLOC_OVER  bytes; cuMemAlloc_v2 / cuMemAllocManaged above this size fail with 2 (OOM)
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <stdarg.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <unistd.h>

typedef int CUresult;
typedef int CUdevice;
typedef unsigned long long CUdeviceptr;
typedef void *CUcontext;
typedef struct { char bytes[16]; } CUuuid;

#define MAXDEV 16
#define MAXALLOC 65536

static int g_ndev = -1;
static uint64_t g_total = 0;
static uint64_t g_used[MAXDEV];
static struct { CUdeviceptr p; uint64_t n; int dev; int kind; } g_alloc[MAXALLOC];
static int g_nalloc = 0;
static CUdeviceptr g_next = 0x7f0000000000ull;
static char g_ctx_storage[MAXDEV][64];
static CUcontext g_cur;
static uint64_t g_fail_over = 0;

static void cfg(void) {
    if (g_ndev >= 0) return;
    const char *e = getenv("MOCK_GPU_COUNT");
    g_ndev = e ? atoi(e) : 1;
    if (g_ndev < 1) g_ndev = 1;
    if (g_ndev > MAXDEV) g_ndev = MAXDEV;
    e = getenv("MOCK_TOTAL_MIB");
    g_total = (e ? strtoull(e, NULL, 0) : 16384ull) << 20;
    e = getenv("MOCK_FAIL_ALLOC_OVER");
    g_fail_over = e ? strtoull(e, NULL, 0) : 0;
    g_cur = g_ctx_storage[0];
}

static void logf_(const char *fmt, ...) {
    const char *path = getenv("MOCK_CUDA_LOG");
    if (!path || !*path) return;
    FILE *f = fopen(path, "a");
    if (!f) return;
    va_list ap;
    va_start(ap, fmt);
    vfprintf(f, fmt, ap);
    va_end(ap);
    fputc('\n', f);
    fclose(f);
}

/* contexts made by cuCtxCreate: distinct handles, each bound to a device */
#define MAXCREATED 32
static char g_created[MAXCREATED][64];
static int g_created_dev[MAXCREATED];  /* -1 = free (set in cfg) / destroyed */
static int g_created_n;
static CUcontext g_stack[32];
static int g_top;

static int ctx_dev(CUcontext c) {
    for (int i = 0; i < MAXDEV; ++i)
        if (c == (CUcontext)g_ctx_storage[i]) return i;
    for (int i = 0; i < g_created_n; ++i)
        if (c == (CUcontext)g_created[i]) return g_created_dev[i];
    return -1;
}
static int cur_dev(void) {
    int d = ctx_dev(g_cur);
    return d < 0 ? 0 : d;
}

static pthread_mutex_t g_mock_mu = PTHREAD_MUTEX_INITIALIZER;  /* the allocation table is shared by the caller's threads */
static CUresult do_alloc_locked(CUdeviceptr *p, uint64_t n, int kind, const char *name, unsigned flags);
static CUresult do_alloc(CUdeviceptr *p, uint64_t n, int kind, const char *name, unsigned flags) {
    pthread_mutex_lock(&g_mock_mu);
    CUresult rc = do_alloc_locked(p, n, kind, name, flags);
    pthread_mutex_unlock(&g_mock_mu);
    return rc;
}
static CUresult do_alloc_locked(CUdeviceptr *p, uint64_t n, int kind, const char *name, unsigned flags) {
    cfg();
    int dev = cur_dev();
    if (g_fail_over && n > g_fail_over) {
        logf_("%s size=%llu flags=%u -> rc=2", name, (unsigned long long)n, flags);
        return 2;
    }
    if (g_nalloc >= MAXALLOC) return 2;
    uint64_t sz = (n + 0x1fffffull) & ~0x1fffffull;
    if (sz == 0) sz = 0x200000;
    *p = g_next;
    g_next += sz;
    g_alloc[g_nalloc].p = *p;
    g_alloc[g_nalloc].n = n;
    g_alloc[g_nalloc].dev = dev;
    g_alloc[g_nalloc].kind = kind;
    ++g_nalloc;
    g_used[dev] += n;
    logf_("%s size=%llu flags=%u -> rc=0", name, (unsigned long long)n, flags);
    return 0;
}

CUresult cuInit(unsigned flags) { (void)flags; cfg(); logf_("cuInit"); return 0; }
CUresult cuDriverGetVersion(int *v) { *v = 12090; return 0; }
CUresult cuDeviceGetCount(int *n) { cfg(); *n = g_ndev; return 0; }
CUresult cuDeviceGet(CUdevice *d, int ordinal) { cfg(); if (ordinal < 0 || ordinal >= g_ndev) return 101; *d = ordinal; return 0; }
CUresult cuDeviceGetName(char *name, int len, CUdevice d) { (void)d; snprintf(name, (size_t)len, "Mock B200"); return 0; }
CUresult cuDeviceGetAttribute(int *pi, int attrib, CUdevice d) {
    (void)d;
    switch (attrib) {
        case 16: *pi = 148; break;    /* MULTIPROCESSOR_COUNT */
        case 39: *pi = 2048; break;   /* MAX_THREADS_PER_MULTIPROCESSOR */
        case 75: *pi = 10; break;     /* COMPUTE_CAPABILITY_MAJOR */
        case 76: *pi = 0; break;
        default: *pi = 0;
    }
    return 0;
}
static void mk_uuid(CUuuid *u, int d) { memset(u, 0, sizeof(*u)); u->bytes[15] = (char)(d + 1); }
CUresult cuDeviceGetUuid(CUuuid *u, CUdevice d) { mk_uuid(u, d); return 0; }
CUresult cuDeviceGetUuid_v2(CUuuid *u, CUdevice d) { mk_uuid(u, d); return 0; }
CUresult cuDeviceGetPCIBusId(char *s, int len, CUdevice d) { snprintf(s, (size_t)len, "0000:%02x:00.0", 0x10 + d); return 0; }
CUresult cuDeviceGetByPCIBusId(CUdevice *d, const char *s) { (void)s; *d = 0; return 0; }
CUresult cuDeviceTotalMem_v2(size_t *b, CUdevice d) { (void)d; cfg(); *b = g_total; return 0; }
CUresult cuDeviceTotalMem(size_t *b, CUdevice d) { return cuDeviceTotalMem_v2(b, d); }

/* live contexts per device, read by the mock NVML (same process) to decide whether this process shows up in
 * nvmlDeviceGetComputeRunningProcesses */
static int g_ctx_refs[MAXDEV];
int mock_cuda_ctx_refs(int d) { return (d >= 0 && d < MAXDEV) ? g_ctx_refs[d] : 0; }
CUresult cuDevicePrimaryCtxRetain(CUcontext *c, CUdevice d) {
    cfg();
    if (d < 0 || d >= g_ndev) return 101;
    *c = g_ctx_storage[d];
    ++g_ctx_refs[d];
    logf_("cuDevicePrimaryCtxRetain dev=%d", d);
    return 0;
}
CUresult cuDevicePrimaryCtxRelease_v2(CUdevice d) { if (d >= 0 && d < MAXDEV && g_ctx_refs[d] > 0) --g_ctx_refs[d]; return 0; }
CUresult cuDevicePrimaryCtxRelease(CUdevice d) { return cuDevicePrimaryCtxRelease_v2(d); }
CUresult cuDevicePrimaryCtxGetState(CUdevice d, unsigned *flags, int *active) {
    *flags = 0;
    *active = (getenv("MOCK_CTX_MIB") && d >= 0 && d < MAXDEV) ? g_ctx_refs[d] > 0 : 1;  /* always-active without MOCK_CTX_MIB, as the first goldens were made */
    return 0;
}
CUresult cuDevicePrimaryCtxSetFlags_v2(CUdevice d, unsigned f) { (void)d; (void)f; return 0; }
CUresult cuDevicePrimaryCtxReset_v2(CUdevice d) { (void)d; return 0; }
CUresult cuCtxCreate_v2(CUcontext *c, unsigned f, CUdevice d) {
    (void)f;
    cfg();
    if (d < 0 || d >= g_ndev) return 101;
    if (g_created_n >= MAXCREATED) return 2;
    g_created_dev[g_created_n] = d;
    *c = g_created[g_created_n++];
    ++g_ctx_refs[d];
    if (g_top < 32) g_stack[g_top++] = g_cur;  /* a created context is pushed */
    g_cur = *c;
    logf_("cuCtxCreate_v2 dev=%d", d);
    return 0;
}
CUresult cuCtxDestroy_v2(CUcontext c) {
    for (int i = 0; i < g_created_n; ++i)
        if (c == (CUcontext)g_created[i] && g_created_dev[i] >= 0) {
            if (g_ctx_refs[g_created_dev[i]] > 0) --g_ctx_refs[g_created_dev[i]];
            g_created_dev[i] = -1;
            if (g_cur == c) g_cur = g_top > 0 ? g_stack[--g_top] : (CUcontext)g_ctx_storage[0];
            logf_("cuCtxDestroy_v2");
            return 0;
        }
    return 201; /* INVALID_CONTEXT */
}
CUresult cuCtxSetCurrent(CUcontext c) { cfg(); if (c) g_cur = c; return 0; }
CUresult cuCtxGetCurrent(CUcontext *c) { cfg(); *c = g_cur; return 0; }
CUresult cuCtxPushCurrent_v2(CUcontext c) { cfg(); if (g_top < 32) g_stack[g_top++] = g_cur; g_cur = c; return 0; }
CUresult cuCtxPopCurrent_v2(CUcontext *c) {
    cfg();
    if (c) *c = g_cur;
    if (g_top > 0) g_cur = g_stack[--g_top];
    return 0;
}
CUresult cuCtxGetDevice(CUdevice *d) { cfg(); *d = cur_dev(); return 0; }
CUresult cuCtxSynchronize(void) { return 0; }
CUresult cuCtxGetApiVersion(CUcontext c, unsigned *v) { (void)c; *v = 12090; return 0; }
CUresult cuCtxGetFlags(unsigned *f) { *f = 0; return 0; }

CUresult cuMemAlloc_v2(CUdeviceptr *p, size_t n) { return do_alloc(p, n, 0, "cuMemAlloc_v2", 0); }
CUresult cuMemAllocManaged(CUdeviceptr *p, size_t n, unsigned flags) { return do_alloc(p, n, 1, "cuMemAllocManaged", flags); }
CUresult cuMemAllocPitch_v2(CUdeviceptr *p, size_t *pitch, size_t w, size_t h, unsigned elem) {
    size_t pw = ((w + 511) / 512) * 512;
    *pitch = pw;
    (void)elem;
    return do_alloc(p, pw * h, 2, "cuMemAllocPitch_v2", elem);
}
CUresult cuMemFree_v2(CUdeviceptr p) {
    cfg();
    pthread_mutex_lock(&g_mock_mu);
    for (int i = 0; i < g_nalloc; ++i)
        if (g_alloc[i].p == p) {
            g_used[g_alloc[i].dev] -= g_alloc[i].n;
            logf_("cuMemFree_v2 size=%llu -> rc=0", (unsigned long long)g_alloc[i].n);
            g_alloc[i] = g_alloc[--g_nalloc];
            pthread_mutex_unlock(&g_mock_mu);
            return 0;
        }
    pthread_mutex_unlock(&g_mock_mu);
    logf_("cuMemFree_v2 unknown -> rc=1");
    return 1;
}
CUresult cuMemGetInfo_v2(size_t *f, size_t *t) {
    cfg();
    *t = g_total;
    *f = g_total - g_used[cur_dev()];
    logf_("cuMemGetInfo_v2");
    return 0;
}
CUresult cuMemAllocHost_v2(void **pp, size_t n) { *pp = malloc(n ? n : 1); logf_("cuMemAllocHost_v2 size=%llu", (unsigned long long)n); return *pp ? 0 : 2; }
CUresult cuMemHostAlloc(void **pp, size_t n, unsigned f) { (void)f; *pp = malloc(n ? n : 1); logf_("cuMemHostAlloc size=%llu", (unsigned long long)n); return *pp ? 0 : 2; }
CUresult cuMemFreeHost(void *p) { free(p); logf_("cuMemFreeHost"); return 0; }
CUresult cuMemHostRegister_v2(void *p, size_t n, unsigned f) { (void)p; (void)n; (void)f; return 0; }
CUresult cuMemHostUnregister(void *p) { (void)p; return 0; }
CUresult cuLaunchKernel(void *f, unsigned gx, unsigned gy, unsigned gz, unsigned bx, unsigned by, unsigned bz, unsigned shm, void *st, void **params, void **extra) {
    (void)f; (void)shm; (void)st; (void)params; (void)extra;
    logf_("cuLaunchKernel grid=%u,%u,%u block=%u,%u,%u", gx, gy, gz, bx, by, bz);
    return 0;
}
CUresult cuArray3DGetDescriptor_v2(void *d, void *a) { (void)d; (void)a; return 0; }
/* stream-ordered allocations: same bookkeeping as cuMemAlloc, logged under their own names */
CUresult cuMemAllocAsync(CUdeviceptr *p, size_t n, void *stream) { (void)stream; return do_alloc(p, n, 3, "cuMemAllocAsync", 0); }
CUresult cuMemAllocFromPoolAsync(CUdeviceptr *p, size_t n, void *pool, void *stream) { (void)pool; (void)stream; return do_alloc(p, n, 3, "cuMemAllocFromPoolAsync", 0); }
CUresult cuMemFreeAsync(CUdeviceptr p, void *stream) {
    (void)stream;
    cfg();
    for (int i = 0; i < g_nalloc; ++i)
        if (g_alloc[i].p == p) {
            g_used[g_alloc[i].dev] -= g_alloc[i].n;
            logf_("cuMemFreeAsync size=%llu -> rc=0", (unsigned long long)g_alloc[i].n);
            g_alloc[i] = g_alloc[--g_nalloc];
            return 0;
        }
    return 1;
}
/* VMM: physical memory lives from cuMemCreate until it is released AND unmapped everywhere */
#define MAXVMM 4096
static struct { unsigned long long h; uint64_t n; int dev, released, maps; } g_vmm[MAXVMM];
static struct { CUdeviceptr p; uint64_t n; int slot; } g_vmap[MAXVMM];
static int g_nvmm, g_nvmap;
static void vmm_maybe_free(int slot) {
    if (g_vmm[slot].h && g_vmm[slot].released && g_vmm[slot].maps == 0) {
        g_used[g_vmm[slot].dev] -= g_vmm[slot].n;
        logf_("vmm physical free size=%llu", (unsigned long long)g_vmm[slot].n);
        g_vmm[slot].h = 0;
    }
}
typedef struct { int type, handle_types, loc_type, loc_id; void *win32; unsigned char flags[8]; } mock_alloc_prop_t;
CUresult cuMemCreate(unsigned long long *h, size_t n, const mock_alloc_prop_t *prop, unsigned long long flags) {
    (void)flags;
    cfg();
    if (!prop || prop->loc_type != 1 || prop->loc_id < 0 || prop->loc_id >= g_ndev) return 1;
    if (g_fail_over && n > g_fail_over) return 2;
    if (g_nvmm >= MAXVMM) return 2;
    g_vmm[g_nvmm].h = 0xabc000ull + (unsigned long long)g_nvmm;
    g_vmm[g_nvmm].n = n;
    g_vmm[g_nvmm].dev = prop->loc_id;
    g_vmm[g_nvmm].released = g_vmm[g_nvmm].maps = 0;
    g_used[prop->loc_id] += n;
    *h = g_vmm[g_nvmm++].h;
    logf_("cuMemCreate size=%llu dev=%d -> rc=0", (unsigned long long)n, prop->loc_id);
    return 0;
}
static int vmm_slot(unsigned long long h) {
    for (int i = 0; i < g_nvmm; ++i)
        if (g_vmm[i].h == h && h) return i;
    return -1;
}
CUresult cuMemRelease(unsigned long long h) {
    int s = vmm_slot(h);
    if (s < 0) return 0;  /* handles this mock never made (the data-path library's): accepted as before */
    g_vmm[s].released = 1;
    vmm_maybe_free(s);
    return 0;
}
CUresult cuMemAddressReserve(CUdeviceptr *p, size_t n, size_t align, CUdeviceptr addr, unsigned long long flags) {
    (void)align; (void)addr; (void)flags;
    *p = g_next;
    g_next += (n + 0x1fffffull) & ~0x1fffffull;
    return 0;
}
CUresult cuMemAddressFree(CUdeviceptr p, size_t n) { (void)p; (void)n; return 0; }
CUresult cuMemMap(CUdeviceptr p, size_t n, size_t off, unsigned long long h, unsigned long long flags) {
    (void)off; (void)flags;
    int s = vmm_slot(h);
    if (s < 0 || g_nvmap >= MAXVMM) return 1;
    g_vmap[g_nvmap].p = p;
    g_vmap[g_nvmap].n = n;
    g_vmap[g_nvmap++].slot = s;
    ++g_vmm[s].maps;
    return 0;
}
CUresult cuMemUnmap(CUdeviceptr p, size_t n) {
    for (int i = 0; i < g_nvmap;)
        if (g_vmap[i].p >= p && g_vmap[i].p + g_vmap[i].n <= p + n) {
            int s = g_vmap[i].slot;
            g_vmap[i] = g_vmap[--g_nvmap];
            --g_vmm[s].maps;
            vmm_maybe_free(s);
        } else {
            ++i;
        }
    return 0;
}
CUresult cuMemSetAccess(CUdeviceptr p, size_t n, const void *d, size_t c) { (void)p; (void)n; (void)d; (void)c; return 0; }
CUresult cuMemGetAllocationGranularity(size_t *g, const void *prop, int opt) { (void)prop; (void)opt; *g = 0x200000; return 0; }
CUresult cuPointerGetAttribute(void *data, int attr, CUdeviceptr p) {
    if (attr == 8) { *(unsigned *)data = 1; return 0; }
    memset(data, 0, 8);
    if (attr == 9)  /* DEVICE_ORDINAL: where the allocation lives */
        for (int i = 0; i < g_nalloc; ++i)
            if (p >= g_alloc[i].p && p < g_alloc[i].p + (g_alloc[i].n ? g_alloc[i].n : 1)) *(int *)data = g_alloc[i].dev;
    return 0;
}
CUresult cuPointerGetAttributes(unsigned n, int *attrs, void **data, CUdeviceptr p) {
    (void)p;
    for (unsigned i = 0; i < n; ++i) {
        if (attrs[i] == 8) *(unsigned *)data[i] = 1;        /* IS_MANAGED: the mock says "managed" for everything */
        else if (attrs[i] == 2) *(unsigned *)data[i] = 2;   /* MEMORY_TYPE: device */
        else memset(data[i], 0, 4);
    }
    logf_("cuPointerGetAttributes n=%u", n);
    return 0;
}
CUresult cuGetErrorString(CUresult e, const char **s) { (void)e; *s = "mock"; return 0; }
CUresult cuGetErrorName(CUresult e, const char **s) { (void)e; *s = "MOCK"; return 0; }

/* private export table: >= 7 valid function pointers on a writable, page-aligned page */
static CUresult table_stub(void) { return 0; }
CUresult cuGetExportTable(const void **table, const void *id) {
    (void)id;
    static void **page;
    if (!page) {
        page = (void **)mmap(NULL, 4096, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        for (int i = 0; i < 64; ++i) page[i] = (void *)table_stub;
        page[0] = (void *)(uintptr_t)(64 * sizeof(void *));
    }
    *table = page;
    return 0;
}

static CUresult proc_lookup(const char *sym, void **pfn) {
    char buf[256];
    /* libc's own dlsym, not whatever is preloaded: the real driver does not look its entry points up through dlsym */
    typedef void *(*dlsym_t)(void *, const char *);
    static dlsym_t real;
    if (!real) real = (dlsym_t)dlvsym(RTLD_NEXT, "dlsym", "GLIBC_2.2.5");
    if (!real) real = (dlsym_t)dlvsym(RTLD_NEXT, "dlsym", "GLIBC_2.34");
    if (!real) real = dlsym;
    void *self = dlopen("libcuda.so.1", RTLD_LAZY | RTLD_NOLOAD);
    snprintf(buf, sizeof(buf), "%s_v2", sym);
    void *f = self ? real(self, buf) : NULL;
    if (!f && self) f = real(self, sym);
    *pfn = f;
    return f ? 0 : 500;
}
CUresult cuGetProcAddress(const char *sym, void **pfn, int ver, unsigned long long flags) { (void)ver; (void)flags; return proc_lookup(sym, pfn); }
CUresult cuGetProcAddress_v2(const char *sym, void **pfn, int ver, unsigned long long flags, int *st) { (void)ver; (void)flags; if (st) *st = 0; return proc_lookup(sym, pfn); }
