/* mock_nvml.c — fake libnvidia-ml.so.1 for the differential quota harness.
 * TEST INFRASTRUCTURE ONLY.  By default no running processes are
 * reported, so the reference's set_task_pid finds nothing, context_size stays 0 and its
 * rate limiter is bypassed: deterministic quota traces.  With MOCK_CTX_MIB=<n> this process is
 * listed on every device it holds a (mock) context on, using n MiB — which is how a real node
 * looks to the host-pid probe, and what makes the reference charge the context to the quota. */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <unistd.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef int nvmlReturn_t;
typedef void *nvmlDevice_t;
#define MAXDEV 16
static char g_dev[MAXDEV][8];

static int ndev(void) { const char *e = getenv("MOCK_GPU_COUNT"); int n = e ? atoi(e) : 1; return n < 1 ? 1 : (n > MAXDEV ? MAXDEV : n); }
static uint64_t total(void) { const char *e = getenv("MOCK_TOTAL_MIB"); return (e ? strtoull(e, NULL, 0) : 16384ull) << 20; }
static int idx_of(nvmlDevice_t d) { for (int i = 0; i < MAXDEV; ++i) if (d == (nvmlDevice_t)g_dev[i]) return i; return 0; }
static void uuid_of(int i, char *out, unsigned len) { snprintf(out, len, "GPU-00000000-0000-0000-0000-%012d", i + 1); }

nvmlReturn_t nvmlInit(void) { return 0; }
nvmlReturn_t nvmlInit_v2(void) { return 0; }
nvmlReturn_t nvmlInitWithFlags(unsigned f) { (void)f; return 0; }
nvmlReturn_t nvmlShutdown(void) { return 0; }
const char *nvmlErrorString(nvmlReturn_t r) { (void)r; return "mock"; }
nvmlReturn_t nvmlDeviceGetCount(unsigned *n) { *n = (unsigned)ndev(); return 0; }
nvmlReturn_t nvmlDeviceGetCount_v2(unsigned *n) { *n = (unsigned)ndev(); return 0; }
nvmlReturn_t nvmlDeviceGetHandleByIndex(unsigned i, nvmlDevice_t *d) { if ((int)i >= ndev()) return 2; *d = g_dev[i]; return 0; }
nvmlReturn_t nvmlDeviceGetHandleByIndex_v2(unsigned i, nvmlDevice_t *d) { return nvmlDeviceGetHandleByIndex(i, d); }
nvmlReturn_t nvmlDeviceGetHandleByUUID(const char *uuid, nvmlDevice_t *d) {
    char u[96];
    for (int i = 0; i < ndev(); ++i) {
        uuid_of(i, u, sizeof(u));
        if (strcmp(u, uuid) == 0) { *d = g_dev[i]; return 0; }
    }
    return 6; /* NOT_FOUND */
}
nvmlReturn_t nvmlDeviceGetIndex(nvmlDevice_t d, unsigned *i) { *i = (unsigned)idx_of(d); return 0; }
nvmlReturn_t nvmlDeviceGetUUID(nvmlDevice_t d, char *uuid, unsigned len) { uuid_of(idx_of(d), uuid, len); return 0; }
nvmlReturn_t nvmlDeviceGetName(nvmlDevice_t d, char *name, unsigned len) { (void)d; snprintf(name, len, "Mock B200"); return 0; }
static uint64_t ctx_bytes(nvmlDevice_t d) {
    const char *e = getenv("MOCK_CTX_MIB");
    uint64_t mib = e ? strtoull(e, NULL, 0) : 0;
    if (!mib) return 0;
    int (*refs)(int) = (int (*)(int))dlsym(RTLD_DEFAULT, "mock_cuda_ctx_refs");
    return (refs && refs(idx_of(d)) > 0) ? mib << 20 : 0;
}
/* v1 layout {u32 pid; u64 usedGpuMemory} = 16 B; v2/v3 add {u32 gpuInstanceId, computeInstanceId} = 24 B */
static nvmlReturn_t procs(nvmlDevice_t d, unsigned *count, void *infos, size_t stride) {
    uint64_t b = ctx_bytes(d);
    if (!b) { *count = 0; return 0; }
    if (*count < 1 || !infos) { *count = 1; return 7; /* INSUFFICIENT_SIZE */ }
    memset(infos, 0, stride);
    *(unsigned *)infos = (unsigned)getpid();
    *(unsigned long long *)((char *)infos + 8) = b;
    *count = 1;
    return 0;
}
nvmlReturn_t nvmlDeviceGetComputeRunningProcesses(nvmlDevice_t d, unsigned *count, void *infos) { return procs(d, count, infos, 16); }
nvmlReturn_t nvmlDeviceGetComputeRunningProcesses_v2(nvmlDevice_t d, unsigned *count, void *infos) { return procs(d, count, infos, 24); }
nvmlReturn_t nvmlDeviceGetComputeRunningProcesses_v3(nvmlDevice_t d, unsigned *count, void *infos) { return procs(d, count, infos, 24); }
nvmlReturn_t nvmlDeviceGetProcessUtilization(nvmlDevice_t d, void *u, unsigned *count, unsigned long long ts) { (void)d; (void)u; (void)ts; *count = 0; return 0; }
typedef struct { unsigned long long total, free, used; } nvmlMemory_t;
nvmlReturn_t nvmlDeviceGetMemoryInfo(nvmlDevice_t d, nvmlMemory_t *m) { (void)d; m->total = total(); m->free = total(); m->used = 0; return 0; }
typedef struct { char busIdLegacy[16]; unsigned domain, bus, device, pciDeviceId, pciSubSystemId; char busId[32]; } nvmlPciInfo_t;
nvmlReturn_t nvmlDeviceGetPciInfo_v3(nvmlDevice_t d, nvmlPciInfo_t *p) {
    memset(p, 0, sizeof(*p));
    p->bus = 0x10u + (unsigned)idx_of(d);
    snprintf(p->busIdLegacy, sizeof(p->busIdLegacy), "0000:%02x:00.0", p->bus);
    snprintf(p->busId, sizeof(p->busId), "00000000:%02x:00.0", p->bus);
    return 0;
}
nvmlReturn_t nvmlDeviceGetPciInfo_v2(nvmlDevice_t d, nvmlPciInfo_t *p) { return nvmlDeviceGetPciInfo_v3(d, p); }
nvmlReturn_t nvmlDeviceGetPciInfo(nvmlDevice_t d, nvmlPciInfo_t *p) { return nvmlDeviceGetPciInfo_v3(d, p); }
