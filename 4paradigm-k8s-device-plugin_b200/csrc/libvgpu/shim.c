/*
 * shim.c — the new libvgpu.so: LD_PRELOAD interposer for the CUDA driver API and NVML.
 *
 * Same boundary as the reference's vgpu/libvgpu.so (Contract B): same hooked
 * symbol names, same environment variables, same CUresult quirks, same shared-cache file.
 * What is behind it differs: no launch throttle (spatial green-context SM partition instead,
 * see greenctx section), no getenv+atoi at every log site, no popen("ps ax") on OOM, and the
 * real dlsym is found with dlvsym (glibc >= 2.34 has no _dl_sym@GLIBC_PRIVATE, which is why the
 * reference binary needs a helper to even start on this image).
 *
 * Sections, in file order (every hook is also listed in g_hooks[] at the end, which is what dlsym and
 * cuGetProcAddress consult):
 *   real entry points, env, init      how the driver's own functions are found; limits from the environment;
 *                                     pre_init (before the real cuInit: NVML, cache file, CUDA_VISIBLE_DEVICES)
 *   oversubscribe backend             CUDA_OVERSUBSCRIBE route: managed memory (reference) or the VMM data path
 *   virtual devices                   NVIDIA_DEVICE_MAP as a device table: ordinals, cache-file columns, merged mode
 *   allocation list, hooks            the quota path proper (★ rows of SURVEY §8a): cuMemAlloc_v2, cuMemAllocManaged,
 *                                     cuMemAllocPitch_v2, cuMemFree_v2, cuMemGetInfo_v2, cuDeviceTotalMem_v2, host
 *                                     allocations, cuPointerGetAttributes
 *   context charge                    host pid probe, context billing (cuDevicePrimaryCtxRetain / Release, cuCtxCreate_v2)
 *   SM cap: green contexts            CUDA_DEVICE_SM_LIMIT as a spatial partition; cuCtx* hooks that swap it in;
 *                                     launch / copy / memset gates for suspend
 *   NVML                              count / handle narrowing, memory info with the quota
 *   virtual devices: entry points     every driver call that takes or returns a device ordinal
 *   strict quota                      opt-in metering of cuMemAllocAsync / cuMemCreate
 *   dispatch                          g_hooks[], versioned lookup, dlsym, cuGetProcAddress(_v2), suspend / resume
 * Calls not listed in g_hooks[] resolve straight to the real driver.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>
#include <unistd.h>

#include "region.h"
#include "shimlog.h"

#define EXPORT __attribute__((visibility("default")))

typedef int CUresult;
typedef int CUdevice;
typedef unsigned long long CUdeviceptr;
typedef void *CUcontext;

#define IPCSIZE 0x200000ull  /* .data@0x620c8: allocations above this take the oversubscribe route */

/* ------------------------------------------------------------------ logging level */
int shim_log_level(void) {
    static int lvl = -2;
    if (lvl == -2) {
        const char *e = getenv("LIBCUDA_LOG_LEVEL");
        lvl = e ? atoi(e) : -1;
    }
    return lvl;
}

/* ------------------------------------------------------------------ real entry points */
typedef void *(*dlsym_fn)(void *, const char *);
static dlsym_fn real_dlsym;
static void *g_cuda, *g_nvml;

static void find_real_dlsym(void) {
    if (real_dlsym) return;
    real_dlsym = (dlsym_fn)dlvsym(RTLD_NEXT, "dlsym", "GLIBC_2.2.5");
    if (!real_dlsym) real_dlsym = (dlsym_fn)dlvsym(RTLD_NEXT, "dlsym", "GLIBC_2.34");
    if (!real_dlsym) {
        SHIM_ERR("cannot find the real dlsym: %s", dlerror());
        abort();
    }
}

static void *cuda_sym(const char *name) {
    find_real_dlsym();
    if (!g_cuda) {
        g_cuda = dlopen("libcuda.so.1", RTLD_NOW | RTLD_NODELETE);  /* load_cuda_libraries@0x2e66c, flags 0x1002 */
        if (!g_cuda) {
            SHIM_ERR("can't dlopen libcuda.so.1: %s", dlerror());
            return NULL;
        }
    }
    void *p = real_dlsym(g_cuda, name);
    if (!p) SHIM_INFO("can't find function %s in libcuda.so.1", name);
    return p;
}

static void *nvml_sym(const char *name) {
    find_real_dlsym();
    if (!g_nvml) {
        g_nvml = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_NODELETE);
        if (!g_nvml) return NULL;
    }
    return real_dlsym(g_nvml, name);
}

#define REAL(ret, name, ...)                                   \
    typedef ret (*name##_t)(__VA_ARGS__);                      \
    static name##_t real_##name;                               \
    static inline name##_t get_##name(void) {                  \
        if (!real_##name) real_##name = (name##_t)cuda_sym(#name); \
        return real_##name;                                    \
    }

REAL(CUresult, cuInit, unsigned)
REAL(CUresult, cuDeviceGetCount, int *)
REAL(CUresult, cuCtxGetDevice, CUdevice *)
REAL(CUresult, cuCtxGetCurrent, CUcontext *)
REAL(CUresult, cuDeviceTotalMem_v2, size_t *, CUdevice)
REAL(CUresult, cuMemGetInfo_v2, size_t *, size_t *)
REAL(CUresult, cuMemAlloc_v2, CUdeviceptr *, size_t)
REAL(CUresult, cuMemAllocManaged, CUdeviceptr *, size_t, unsigned)
REAL(CUresult, cuMemAllocPitch_v2, CUdeviceptr *, size_t *, size_t, size_t, unsigned)
REAL(CUresult, cuMemFree_v2, CUdeviceptr)
REAL(CUresult, cuMemAllocHost_v2, void **, size_t)
REAL(CUresult, cuMemHostAlloc, void **, size_t, unsigned)
REAL(CUresult, cuMemFreeHost, void *)
REAL(CUresult, cuMemHostRegister_v2, void *, size_t, unsigned)
REAL(CUresult, cuMemHostUnregister, void *)
REAL(CUresult, cuDevicePrimaryCtxRetain, CUcontext *, CUdevice)
REAL(CUresult, cuDevicePrimaryCtxRelease_v2, CUdevice)
REAL(CUresult, cuCtxCreate_v2, CUcontext *, unsigned, CUdevice)
REAL(CUresult, cuDeviceGetUuid, unsigned char *, CUdevice)
REAL(CUresult, cuCtxSetCurrent, CUcontext)
REAL(CUresult, cuCtxPushCurrent_v2, CUcontext)
REAL(CUresult, cuLaunchKernel, void *, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, void *, void **, void **)
REAL(CUresult, cuGetProcAddress, const char *, void **, int, unsigned long long)
REAL(CUresult, cuGetProcAddress_v2, const char *, void **, int, unsigned long long, int *)

/* ------------------------------------------------------------------ env (a10, a11) */
/* get_limit_from_env@0x4186c: <n>[GgMmKk], strtoul base 0; unset/empty/0/overflow -> 0 */
static uint64_t limit_from_env(const char *name) {
    const char *v = getenv(name);
    if (!v || !*v) return 0;
    char *end = NULL;
    errno = 0;
    unsigned long n = strtoul(v, &end, 0);
    if (end == v) return 0;
    uint64_t scale = 1;
    switch (*end) {
        case 'G': case 'g': scale = 1ull << 30; break;
        case 'M': case 'm': scale = 1ull << 20; break;
        case 'K': case 'k': scale = 1ull << 10; break;
        default: break;
    }
    uint64_t res = (uint64_t)n * scale;
    if (n != 0 && res / scale != n) return 0;  /* overflow check, same as the reference */
    if (res == 0) {
        if (name[12] == 'S') SHIM_INFO("device core util limit set to 0, which means no limit: %s=%s", name, v);
        else if (name[12] == 'M') SHIM_WARN("invalid device memory limit %s=%s", name, v);
    }
    return res;
}

static void init_limits(vgpu_region_init_t *init) {
    /* do_init_device_memory_limits@0x4232e / do_init_device_sm_limits@0x4246e */
    uint64_t mem_all = limit_from_env("CUDA_DEVICE_MEMORY_LIMIT");
    uint64_t sm_all = limit_from_env("CUDA_DEVICE_SM_LIMIT");
    if (sm_all == 0) sm_all = 100;
    for (int i = 0; i < VGPU_MAX_DEVICES; ++i) {
        char key[64];
        snprintf(key, sizeof(key), "CUDA_DEVICE_MEMORY_LIMIT_%d", i);
        uint64_t v = limit_from_env(key);
        init->limit[i] = v ? v : mem_all;
        snprintf(key, sizeof(key), "CUDA_DEVICE_SM_LIMIT_%d", i);
        v = limit_from_env(key);
        init->sm_limit[i] = v ? v : sm_all;
    }
}

/* ------------------------------------------------------------------ init (a26) */
static pthread_once_t g_pre_once = PTHREAD_ONCE_INIT, g_post_once = PTHREAD_ONCE_INIT;
static int g_virtual_devices = -1;  /* entries of NVIDIA_DEVICE_MAP, else the real device count */
static int g_allocmode = 2;         /* 0: oversubscribe route, 2: plain cuMemAlloc (postInit@0x16ded) */
static int g_map_has_duplicates;    /* duplicate_devices: a GPU named twice in NVIDIA_DEVICE_MAP */
static int g_strict;                /* VGPU_STRICT_QUOTA=1: also meter the allocation paths the reference lets through */
static int g_region_ok = 0;

static int count_device_map(void) {
    const char *m = getenv("NVIDIA_DEVICE_MAP");
    if (!m || !*m) return -1;
    int n = 0;
    const char *p = m;
    while (*p) {
        while (*p == ' ') ++p;
        if (!*p) break;
        ++n;
        while (*p && *p != ' ') ++p;
    }
    return n;
}

static int device_map_uuids(char out[][VGPU_UUID_LEN], int max);

/* GPUs the container can see but was not given: let the DRIVER hide them (it reads CUDA_VISIBLE_DEVICES at cuInit,
 * which has not run yet), so that the private driver paths of the CUDA runtime agree with what the device hooks
 * answer — with the hooks alone the runtime fails to start ("Error 103: integrity checks failed", seen on a 2-GPU
 * box).  A CUDA_VISIBLE_DEVICES the user (or the container runtime) already set is read the way it is meant inside a
 * vGPU container — ordinals count the container's OWN devices, i.e. the map's — and rewritten as UUIDs; as with the
 * driver, the list ends at the first entry that names nothing.  (Reference: parse_cuda_visible_env@0x18b15.) */
static void list_append(char *out, size_t cap, const char *s) {
    const size_t l = strlen(out);
    if (l + strlen(s) + 2 > cap) return;
    if (l) strcat(out, ",");
    strcat(out, s);
}
static void narrow_visible_devices(char vis[][VGPU_UUID_LEN], int nvis, int nreal) {
    if (nvis <= 0) return;
    const char *user = getenv("CUDA_VISIBLE_DEVICES");
    char out[VGPU_MAX_DEVICES * VGPU_UUID_LEN] = "";
    if (!user) {
        if (nvis >= nreal) return;  /* nothing to hide */
        for (int i = 0; i < nvis; ++i) list_append(out, sizeof(out), vis[i]);
    } else {
        if (strstr(user, "MIG-")) return;  /* MIG instances are not vGPUs of this plugin: left alone */
        int used[VGPU_MAX_DEVICES] = {0}, n = 0;
        const char *p = user;
        while (*p && n < VGPU_MAX_DEVICES) {
            const char *end = strchr(p, ',');
            size_t len = end ? (size_t)(end - p) : strlen(p);
            int pick = -1;
            if (len > 0 && len < 8 && strspn(p, "0123456789") == len) {
                int idx = atoi(p);
                if (idx < nvis) pick = idx;
            } else if (len >= 5 && strncasecmp(p, "GPU-", 4) == 0) {
                for (int i = 0; i < nvis && pick < 0; ++i)
                    if (strncasecmp(vis[i], p, len) == 0) pick = i;  /* the driver accepts unique prefixes */
            }
            if (pick < 0 || used[pick]) break;
            used[pick] = 1;
            list_append(out, sizeof(out), vis[pick]);
            ++n;
            if (!end) break;
            p = end + 1;
        }
        if (strcmp(out, user) == 0) return;
    }
    /* The spill tier's peer GPU (VGPU_OVERSUB_PEER=<UUID>) must stay open to the DRIVER of this process — P2P needs
     * both ends — so it rides behind the vGPU's own GPUs; the virtual-device table keeps it away from the application
     * (build_vdev). */
    const char *peer = getenv("VGPU_OVERSUB_PEER");
    if (peer && strncasecmp(peer, "GPU-", 4) == 0 && !strcasestr(out, peer)) list_append(out, sizeof(out), peer);
    setenv("CUDA_VISIBLE_DEVICES", out, 1);
    SHIM_INFO("CUDA_VISIBLE_DEVICES=%s (from NVIDIA_DEVICE_MAP%s%s)", out, user ? " and " : "", user ? user : "");
}

/* load_env_from_file@0x420cc (the reference calls it with "/overrideEnv" from its NVML pre-init): KEY=VALUE lines, each
 * put into the environment over whatever is there (setenv(.., 1)); the value is everything after the FIRST '=', the
 * trailing newline goes; reading stops at the first line that has no '='.  Always answers 0. */
EXPORT int load_env_from_file(const char *path) {
    FILE *f = path ? fopen(path, "r") : NULL;
    if (!f) return 0;
    char line[10000];
    while (fgets(line, sizeof(line), f)) {
        char *eq = strchr(line, '=');
        if (!eq) break;
        size_t n = strlen(line);
        if (n && line[n - 1] == '\n') line[n - 1] = 0;
        *eq = 0;
        setenv(line, eq + 1, 1);
    }
    fclose(f);
    return 0;
}

static void pre_init(void) {
    vgpu_region_init_t init;
    memset(&init, 0, sizeof(init));
    load_env_from_file("/overrideEnv");  /* an admin-provided file in the container wins over the pod's environment */
    init_limits(&init);
    const char *prio = getenv("CUDA_TASK_PRIORITY");
    init.priority = prio ? atoi(prio) : 1;

    int real_count = 0;
    if (get_cuDeviceGetCount()) {
        /* cuDeviceGetCount before cuInit fails on a real driver; retried in post_init */
        if (real_cuDeviceGetCount(&real_count) != 0) real_count = 0;
    }
    int mapped = count_device_map();
    g_virtual_devices = mapped > 0 ? mapped : real_count;

    /* put_device_info@0x41e2e: UUIDs through NVML */
    typedef int (*nv_init_t)(void);
    typedef int (*nv_count_t)(unsigned *);
    typedef int (*nv_byidx_t)(unsigned, void **);
    typedef int (*nv_uuid_t)(void *, char *, unsigned);
    nv_init_t nv_init = (nv_init_t)nvml_sym("nvmlInit_v2");
    if (!nv_init) nv_init = (nv_init_t)nvml_sym("nvmlInit");
    nv_byidx_t nv_byidx = (nv_byidx_t)nvml_sym("nvmlDeviceGetHandleByIndex_v2");
    if (!nv_byidx) nv_byidx = (nv_byidx_t)nvml_sym("nvmlDeviceGetHandleByIndex");
    nv_uuid_t nv_uuid = (nv_uuid_t)nvml_sym("nvmlDeviceGetUUID");
    nv_count_t nv_count = (nv_count_t)nvml_sym("nvmlDeviceGetCount_v2");
    if (nv_init && nv_byidx && nv_uuid && nv_init() == 0) {
        unsigned n = 0;
        if (g_virtual_devices <= 0 && nv_count && nv_count(&n) == 0) g_virtual_devices = (int)n;
        char want[VGPU_MAX_DEVICES][VGPU_UUID_LEN];
        const int nwant = device_map_uuids(want, VGPU_MAX_DEVICES);
        for (int i = 0; i < nwant; ++i)
            for (int k = 0; k < i; ++k) g_map_has_duplicates |= strcasecmp(want[i], want[k]) == 0;
        int nreal = g_virtual_devices, nvisible = 0;
        char visible[VGPU_MAX_DEVICES][VGPU_UUID_LEN];  /* the distinct GPUs the map names, physical order */
        if (nwant > 0 && nv_count && nv_count(&n) == 0) nreal = (int)n;
        for (int i = 0, v = 0; i < nreal && i < VGPU_MAX_DEVICES && v < VGPU_MAX_DEVICES; ++i) {
            void *h = NULL;
            char u[VGPU_UUID_LEN] = {0};  /* copied whole into the cache file: no stack bytes after the terminator */
            if (nv_byidx((unsigned)i, &h) != 0 || nv_uuid(h, u, VGPU_UUID_LEN) != 0) break;
            if (nwant == 0) {
                memcpy(init.uuids[v++], u, VGPU_UUID_LEN);
                continue;
            }
            int named = 0;
            for (int k = 0; k < nwant && v < VGPU_MAX_DEVICES; ++k)  /* one line per virtual device, physical order */
                if (strcasecmp(want[k], u) == 0) {
                    memcpy(init.uuids[v++], u, VGPU_UUID_LEN);
                    ++named;
                }
            if (named && nvisible < VGPU_MAX_DEVICES) memcpy(visible[nvisible++], u, VGPU_UUID_LEN);
        }
        narrow_visible_devices(visible, nvisible, nreal);
    }
    init.device_num = g_virtual_devices > 0 ? (uint64_t)g_virtual_devices : 0;

    const char *path = getenv("CUDA_DEVICE_MEMORY_SHARED_CACHE");
    if (!path) path = "/tmp/cudevshr.cache";  /* .rodata@0x56ac0 */
    g_region_ok = region_attach(path, &init) == 0;
    if (!g_region_ok) {
        /* Fail closed.  The path is the tenant's to tamper with (an environment variable); carrying on without the
         * file would mean carrying on without a quota.  The reference logs the same errors and then crashes in its
         * init (multiprocess_memory_limit.c:680-697, segfault); here the process ends with a message. */
        SHIM_ERR("shared cache %s unavailable: refusing to run CUDA without the vGPU quota", path);
        exit(-1);
    }
}

static void load_vmm_backend(void);
static void start_mover(void);

static void post_init(void) {
    const char *o = getenv("CUDA_OVERSUBSCRIBE");
    g_allocmode = (o && strcmp(o, "true") == 0) ? 0 : 2;
    const char *st = getenv("VGPU_STRICT_QUOTA");
    g_strict = st && *st && strcmp(st, "0") != 0;
    if (g_allocmode == 0) load_vmm_backend();
    if (g_virtual_devices <= 0 && get_cuDeviceGetCount()) {
        int n = 0;
        if (real_cuDeviceGetCount(&n) == 0) g_virtual_devices = n;
    }
}

static void ensure_initialized(void) { pthread_once(&g_pre_once, pre_init); }

/* suspend gate: wait_status_self@0x45ff5 — wait while an external monitor has us at status 2,
 * and (vmm backend) until the data that was spilled on suspend is back behind its pointers */
static volatile int g_data_out;  /* 1 while VMM-backed allocations are (being) spilled */
static void wait_running(void) {
    while (g_region_ok && (region_my_status() != 1 || g_data_out)) usleep(g_data_out ? 1000 : 1000000);
}
/* A real gate for the calls that touch device memory or submit work: a caller announces itself (in-flight count)
 * AFTER it has seen the gate open and re-checks; the mover closes the gate first and then waits for the count to
 * reach zero, so nobody can be between "passed the gate" and "called the driver" when the HBM is unmapped. */
static volatile int g_inflight;
static void gate_enter(void) {
    for (;;) {
        wait_running();
        __atomic_add_fetch(&g_inflight, 1, __ATOMIC_SEQ_CST);
        if (!g_data_out) return;
        __atomic_sub_fetch(&g_inflight, 1, __ATOMIC_SEQ_CST);
    }
}
static void gate_exit(void) { __atomic_sub_fetch(&g_inflight, 1, __ATOMIC_SEQ_CST); }

/* ------------------------------------------------------------------ oversubscribe backend
 * CUDA_OVERSUBSCRIBE=true routes allocations above 2 MiB away from plain cuMemAlloc.  The
 * reference hands them to UVM (cuMemAllocManaged).  With VGPU_OVERSUB_BACKEND=vmm they are
 * backed by the VMM manager of libvgpu_dp.so instead: the pointer the application holds stays
 * valid while the chunks behind it are spilled to pinned host RAM by the spill kernel (suspend,
 * SIGUSR2 / cuVGPUSuspend) and restored + checksum-verified by the restore kernel (resume,
 * SIGUSR1 / cuVGPUResume), releasing the HBM to other vGPUs in between.
 * While the shim itself calls into libvgpu_dp.so, that library's own driver calls must not be
 * metered or re-enter the allocator lock: tl_bypass makes every hook a plain pass-through. */
static __thread int tl_bypass;
static struct {
    void *lib;
    int (*create)(int, uint64_t, uint64_t, int, int, uint64_t, void **);
    int (*create_tiered)(int, uint64_t, uint64_t, int, uint64_t, uint64_t, void **);
    int (*base)(void *, void **);
    int (*ensure)(void *, uint64_t, uint64_t);
    int (*evict)(void *, uint64_t, uint64_t);
    int (*missing)(void *, uint64_t, uint64_t, uint64_t *);
    int (*swap)(void *, void *, uint64_t, uint64_t);  /* optional: a data-path library without it evicts, then ensures */
    int (*destroy)(void *);
    const char *(*last_error)(void);
} g_dp;
static int g_backend_vmm;

static void load_vmm_backend(void) {
    const char *b = getenv("VGPU_OVERSUB_BACKEND");
    if (!b || strcmp(b, "vmm") != 0) return;
    find_real_dlsym();
    Dl_info info;
    char path[4096] = "libvgpu_dp.so";
    if (dladdr((void *)load_vmm_backend, &info) && info.dli_fname) {  /* next to libvgpu.so */
        const char *slash = strrchr(info.dli_fname, '/');
        if (slash) snprintf(path, sizeof(path), "%.*s/libvgpu_dp.so", (int)(slash - info.dli_fname), info.dli_fname);
    }
    tl_bypass = 1;
    g_dp.lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    tl_bypass = 0;
    if (!g_dp.lib) {
        SHIM_ERR("VGPU_OVERSUB_BACKEND=vmm but %s cannot be loaded (%s): staying with managed memory", path, dlerror());
        return;
    }
    g_dp.create = (int (*)(int, uint64_t, uint64_t, int, int, uint64_t, void **))real_dlsym(g_dp.lib, "vgpu_vmem_create");
    g_dp.create_tiered = (int (*)(int, uint64_t, uint64_t, int, uint64_t, uint64_t, void **))real_dlsym(g_dp.lib, "vgpu_vmem_create_tiered");
    g_dp.base = (int (*)(void *, void **))real_dlsym(g_dp.lib, "vgpu_vmem_base");
    g_dp.ensure = (int (*)(void *, uint64_t, uint64_t))real_dlsym(g_dp.lib, "vgpu_vmem_ensure");
    g_dp.evict = (int (*)(void *, uint64_t, uint64_t))real_dlsym(g_dp.lib, "vgpu_vmem_evict");
    g_dp.missing = (int (*)(void *, uint64_t, uint64_t, uint64_t *))real_dlsym(g_dp.lib, "vgpu_vmem_missing");
    g_dp.swap = getenv("VGPU_OVERSUB_NO_SWAP") ? NULL : (int (*)(void *, void *, uint64_t, uint64_t))real_dlsym(g_dp.lib, "vgpu_vmem_swap");
    g_dp.destroy = (int (*)(void *))real_dlsym(g_dp.lib, "vgpu_vmem_destroy");
    g_dp.last_error = (const char *(*)(void))real_dlsym(g_dp.lib, "vgpu_last_error");
    g_backend_vmm = g_dp.create && g_dp.create_tiered && g_dp.base && g_dp.ensure && g_dp.evict && g_dp.missing && g_dp.destroy && g_dp.last_error;
    if (!g_backend_vmm) SHIM_ERR("%s lacks the vgpu_vmem_* entry points", path);
    else start_mover();
}

/* ------------------------------------------------------------------ virtual devices (a28)
 * NVIDIA_DEVICE_MAP="0:<UUID> 1:<UUID> ..." (Contract A, server.go:487-496) names the vGPUs of this
 * container.  Two of them may sit on the same physical GPU (a pod that asked for two vGPUs usually gets
 * neighbours of one split), and the container may see GPUs it was not given.  Observed from the
 * reference (nvml/util.c parser@0x28a39, initial_virtual_devices@0x27f0d, cuda/device.c, cuda/context.c)
 * under the mock driver, vectors in tests/golden/quota_remap_*.json.gz:
 *   - the application sees one CUDA device per map entry, enumerated in physical order whatever the
 *     order of the map; unlisted GPUs disappear;
 *   - the 2nd, 3rd ... entry of one physical GPU is a device of its own: same UUID and attributes, PCI
 *     function .1, .2 ... in its bus id, its own CUDA_DEVICE_MEMORY_LIMIT_<i> / SM limit / usage
 *     column, and — because a physical device has one primary context — a context of its own, made
 *     with cuCtxCreate the first time its "primary context" is retained (cuDevicePrimaryCtxRetain
 *     @0x297be) and billed like any first context.
 * CUdevice values handed to the application are the virtual ordinals; every entry point below that
 * takes or returns one translates.  With no map, or a map that is the identity, nothing is translated.
 * Not reproduced: the reference's rewritten UUID for the duplicate, its crash on an out-of-range
 * ordinal (here CUDA_ERROR_INVALID_DEVICE), and its virtual NVML handles (here NVML keeps physical
 * devices; memory info of a shared GPU is the sum of its vGPUs). */
typedef struct {
    int phys;       /* real CUDA ordinal */
    int dup;        /* 0 = first vGPU of that GPU (owns the primary context), k = k-th further one */
    CUcontext ctx;  /* dup > 0: the context standing in for a primary context */
    int refs;
    int col0, ncols; /* its columns (limit[], sm_limit[], used[]) of the cache file: one, or — merged — all of that GPU's */
    int spill_only;  /* the peer GPU of the spill tier where it cannot be hidden (see build_vdev): a device with no memory to give */
} vdev_t;
static vdev_t g_vdev[VGPU_MAX_DEVICES];
static int g_nvdev;
static int g_ncols;  /* map entries that found their GPU = columns of the cache file in use */
static int g_remap;  /* 0: identity, nothing to translate */
static int g_vdev_built;

/* The CUDA runtime cannot live with two devices on one GPU: besides the public API it asks the driver about
 * devices through private tables (cuGetExportTable) that no interposer sees, and fails to initialise when those
 * disagree with the public answers (seen on CUDA 12.9: "CUDA driver initialization failed").  So a process that
 * fetches an export table before it enumerates devices — every runtime-API program does, first thing after cuInit —
 * gets the vGPUs of one GPU MERGED into one device that owns all their columns: limit = the sum, usage = the sum,
 * SM share = the sum.  Pure driver-API programs get them split, as the reference does.
 * VGPU_SHARED_GPU=split|merge overrides the detection. */
static int g_runtime_client;
static int merge_shared(void) {
    const char *e = getenv("VGPU_SHARED_GPU");
    if (e && strcmp(e, "split") == 0) return 0;
    if (e && strcmp(e, "merge") == 0) return 1;
    return g_runtime_client;
}

static void uuid_text(const unsigned char *u, char *s, size_t n) {
    snprintf(s, n, "GPU-%02x%02x%02x%02x-%02x%02x-%02x%02x-%02x%02x-%02x%02x%02x%02x%02x%02x", u[0], u[1], u[2], u[3], u[4],
             u[5], u[6], u[7], u[8], u[9], u[10], u[11], u[12], u[13], u[14], u[15]);
}

/* the UUIDs of NVIDIA_DEVICE_MAP in map order; -> count */
static int device_map_uuids(char out[][VGPU_UUID_LEN], int max) {
    const char *m = getenv("NVIDIA_DEVICE_MAP");
    int n = 0;
    while (m && *m && n < max) {
        while (*m == ' ') ++m;
        if (!*m) break;
        const char *end = m;
        while (*end && *end != ' ') ++end;
        const char *colon = memchr(m, ':', (size_t)(end - m));
        const char *u = colon ? colon + 1 : m;
        size_t len = (size_t)(end - u);
        if (len > 0 && len < VGPU_UUID_LEN) {
            memcpy(out[n], u, len);
            out[n][len] = 0;
            ++n;
        }
        m = end;
    }
    return n;
}

static int build_vdev(void) {  /* needs a successful cuInit: -> 0 (and nothing latched) before that */
    int real = 0;
    if (!get_cuDeviceGetCount() || real_cuDeviceGetCount(&real) != 0) return 0;
    if (real > VGPU_MAX_DEVICES) real = VGPU_MAX_DEVICES;
    char want[VGPU_MAX_DEVICES][VGPU_UUID_LEN];
    int nwant = device_map_uuids(want, VGPU_MAX_DEVICES);
    const int merge = merge_shared();
    g_nvdev = g_ncols = 0;
    for (int p = 0; p < real && nwant > 0; ++p) {  /* physical order, whatever the order of the map */
        unsigned char u[16];
        char text[64];
        if (!get_cuDeviceGetUuid() || real_cuDeviceGetUuid(u, p) != 0) continue;
        uuid_text(u, text, sizeof(text));
        int k = 0;
        for (int i = 0; i < nwant && g_ncols < VGPU_MAX_DEVICES; ++i) {
            if (strcasecmp(want[i], text) != 0) continue;
            if (merge && k > 0) ++g_vdev[g_nvdev - 1].ncols;
            else g_vdev[g_nvdev++] = (vdev_t){p, k, NULL, 0, g_ncols, 1, 0};
            ++k;
            ++g_ncols;
        }
    }
    /* The spill tier's peer GPU is open to the driver but is not one of the container's vGPUs.  A driver-API program
     * simply does not get it (it is not in the table).  The CUDA runtime cannot be lied to about the device count (it
     * cross-checks private driver tables, see merge_shared), so a runtime client sees it as one more device at the end
     * with a quota of one byte: nothing can be allocated there, and NVML — which frameworks count GPUs with — still
     * shows the vGPUs only. */
    const char *peer = getenv("VGPU_OVERSUB_PEER");
    if (g_nvdev > 0 && merge && peer && strncasecmp(peer, "GPU-", 4) == 0)
        for (int p = 0; p < real && g_nvdev < VGPU_MAX_DEVICES; ++p) {
            unsigned char u[16];
            char text[64];
            if (!get_cuDeviceGetUuid() || real_cuDeviceGetUuid(u, p) != 0) continue;
            uuid_text(u, text, sizeof(text));
            int named = strcasecmp(text, peer) != 0;
            for (int i = 0; i < nwant; ++i) named |= strcasecmp(want[i], text) == 0;
            if (!named) g_vdev[g_nvdev++] = (vdev_t){p, 0, NULL, 0, VGPU_MAX_DEVICES - 1, 0, 1};
        }
    if (g_nvdev == 0) {
        if (nwant > 0) SHIM_ERR("NVIDIA_DEVICE_MAP names no GPU this container can see: devices are not remapped");
        for (int p = 0; p < real; ++p) g_vdev[g_nvdev++] = (vdev_t){p, 0, NULL, 0, p, 1, 0};
        g_ncols = nwant > 0 ? 0 : real;  /* a map that matches nothing: the column count differs from the map's, quota off */
    }
    g_remap = g_nvdev != real;
    for (int i = 0; i < g_nvdev; ++i) g_remap |= g_vdev[i].phys != i || g_vdev[i].col0 != i || g_vdev[i].ncols != 1;
    if (g_remap)
        for (int i = 0; i < g_nvdev; ++i)
            SHIM_INFO("virtual device %d -> GPU %d%s, columns %d..%d", i, g_vdev[i].phys, g_vdev[i].dup ? " (shared)" : "", g_vdev[i].col0,
                      g_vdev[i].col0 + g_vdev[i].ncols - 1);
    return 1;
}
/* built once, by the first caller that comes after cuInit (an NVML-only call before that must not freeze an empty table) */
static void ensure_vdev(void) {
    static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    if (__atomic_load_n(&g_vdev_built, __ATOMIC_ACQUIRE)) return;
    pthread_mutex_lock(&mu);
    if (!g_vdev_built && build_vdev()) __atomic_store_n(&g_vdev_built, 1, __ATOMIC_RELEASE);
    pthread_mutex_unlock(&mu);
}

/* virtual ordinal -> real ordinal, in place; 0 when there is no such device */
static int to_phys(CUdevice *d) {
    ensure_vdev();
    if (!g_remap) return 1;
    if (*d < 0 || *d >= g_nvdev) return 0;
    *d = g_vdev[*d].phys;
    return 1;
}
static int phys_of(int vdev) { return (g_remap && vdev >= 0 && vdev < g_nvdev) ? g_vdev[vdev].phys : vdev; }
static int first_vdev_of(int phys) {
    if (!g_remap) return phys;
    for (int i = 0; i < g_nvdev; ++i)
        if (g_vdev[i].phys == phys) return i;
    return -1;
}

/* limit / usage / SM share of a virtual device = those of its columns */
static int col_of(int vdev) { return (g_remap && vdev >= 0 && vdev < g_nvdev) ? g_vdev[vdev].col0 : vdev; }
static uint64_t q_limit(int vdev) {
    if (!g_remap || vdev < 0 || vdev >= g_nvdev) return region_limit(vdev);
    if (g_vdev[vdev].spill_only) return 1;  /* one byte: every allocation is over quota */
    uint64_t t = 0;
    for (int c = g_vdev[vdev].col0; c < g_vdev[vdev].col0 + g_vdev[vdev].ncols; ++c) {
        if (region_limit(c) == 0) return 0;  /* one unlimited vGPU makes the merged device unlimited */
        t += region_limit(c);
    }
    return t;
}
static uint64_t q_usage(int vdev) {
    if (!g_remap || vdev < 0 || vdev >= g_nvdev) return region_usage(vdev);
    uint64_t t = 0;
    for (int c = g_vdev[vdev].col0; c < g_vdev[vdev].col0 + g_vdev[vdev].ncols; ++c) t += region_usage(c);
    return t;
}
static uint64_t q_sm_limit(int vdev) {
    if (!g_remap || vdev < 0 || vdev >= g_nvdev) return region_sm_limit(vdev);
    if (g_vdev[vdev].spill_only) return 100;
    uint64_t t = 0;
    for (int c = g_vdev[vdev].col0; c < g_vdev[vdev].col0 + g_vdev[vdev].ncols; ++c) {
        uint64_t p = region_sm_limit(c);
        if (p == 0 || p >= 100) return 100;
        t += p;
    }
    return t > 100 ? 100 : t;
}

/* ------------------------------------------------------------------ allocation list (a13, a15) */
typedef struct alloc_node {
    CUdeviceptr addr;
    size_t len;
    int dev;
    void *vmem;  /* VMM-backed (oversubscribe backend) */
    uint64_t last_use;   /* residency clock (paged mode) */
    uint64_t epoch;      /* the launch / copy that needs it right now: not a victim */
    int windowed;        /* larger than the GPU can hold at once: residency by prefetch hints only */
    int raw;             /* made by cuMemoryAllocate (the route under the quota): paged like the others, never billed */
    uint64_t peer_bytes; /* its share of the VGPU_OVERSUB_PEER_MB budget, given back when it is freed */
    int prefer;          /* cuMemAdvise: +1 preferred location = the GPU (spilled last), -1 = the CPU (spilled first), 0 none */
    struct alloc_node *next, *prev;
} alloc_node_t;
static alloc_node_t *g_head, *g_tail;
static size_t g_nalloc;
static pthread_mutex_t g_alloc_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_mutex_t g_res_mu = PTHREAD_MUTEX_INITIALIZER;  /* residency of VMM-backed allocations (taken before g_alloc_mu) */
static uint64_t g_peer_left = ~0ull;  /* VGPU_OVERSUB_PEER_MB: what this process may still put on the peer (under g_res_mu) */
static alloc_node_t **g_idx;           /* the managed ranges sorted by address (vmm_find's binary search), under g_alloc_mu */
static size_t g_idx_n, g_idx_cap;
static void idx_del(alloc_node_t *a);
static void vmm_quiesce(void);
static int vmm_make_room(void);
static void vmm_keep_reserve(void);

static void track_vmem(CUdeviceptr p, size_t n, int dev, void *vmem) {
    alloc_node_t *a = (alloc_node_t *)malloc(sizeof(*a));
    a->addr = p;
    a->len = n;
    a->dev = dev;
    a->vmem = vmem;
    a->last_use = a->epoch = 0;
    a->windowed = a->raw = 0;
    a->peer_bytes = 0;
    a->prefer = 0;
    a->next = NULL;
    a->prev = g_tail;
    if (g_tail) g_tail->next = a; else g_head = a;
    g_tail = a;
    ++g_nalloc;
}

static int vdev_of_ctx(CUcontext ctx, int phys);
/* the (virtual) device of the current context */
static int cur_dev(int *dev) {
    cuCtxGetDevice_t f = get_cuCtxGetDevice();
    if (!f) return 3;
    int rc = f(dev);
    if (rc != 0) return rc;
    ensure_vdev();
    if (g_remap) {
        CUcontext ctx = NULL;
        if (get_cuCtxGetCurrent()) real_cuCtxGetCurrent(&ctx);
        int v = vdev_of_ctx(ctx, *dev);
        if (v < 0) return 101;  /* a context on a GPU that is not ours */
        *dev = v;
    }
    return 0;
}

static void track(CUdeviceptr p, size_t n, int dev) { track_vmem(p, n, dev, NULL); }

/* oom_check@0x40528 */
static int oom_check(int dev, size_t addon) {
    if (!g_region_ok) return 0;
    ensure_vdev();
    if (!g_vdev_built || g_ncols != g_virtual_devices) return 0;  /* every map entry must have found its GPU (the reference compares device counts, @0x40567) */
    if (dev == -1 && cur_dev(&dev) != 0) return 0;
    for (;;) {
        uint64_t limit = q_limit(dev);
        if (limit == 0) return 0;
        uint64_t usage = q_usage(dev);
        if (usage + addon > limit) {  /* u64 add, no overflow guard, as in the reference */
            SHIM_ERR("Device %d OOM %lu / %lu", dev, (unsigned long)(usage + addon), (unsigned long)limit);
            if (region_reap_dead() > 0) continue;  /* rm_quitted_process: retry once per reaped process */
            return 1;
        }
        return 0;
    }
}

static int vmm_alloc(int dev, size_t n, CUdeviceptr *p);  /* residency section, below */

/* add_chunk@0x40b73 (route = 0) and add_chunk_only@0x41011 (route = 1: caller already allocated) */
static int add_chunk(CUdeviceptr *p, size_t n) {
    int dev = 0;
    CUresult rc = cur_dev(&dev);
    if (rc != 0) return rc;
    if (oom_check(dev, n)) return -1;
    if (n <= IPCSIZE || g_allocmode == 2) {
        cuMemAlloc_v2_t f = get_cuMemAlloc_v2();
        rc = f ? f(p, n) : 3;
        /* VMM backend: small allocations live outside the managed ranges, in the HBM the residency engine keeps free
         * (VGPU_OVERSUB_RESERVE_MB).  A burst of them can use that up between two launches: spill and try again. */
        while (rc == 2 && g_backend_vmm && vmm_make_room()) rc = f(p, n);
        if (rc == 0 && g_backend_vmm) vmm_keep_reserve();
    } else if (g_backend_vmm) {
        if (vmm_alloc(dev, n, p) != 0) return 2;
        region_add_usage(col_of(dev), n, VGPU_USE_DATA);
        return 0;
    } else {
        /* cuMemoryAllocate@0x32146, allocmode 0.  The reference hands the allocation to
         * UVM here; the explicit spill/restore data path (libvgpu_dp.so) takes this route
         * over when the VMM manager is attached (vmem section). */
        cuMemAllocManaged_t f = get_cuMemAllocManaged();
        rc = f ? f(p, n, 1 /* CU_MEM_ATTACH_GLOBAL */) : 3;
    }
    if (rc != 0) return rc;
    track(*p, n, dev);
    region_add_usage(col_of(dev), n, VGPU_USE_DATA);
    return 0;
}

static int add_chunk_only(CUdeviceptr p, size_t n) {
    int dev = 0;
    CUresult rc = cur_dev(&dev);
    if (rc != 0) return rc;
    track(p, n, dev);
    region_add_usage(col_of(dev), n, VGPU_USE_DATA);
    return 0;
}

/* ------------------------------------------------------------------ hooks */
EXPORT CUresult cuInit(unsigned flags) {
    ensure_initialized();
    cuInit_t f = get_cuInit();
    CUresult rc = f ? f(flags) : 3;
    if (rc == 0) pthread_once(&g_post_once, post_init);
    return rc;
}

EXPORT CUresult cuMemAlloc_v2(CUdeviceptr *dptr, size_t bytesize) {
    if (tl_bypass) return get_cuMemAlloc_v2()(dptr, bytesize);
    ensure_initialized();
    pthread_once(&g_post_once, post_init);
    gate_enter();
    if (g_backend_vmm) pthread_mutex_lock(&g_res_mu);  /* lock order: residency, then the allocation list */
    pthread_mutex_lock(&g_alloc_mu);
    int rc = add_chunk(dptr, bytesize);  /* allocate_raw@0x4157c: -1 on quota OOM, returned verbatim */
    pthread_mutex_unlock(&g_alloc_mu);
    if (g_backend_vmm) pthread_mutex_unlock(&g_res_mu);
    gate_exit();
    return rc;
}
EXPORT CUresult cuMemAlloc(CUdeviceptr *dptr, size_t bytesize) { return cuMemAlloc_v2(dptr, bytesize); }

EXPORT CUresult cuMemAllocManaged(CUdeviceptr *dptr, size_t bytesize, unsigned flags) {
    if (tl_bypass) return get_cuMemAllocManaged()(dptr, bytesize, flags);
    ensure_initialized();
    pthread_once(&g_post_once, post_init);
    wait_running();
    if (oom_check(-1, bytesize)) return 2;  /* CUDA_ERROR_OUT_OF_MEMORY, @0x3281d */
    cuMemAllocManaged_t f = get_cuMemAllocManaged();
    CUresult rc = f ? f(dptr, bytesize, flags) : 3;
    if (rc != 0) return rc;
    pthread_mutex_lock(&g_alloc_mu);
    add_chunk_only(*dptr, bytesize);
    pthread_mutex_unlock(&g_alloc_mu);
    return 0;
}

EXPORT CUresult cuMemAllocPitch_v2(CUdeviceptr *dptr, size_t *pitch, size_t w, size_t h, unsigned elem) {
    if (tl_bypass) return get_cuMemAllocPitch_v2()(dptr, pitch, w, h, elem);
    ensure_initialized();
    pthread_once(&g_post_once, post_init);
    wait_running();
    /* bytes = ((W-1)/Elem+1)*Elem * H, @0x32bd7-0x32c02 */
    size_t bytes = elem ? (((w - 1) / elem + 1) * elem) * h : w * h;
    if (oom_check(-1, bytes)) return 2;
    cuMemAllocPitch_v2_t f = get_cuMemAllocPitch_v2();
    CUresult rc = f ? f(dptr, pitch, w, h, elem) : 3;
    if (rc != 0) return rc;
    pthread_mutex_lock(&g_alloc_mu);
    add_chunk_only(*dptr, bytes);
    pthread_mutex_unlock(&g_alloc_mu);
    return 0;
}
EXPORT CUresult cuMemAllocPitch(CUdeviceptr *dptr, size_t *pitch, size_t w, size_t h, unsigned elem) {
    return cuMemAllocPitch_v2(dptr, pitch, w, h, elem);
}

EXPORT CUresult cuMemFree_v2(CUdeviceptr dptr) {
    if (tl_bypass) return get_cuMemFree_v2()(dptr);
    ensure_initialized();
    if (dptr == 0) return 0;  /* cuMemFree_v2@0x32e5d */
    const int vmm = g_backend_vmm;
    if (vmm) pthread_mutex_lock(&g_res_mu);
    pthread_mutex_lock(&g_alloc_mu);
    alloc_node_t *a = g_head;
    while (a && (a->addr != dptr || a->raw)) a = a->next;  /* remove_chunk@0x413dd: exact base match */
    if (!a) {
        pthread_mutex_unlock(&g_alloc_mu);
        if (vmm) pthread_mutex_unlock(&g_res_mu);
        return -1;
    }
    CUresult rc;
    if (a->vmem) {
        vmm_quiesce();  /* like cuMemFree: what was queued on the memory is done before it goes */
        tl_bypass = 1;
        rc = g_dp.destroy(a->vmem) == 0 ? 0 : 999;
        tl_bypass = 0;
        if (g_peer_left != ~0ull) g_peer_left += a->peer_bytes;
    } else {
        cuMemFree_v2_t f = get_cuMemFree_v2();
        rc = f ? f(dptr) : 3;
    }
    if (a->prev) a->prev->next = a->next; else g_head = a->next;
    if (a->next) a->next->prev = a->prev; else g_tail = a->prev;
    --g_nalloc;
    idx_del(a);
    int dev = a->dev;
    size_t len = a->len;
    free(a);
    pthread_mutex_unlock(&g_alloc_mu);
    if (vmm) pthread_mutex_unlock(&g_res_mu);
    region_rm_usage(col_of(dev), len, VGPU_USE_DATA);
    return rc;
}
EXPORT CUresult cuMemFree(CUdeviceptr dptr) { return cuMemFree_v2(dptr); }

EXPORT CUresult cuMemGetInfo_v2(size_t *free_b, size_t *total_b) {
    if (tl_bypass) return get_cuMemGetInfo_v2()(free_b, total_b);
    ensure_initialized();
    int dev = 0;
    CUresult rc = cur_dev(&dev);
    if (rc != 0) return rc;
    uint64_t usage = g_region_ok ? q_usage(dev) : 0;
    uint64_t limit = g_region_ok ? q_limit(dev) : 0;
    cuMemGetInfo_v2_t f = get_cuMemGetInfo_v2();
    if (limit == 0) {
        rc = f ? f(free_b, total_b) : 3;
        if (rc != 0) return rc;      /* nothing was written: leave the caller's values alone */
        *free_b = *total_b - usage;  /* quirk a17: free = total - usage when unlimited */
        return 0;
    }
    if (limit < usage) return 1;
    if (f) f(free_b, total_b);  /* the reference calls through and overwrites (@0x377c9) */
    *free_b = limit - usage;
    *total_b = limit;
    return 0;
}
EXPORT CUresult cuMemGetInfo(size_t *free_b, size_t *total_b) { return cuMemGetInfo_v2(free_b, total_b); }

EXPORT CUresult cuDeviceTotalMem_v2(size_t *bytes, CUdevice dev) {
    ensure_initialized();
    CUdevice phys = dev;
    if (!to_phys(&phys)) return 101;
    if (!g_region_ok) {
        cuDeviceTotalMem_v2_t f = get_cuDeviceTotalMem_v2();
        return f ? f(bytes, phys) : 3;
    }
    *bytes = q_limit(dev);  /* quirk a18: the limit, unconditionally (0 when unlimited) */
    return 0;
}
EXPORT CUresult cuDeviceTotalMem(size_t *bytes, CUdevice dev) { return cuDeviceTotalMem_v2(bytes, dev); }

/* which virtual device a pointer belongs to: the one it was allocated on if the allocation is tracked, else the
 * first vGPU of the physical GPU the driver names */
static int vdev_of_pointer(CUdeviceptr ptr, int phys) {
    ensure_vdev();
    if (!g_remap) return phys;
    int v = -1;
    pthread_mutex_lock(&g_alloc_mu);
    for (alloc_node_t *a = g_head; a; a = a->next)
        if (ptr >= a->addr && ptr < a->addr + (a->len ? a->len : 1)) {
            v = a->dev;
            break;
        }
    pthread_mutex_unlock(&g_alloc_mu);
    if (v < 0) v = first_vdev_of(phys);
    return v < 0 ? phys : v;
}

/* cuPointerGetAttributes@0x33cf3 (a24): after the real call, IS_MANAGED (8) is forced to 0 so
 * that frameworks do not notice the managed memory behind oversubscribed allocations.
 * VMM backend: a query about a managed range (cudaPointerGetAttributes of DLPack / array-interface imports, NCCL's
 * buffer checks) must see device memory whether the range is in HBM or spilled, and one allocation where the driver has
 * one mapping per block: the chunk asked about is made resident first, RANGE_START_ADDR (11) / RANGE_SIZE (12) are the
 * allocation's.  vmm_query_begin / _end live with the residency engine. */
static int vmm_query_begin(CUdeviceptr ptr, CUdeviceptr *base, size_t *len);
static void vmm_query_end(int mode);
static void pointer_attr_fix(int attr, void *data, CUresult rc, CUdeviceptr ptr, CUdeviceptr base, size_t len) {
    if (!data) return;
    if (attr == 8) *(unsigned *)data = 0;
    if (rc != 0) return;
    if (attr == 9) *(int *)data = vdev_of_pointer(ptr, *(int *)data);
    if (len && attr == 11) *(CUdeviceptr *)data = base;
    if (len && attr == 12) *(size_t *)data = len;
}
EXPORT CUresult cuPointerGetAttributes(unsigned n, int *attrs, void **data, CUdeviceptr ptr) {
    typedef CUresult (*fn_t)(unsigned, int *, void **, CUdeviceptr);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuPointerGetAttributes");
    if (tl_bypass) return f ? f(n, attrs, data, ptr) : 3;
    ensure_initialized();
    wait_running();
    CUdeviceptr base = 0;
    size_t len = 0;
    const int m = vmm_query_begin(ptr, &base, &len);
    CUresult rc = f ? f(n, attrs, data, ptr) : 3;
    vmm_query_end(m);
    for (unsigned i = 0; i < n; ++i) pointer_attr_fix(attrs[i], data[i], rc, ptr, base, len);
    return rc;
}

/* CU_POINTER_ATTRIBUTE_DEVICE_ORDINAL (9) answers with a virtual ordinal as well */
EXPORT CUresult cuPointerGetAttribute(void *data, int attr, CUdeviceptr ptr) {
    typedef CUresult (*fn_t)(void *, int, CUdeviceptr);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuPointerGetAttribute");
    if (tl_bypass) return f ? f(data, attr, ptr) : 3;
    CUdeviceptr base = 0;
    size_t len = 0;
    const int m = vmm_query_begin(ptr, &base, &len);
    CUresult rc = f ? f(data, attr, ptr) : 3;
    vmm_query_end(m);
    if (attr != 8) pointer_attr_fix(attr, data, rc, ptr, base, len);  /* the reference masks IS_MANAGED in the plural form only */
    return rc;
}
/* cuMemGetAddressRange of a pointer into a managed range: the allocation, not the block the driver mapped there */
EXPORT CUresult cuMemGetAddressRange_v2(CUdeviceptr *pbase, size_t *psize, CUdeviceptr ptr) {
    typedef CUresult (*fn_t)(CUdeviceptr *, size_t *, CUdeviceptr);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemGetAddressRange_v2");
    if (!tl_bypass) {
        CUdeviceptr base = 0;
        size_t len = 0;
        vmm_query_end(vmm_query_begin(ptr, &base, &len));
        if (len) {
            if (pbase) *pbase = base;
            if (psize) *psize = len;
            return 0;
        }
    }
    return f ? f(pbase, psize, ptr) : 3;
}
EXPORT CUresult cuMemGetAddressRange(CUdeviceptr *pbase, size_t *psize, CUdeviceptr ptr) { return cuMemGetAddressRange_v2(pbase, psize, ptr); }

/* host-memory hooks (a20): allocate first, then check_oom@0x311bd == oom_check(cur, 0) */
EXPORT CUresult cuMemAllocHost_v2(void **pp, size_t n) {
    if (tl_bypass) return get_cuMemAllocHost_v2()(pp, n);
    ensure_initialized();
    wait_running();  /* host allocations park while the vGPU is suspended, as in the reference */
    cuMemAllocHost_v2_t f = get_cuMemAllocHost_v2();
    CUresult rc = f ? f(pp, n) : 3;
    if (rc != 0) return rc;
    if (oom_check(-1, 0)) {
        if (get_cuMemFreeHost()) real_cuMemFreeHost(*pp);
        return 2;
    }
    return 0;
}
EXPORT CUresult cuMemAllocHost(void **pp, size_t n) { return cuMemAllocHost_v2(pp, n); }

EXPORT CUresult cuMemHostAlloc(void **pp, size_t n, unsigned flags) {
    if (tl_bypass) return get_cuMemHostAlloc()(pp, n, flags);
    ensure_initialized();
    wait_running();  /* host allocations park while the vGPU is suspended, as in the reference */
    cuMemHostAlloc_t f = get_cuMemHostAlloc();
    CUresult rc = f ? f(pp, n, flags) : 3;
    if (rc != 0) return rc;
    if (oom_check(-1, 0)) {
        if (get_cuMemFreeHost()) real_cuMemFreeHost(*pp);
        return 2;
    }
    return 0;
}

EXPORT CUresult cuMemHostRegister_v2(void *p, size_t n, unsigned flags) {
    if (tl_bypass) return get_cuMemHostRegister_v2()(p, n, flags);
    ensure_initialized();
    wait_running();  /* host allocations park while the vGPU is suspended, as in the reference */
    cuMemHostRegister_v2_t f = get_cuMemHostRegister_v2();
    CUresult rc = f ? f(p, n, flags) : 3;
    if (rc != 0) return rc;
    if (oom_check(-1, 0)) {
        if (get_cuMemHostUnregister()) real_cuMemHostUnregister(p);
        return 2;
    }
    return 0;
}
EXPORT CUresult cuMemHostRegister(void *p, size_t n, unsigned flags) { return cuMemHostRegister_v2(p, n, flags); }

/* ------------------------------------------------------------------ context charge (a27)
 * On a real node the reference bills every vGPU process for its CUDA context: set_task_pid@0x17de8
 * lists the GPU's compute processes through NVML, retains the primary context, lists them again; the
 * pid that appeared is this process as the HOST sees it (stored in the slot: the monitor needs it to
 * match NVML's per-process figures to containers) and its usedGpuMemory is `context_size`
 * (store @0x189de).  From then on the first cuDevicePrimaryCtxRetain@0x294bd / cuCtxCreate_v2@0x2a4dd
 * on a device adds context_size to the usage as kind 0 (@0x29c86, @0x2a80e) and sets a per-device flag;
 * cuDevicePrimaryCtxRelease_v2@0x29f08 clears the flag WITHOUT giving the bytes back (@0x2a0bf), so a
 * retain / release / retain sequence is billed twice; cuCtxDestroy gives nothing back either.  All of
 * that is reproduced (tests/golden/quota_ctx_*.json.gz, captured from the reference with the mock NVML
 * listing the process).  What differs: the reference probes at cuInit with an extra retain + release
 * of device 0 (one context creation and teardown per process start, ~0.2 s on a B200); here the probe
 * wraps the first retain / create the application makes itself, so nothing is created twice. */
static uint64_t g_context_size;
static int g_hostpid_known;  /* pidfound: the monitor can see this process, so its launch lever applies */
static int g_ctx_probed;
static int g_ctx_flag[VGPU_MAX_DEVICES];  /* vdevices[d].flag (+0x14) */
static pthread_mutex_t g_ctx_mu = PTHREAD_MUTEX_INITIALIZER;

typedef struct { unsigned pid; unsigned long long used; unsigned gi, ci; } nv_proc_t;  /* _v2/_v3 layout, 24 B */
#define PROBE_MAX 256

static void *nvml_handle_for(CUdevice dev) {
    typedef int (*byuuid_t)(const char *, void **);
    typedef int (*byidx_t)(unsigned, void **);
    void *h = NULL;
    unsigned char u[16];
    byuuid_t byuuid = (byuuid_t)nvml_sym("nvmlDeviceGetHandleByUUID");
    if (byuuid && get_cuDeviceGetUuid() && real_cuDeviceGetUuid(u, dev) == 0) {
        char s[64];
        snprintf(s, sizeof(s), "GPU-%02x%02x%02x%02x-%02x%02x-%02x%02x-%02x%02x-%02x%02x%02x%02x%02x%02x", u[0], u[1], u[2],
                 u[3], u[4], u[5], u[6], u[7], u[8], u[9], u[10], u[11], u[12], u[13], u[14], u[15]);
        if (byuuid(s, &h) == 0) return h;
    }
    byidx_t byidx = (byidx_t)nvml_sym("nvmlDeviceGetHandleByIndex_v2");
    if (!byidx) byidx = (byidx_t)nvml_sym("nvmlDeviceGetHandleByIndex");
    return (byidx && byidx((unsigned)dev, &h) == 0) ? h : NULL;
}

/* -> number of compute processes on the device (pid / used filled), -1 if NVML cannot tell */
static int list_procs(void *h, nv_proc_t *out) {
    typedef int (*procs_t)(void *, unsigned *, void *);
    unsigned n = PROBE_MAX;
    procs_t f = (procs_t)nvml_sym("nvmlDeviceGetComputeRunningProcesses_v3");
    if (!f) f = (procs_t)nvml_sym("nvmlDeviceGetComputeRunningProcesses_v2");
    if (f) return f(h, &n, out) == 0 ? (int)n : -1;
    f = (procs_t)nvml_sym("nvmlDeviceGetComputeRunningProcesses");  /* 16-byte entries */
    struct { unsigned pid; unsigned long long used; } v1[PROBE_MAX];
    if (!f || f(h, &n, v1) != 0) return -1;
    for (unsigned i = 0; i < n; ++i) {
        out[i].pid = v1[i].pid;
        out[i].used = v1[i].used;
    }
    return (int)n;
}

/* try_lock_unified_lock@0x171ac: /tmp/vgpulock is a host directory every vGPU container mounts, so
 * probes of different containers on one GPU do not see each other's new process.  The reference spins
 * on it forever; here a missing directory or a lock nobody releases within 10 s is passed by. */
static int unified_lock(void) {
    for (int i = 0; i < 10000; ++i) {
        int fd = open("/tmp/vgpulock/lock", O_CREAT | O_EXCL | O_WRONLY, 0666);
        if (fd >= 0) {
            close(fd);
            return 1;
        }
        if (errno != EEXIST) return 0;
        usleep(1000);
    }
    SHIM_WARN("/tmp/vgpulock/lock held for 10 s: probing without it");
    return 0;
}

/* Runs `make_ctx` (the application's own first retain / create) between two NVML process listings. */
static CUresult probe_around(CUdevice dev, CUresult (*make_ctx)(void *), void *arg) {
    static nv_proc_t before[PROBE_MAX], after[PROBE_MAX];
    g_ctx_probed = 1;
    void *h = nvml_handle_for(dev);
    int locked = h ? unified_lock() : 0;
    int nb = h ? list_procs(h, before) : -1;
    CUresult rc = make_ctx(arg);
    if (rc == 0 && nb >= 0) {
        int na = list_procs(h, after);
        for (int i = 0; i < na; ++i) {
            int known = 0;
            for (int j = 0; j < nb; ++j) known |= before[j].pid == after[i].pid;
            if (known) continue;
            /* getextrapid@0x17a4f + set_host_pid@0x4497c; ~0ull = NVML_VALUE_NOT_AVAILABLE (e.g. under WDDM / vGPU hosts) */
            g_context_size = after[i].used == ~0ull ? 0 : after[i].used;
            region_set_hostpid((int)after[i].pid);
            g_hostpid_known = 1;
            SHIM_INFO("host pid %u, context size %lu", after[i].pid, (unsigned long)g_context_size);
            break;
        }
    }
    if (rc == 0 && !g_context_size) SHIM_WARN("SET_TASK_PID FAILED.");  /* libvgpu.c:905 */
    if (locked) unlink("/tmp/vgpulock/lock");
    return rc;
}

static void charge_context(CUdevice dev, int created) {
    if (dev < 0 || dev >= VGPU_MAX_DEVICES) return;
    if (g_ctx_flag[dev]) {
        if (created) SHIM_ERR("Duplicate cuCtxCreate, this may indicate errors");  /* context.c:144 */
        return;
    }
    if (g_context_size && g_region_ok) region_add_usage(col_of(dev), g_context_size, VGPU_USE_CONTEXT);
    g_ctx_flag[dev] = 1;
}

struct retain_args { CUcontext *pctx; CUdevice dev; unsigned flags; };
static CUresult do_retain(void *a) {
    struct retain_args *r = (struct retain_args *)a;
    cuDevicePrimaryCtxRetain_t f = get_cuDevicePrimaryCtxRetain();
    return f ? f(r->pctx, r->dev) : 3;
}
static CUresult do_create(void *a) {
    struct retain_args *r = (struct retain_args *)a;
    cuCtxCreate_v2_t f = get_cuCtxCreate_v2();
    return f ? f(r->pctx, r->flags, r->dev) : 3;
}

/* ------------------------------------------------------------------ SM cap: green contexts
 * Replaces rate_limiter@0x463b1 / utilization_watcher@0x471a7 (a25).  CUDA_DEVICE_SM_LIMIT=p
 * with 0 < p < 100 turns into a green context of floor(SMs*p/100) SMs (stepped down to the
 * driver's granule so the vGPU never exceeds its share); its CUcontext is handed to the
 * application wherever it asks for the device's primary context, so the runtime API and
 * every library on top of it run inside the partition.  Which of the disjoint partitions a
 * container gets: VGPU_SM_PARTITION_INDEX[_<dev>] from the plugin (vgpu_plugin_cfg_t.emit_partition_index); without it
 * the placement is not coordinated (Contract A carries no vGPU index into the pod) and the shim says so.
 * VGPU_SM_PARTITION=off disables the cap. */
typedef struct {
    int type;
    unsigned char pad[92];
    union { struct { unsigned smCount; } sm; unsigned char oversize[48]; };
} shim_dev_resource_t;
_Static_assert(sizeof(shim_dev_resource_t) == 144, "CUdevResource ABI v1");
typedef void *CUgreenCtx;
typedef void *CUdevResourceDesc;

static CUcontext g_green_ctx[VGPU_MAX_DEVICES];
static CUcontext g_primary_ctx[VGPU_MAX_DEVICES];  /* the real primary context the green one stands in for */
static int g_green_state[VGPU_MAX_DEVICES];  /* 0 untried, 1 active, -1 unavailable */
static unsigned g_green_sms[VGPU_MAX_DEVICES];
static pthread_mutex_t g_green_mu = PTHREAD_MUTEX_INITIALIZER;

/* contexts the application made with cuCtxCreate on a virtual device */
#define CTX_REG_MAX 64
static struct { CUcontext ctx; int vdev; } g_ctx_reg[CTX_REG_MAX];
static int g_ctx_reg_n;
static void ctx_register(CUcontext ctx, int vdev) {
    for (int i = 0; i < g_ctx_reg_n; ++i)
        if (g_ctx_reg[i].ctx == ctx) { g_ctx_reg[i].vdev = vdev; return; }
    if (g_ctx_reg_n < CTX_REG_MAX) {
        g_ctx_reg[g_ctx_reg_n].ctx = ctx;
        g_ctx_reg[g_ctx_reg_n++].vdev = vdev;
    }
}
static int vdev_of_ctx(CUcontext ctx, int phys) {
    int v = -1;
    if (ctx) {
        pthread_mutex_lock(&g_ctx_mu);  /* the tables grow under it (retain / create hooks) */
        for (int i = 0; i < g_nvdev && v < 0; ++i)
            if (g_vdev[i].ctx == ctx || (g_green_state[i] == 1 && g_green_ctx[i] == ctx)) v = i;
        for (int i = 0; i < g_ctx_reg_n && v < 0; ++i)
            if (g_ctx_reg[i].ctx == ctx) v = g_ctx_reg[i].vdev;
        pthread_mutex_unlock(&g_ctx_mu);
    }
    return v >= 0 ? v : first_vdev_of(phys);
}

/* GPU_CORE_UTILIZATION_POLICY (set_env_utilization_switch@0x44b3b): DISABLE / disable switches the reference's launch
 * throttle off, FORCE / force keeps it on whatever the monitor says.  The cap here is spatial and always on, so FORCE
 * is the normal case and DISABLE means: no partition. */
static int sm_policy_disabled(void) {
    const char *p = getenv("GPU_CORE_UTILIZATION_POLICY");
    return p && (strcmp(p, "DISABLE") == 0 || strcmp(p, "disable") == 0);
}

/* Which of the nb disjoint groups is this vGPU's?  VGPU_SM_PARTITION_INDEX_<dev> (emitted by the plugin when
 * vgpu_plugin_cfg_t.emit_partition_index is set: the i of "<UUID>-<i>"), else VGPU_SM_PARTITION_INDEX for all
 * devices.  Without either nothing in the pod says which vGPU of the GPU this is (Contract A carries no index),
 * and guessing would put containers on top of each other while claiming isolation: say so and take group 0. */
static unsigned partition_pick(int dev, unsigned nb) {
    char name[64];
    snprintf(name, sizeof(name), "VGPU_SM_PARTITION_INDEX_%d", dev);
    const char *e = getenv(name);
    if (!e || !*e) e = getenv("VGPU_SM_PARTITION_INDEX");
    if (e && *e) return (unsigned)strtoul(e, NULL, 0) % nb;
    SHIM_ERR("device %d: no VGPU_SM_PARTITION_INDEX: SM partition placement is NOT coordinated with the other vGPUs "
             "of this GPU (all such containers share group 0); the size of the cap holds, the isolation does not", dev);
    return 0;
}

static CUcontext green_ctx_for(CUdevice dev) {
    if (dev < 0 || dev >= VGPU_MAX_DEVICES || !g_region_ok) return NULL;
    uint64_t pct = q_sm_limit(dev);
    const char *sw = getenv("VGPU_SM_PARTITION");
    if (pct == 0 || pct >= 100 || (sw && strcmp(sw, "off") == 0) || sm_policy_disabled()) return NULL;
    pthread_mutex_lock(&g_green_mu);
    if (g_green_state[dev] == 0) {
        g_green_state[dev] = -1;
        typedef CUresult (*getres_t)(CUdevice, shim_dev_resource_t *, int);
        typedef CUresult (*split_t)(shim_dev_resource_t *, unsigned *, const shim_dev_resource_t *, shim_dev_resource_t *, unsigned, unsigned);
        typedef CUresult (*gendesc_t)(CUdevResourceDesc *, shim_dev_resource_t *, unsigned);
        typedef CUresult (*create_t)(CUgreenCtx *, CUdevResourceDesc, CUdevice, unsigned);
        typedef CUresult (*fromg_t)(CUcontext *, CUgreenCtx);
        getres_t getres = (getres_t)cuda_sym("cuDeviceGetDevResource");
        split_t split = (split_t)cuda_sym("cuDevSmResourceSplitByCount");
        gendesc_t gendesc = (gendesc_t)cuda_sym("cuDevResourceGenerateDesc");
        create_t create = (create_t)cuda_sym("cuGreenCtxCreate");
        fromg_t fromg = (fromg_t)cuda_sym("cuCtxFromGreenCtx");
        shim_dev_resource_t all, groups[64], rest;
        if (getres && split && gendesc && create && fromg && getres(phys_of(dev), &all, 1 /* CU_DEV_RESOURCE_TYPE_SM */) == 0) {
            unsigned total = all.sm.smCount, share = (unsigned)(total * pct / 100), nb = 0, want;
            /* VGPU_SM_PARTITION_GRANULE=fine: SM-granular groups (IGNORE_SM_COSCHEDULING = 1), e.g.
             * 10 % of 148 SMs -> 14 instead of 8; default keeps cluster co-scheduling inside the vGPU */
            const char *gr = getenv("VGPU_SM_PARTITION_GRANULE");
            const unsigned sflags = (gr && strcmp(gr, "fine") == 0) ? 1u : 0u;
            for (want = share ? share : 1; want >= 1; --want) {
                nb = 64;
                if (split(groups, &nb, &all, &rest, sflags, want) == 0 && nb >= 1 && (groups[0].sm.smCount <= share || want == 1)) break;
                nb = 0;
            }
            if (nb) {
                unsigned k = partition_pick(dev, nb);
                CUdevResourceDesc desc;
                CUgreenCtx g;
                CUcontext c;
                if (gendesc(&desc, &groups[k], 1) == 0 && create(&g, desc, phys_of(dev), 1 /* CU_GREEN_CTX_DEFAULT_STREAM */) == 0 &&
                    fromg(&c, g) == 0) {
                    g_green_ctx[dev] = c;
                    g_green_sms[dev] = groups[k].sm.smCount;
                    g_green_state[dev] = 1;
                    SHIM_INFO("device %d: SM limit %lu%% -> green context %u of %u, %u of %u SMs", dev, (unsigned long)pct,
                              k, nb, groups[k].sm.smCount, total);
                }
            }
        }
        if (g_green_state[dev] != 1)
            SHIM_ERR("device %d: cannot create a green context for SM limit %lu%%: the SM cap is NOT enforced", dev,
                     (unsigned long)pct);
    }
    CUcontext c = g_green_state[dev] == 1 ? g_green_ctx[dev] : NULL;
    pthread_mutex_unlock(&g_green_mu);
    return c;
}

/* SMs the partition of `dev` has, or will have once its primary context is retained (0 = no cap): the same split the
 * creation does, without creating anything — what device attributes must answer before the context exists. */
static unsigned green_planned_sms(CUdevice dev) {
    static unsigned planned[VGPU_MAX_DEVICES];
    static int done[VGPU_MAX_DEVICES];
    if (dev < 0 || dev >= VGPU_MAX_DEVICES || !g_region_ok) return 0;
    if (g_green_state[dev] == 1) return g_green_sms[dev];
    if (g_green_state[dev] == -1) return 0;
    uint64_t pct = q_sm_limit(dev);
    const char *sw = getenv("VGPU_SM_PARTITION");
    if (pct == 0 || pct >= 100 || (sw && strcmp(sw, "off") == 0) || sm_policy_disabled()) return 0;
    if (done[dev]) return planned[dev];
    typedef CUresult (*getres_t)(CUdevice, shim_dev_resource_t *, int);
    typedef CUresult (*split_t)(shim_dev_resource_t *, unsigned *, const shim_dev_resource_t *, shim_dev_resource_t *, unsigned, unsigned);
    getres_t getres = (getres_t)cuda_sym("cuDeviceGetDevResource");
    split_t split = (split_t)cuda_sym("cuDevSmResourceSplitByCount");
    shim_dev_resource_t all, groups[64], rest;
    if (getres && split && getres(phys_of(dev), &all, 1) == 0) {
        const unsigned total = all.sm.smCount, share = (unsigned)(total * pct / 100);
        const char *gr = getenv("VGPU_SM_PARTITION_GRANULE");
        const unsigned sflags = (gr && strcmp(gr, "fine") == 0) ? 1u : 0u;
        for (unsigned want = share ? share : 1; want >= 1; --want) {
            unsigned nb = 64;
            if (split(groups, &nb, &all, &rest, sflags, want) == 0 && nb >= 1 && (groups[0].sm.smCount <= share || want == 1)) {
                planned[dev] = groups[0].sm.smCount;
                break;
            }
        }
        done[dev] = 1;
    }
    return planned[dev];
}

/* exported for tests / vgpuvalidator: SMs of the partition this process runs in (0 = uncapped) */
EXPORT unsigned cuVGPUPartitionSmCount(CUdevice dev) {
    return (dev >= 0 && dev < VGPU_MAX_DEVICES && g_green_state[dev] == 1) ? g_green_sms[dev] : 0;
}

EXPORT void *cuVGPUPartitionCtx(CUdevice dev) {
    return (dev >= 0 && dev < VGPU_MAX_DEVICES && g_green_state[dev] == 1) ? g_green_ctx[dev] : NULL;
}

REAL(CUresult, cuCtxPopCurrent_v2, CUcontext *)
EXPORT CUresult cuDevicePrimaryCtxRetain(CUcontext *pctx, CUdevice dev) {
    ensure_initialized();
    SHIM_DEBUG("cuDevicePrimaryCtxRetain(dev=%d)", dev);
    if (tl_bypass) return get_cuDevicePrimaryCtxRetain()(pctx, dev);
    CUdevice phys = dev;
    if (!to_phys(&phys)) return 101;
    const int shared = g_remap && g_vdev[dev].dup > 0;
    struct retain_args a = {pctx, phys, 0};
    CUresult rc = 0;
    pthread_mutex_lock(&g_ctx_mu);
    if (shared && g_vdev[dev].ctx) {
        *pctx = g_vdev[dev].ctx;
    } else {
        /* a further vGPU of a GPU whose primary context belongs to the first one gets a context of its own
         * (@0x297be); cuCtxCreate makes it current, a primary retain must not: pop it again */
        CUresult (*make)(void *) = shared ? do_create : do_retain;
        rc = (!g_ctx_probed && g_region_ok) ? probe_around(phys, make, &a) : make(&a);
        if (rc == 0 && shared) {
            CUcontext popped;
            if (get_cuCtxPopCurrent_v2()) real_cuCtxPopCurrent_v2(&popped);
            g_vdev[dev].ctx = *pctx;
        }
    }
    if (rc == 0) {
        if (shared) ++g_vdev[dev].refs;
        charge_context(dev, 0);
    }
    pthread_mutex_unlock(&g_ctx_mu);
    if (rc != 0) return rc;
    if (dev >= 0 && dev < VGPU_MAX_DEVICES) {
        g_primary_ctx[dev] = *pctx;
        green_ctx_for(dev);  /* created here; bound in place of the primary by the cuCtx* hooks */
    }
    return 0;
}

EXPORT CUresult cuDevicePrimaryCtxRelease_v2(CUdevice dev) {
    if (tl_bypass) return get_cuDevicePrimaryCtxRelease_v2()(dev);
    CUdevice phys = dev;
    if (!to_phys(&phys)) return 101;
    CUresult rc = 0;
    if (g_remap && g_vdev[dev].dup > 0) {  /* the stand-in context lives as long as the process */
        pthread_mutex_lock(&g_ctx_mu);
        if (g_vdev[dev].refs > 0) --g_vdev[dev].refs; else rc = 201;
        pthread_mutex_unlock(&g_ctx_mu);
    } else {
        cuDevicePrimaryCtxRelease_v2_t f = get_cuDevicePrimaryCtxRelease_v2();
        rc = f ? f(phys) : 3;
    }
    if (dev >= 0 && dev < VGPU_MAX_DEVICES) {
        pthread_mutex_lock(&g_ctx_mu);
        g_ctx_flag[dev] = 0;  /* @0x2a0bf: the flag goes, the bytes stay billed */
        pthread_mutex_unlock(&g_ctx_mu);
    }
    return rc;
}
EXPORT CUresult cuDevicePrimaryCtxRelease(CUdevice dev) { return cuDevicePrimaryCtxRelease_v2(dev); }

EXPORT CUresult cuCtxCreate_v2(CUcontext *pctx, unsigned flags, CUdevice dev) {
    if (tl_bypass) return get_cuCtxCreate_v2()(pctx, flags, dev);
    ensure_initialized();
    CUdevice phys = dev;
    if (!to_phys(&phys)) return 101;
    struct retain_args a = {pctx, phys, flags};
    pthread_mutex_lock(&g_ctx_mu);
    CUresult rc = (!g_ctx_probed && g_region_ok) ? probe_around(phys, do_create, &a) : do_create(&a);
    if (rc == 0) {
        charge_context(dev, 1);
        if (g_remap) ctx_register(*pctx, dev);
    }
    pthread_mutex_unlock(&g_ctx_mu);
    return rc;
}
EXPORT CUresult cuCtxCreate(CUcontext *pctx, unsigned flags, CUdevice dev) { return cuCtxCreate_v2(pctx, flags, dev); }

/* cudart keeps its own handle on the primary context (it has it before it ever calls
 * cuDevicePrimaryCtxRetain) and binds THAT.  So: whenever a context is bound, find out whether
 * it is a capped device's primary context and, if so, bind the partition in its place; and
 * when the application asks which context is current, answer with the primary it expects. */
static CUcontext partition_after_bind(CUcontext ctx) {
    /* called with `ctx` already current */
    int dev = -1;
    if (!ctx || cur_dev(&dev) != 0 || dev < 0 || dev >= VGPU_MAX_DEVICES) return NULL;
    if (g_green_state[dev] == 0) {
        CUcontext p = NULL;
        cuDevicePrimaryCtxRetain_t f = get_cuDevicePrimaryCtxRetain();
        if (g_remap && g_vdev[dev].dup > 0) p = g_vdev[dev].ctx;  /* its stand-in for a primary context, if made yet */
        else if (!f || f(&p, phys_of(dev)) != 0) p = NULL;
        if (p) {
            g_primary_ctx[dev] = p;
            green_ctx_for(dev);
        }
    }
    return (g_green_state[dev] == 1 && g_primary_ctx[dev] == ctx) ? g_green_ctx[dev] : NULL;
}

static int caps_possible(void) {
    static int v = -1;
    if (v < 0) {
        const char *sw = getenv("VGPU_SM_PARTITION");
        v = 0;
        if (g_region_ok && !(sw && strcmp(sw, "off") == 0) && !sm_policy_disabled())
            for (int d = 0; d < VGPU_MAX_DEVICES; ++d) {
                uint64_t p = region_sm_limit(d);
                if (p > 0 && p < 100) v = 1;
            }
    }
    return v;
}

EXPORT CUresult cuCtxSetCurrent(CUcontext ctx) {
    ensure_initialized();
    cuCtxSetCurrent_t f = get_cuCtxSetCurrent();
    CUresult rc = f ? f(ctx) : 3;
    if (rc != 0 || !caps_possible()) return rc;
    CUcontext g = partition_after_bind(ctx);
    SHIM_DEBUG("cuCtxSetCurrent(%p) -> %p", ctx, g ? g : ctx);
    return g ? f(g) : rc;
}

EXPORT CUresult cuCtxPushCurrent_v2(CUcontext ctx) {
    ensure_initialized();
    cuCtxPushCurrent_v2_t f = get_cuCtxPushCurrent_v2();
    CUresult rc = f ? f(ctx) : 3;
    if (rc != 0 || !caps_possible()) return rc;
    CUcontext g = partition_after_bind(ctx);
    SHIM_DEBUG("cuCtxPushCurrent(%p) -> %p", ctx, g ? g : ctx);
    if (g) {  /* replace the top of the stack */
        cuCtxSetCurrent_t set = get_cuCtxSetCurrent();
        if (set) return set(g);
    }
    return rc;
}

EXPORT CUresult cuCtxGetCurrent(CUcontext *pctx) {
    ensure_initialized();
    cuCtxGetCurrent_t f = get_cuCtxGetCurrent();
    CUresult rc = f ? f(pctx) : 3;
    if (rc == 0 && *pctx)
        for (int d = 0; d < VGPU_MAX_DEVICES; ++d)
            if (g_green_state[d] == 1 && g_green_ctx[d] == *pctx) *pctx = g_primary_ctx[d];
    return rc;
}

EXPORT CUresult cuCtxPopCurrent_v2(CUcontext *pctx) {
    ensure_initialized();
    typedef CUresult (*pop_t)(CUcontext *);
    static pop_t f;
    if (!f) f = (pop_t)cuda_sym("cuCtxPopCurrent_v2");
    CUresult rc = f ? f(pctx) : 3;
    if (rc == 0 && pctx && *pctx)
        for (int d = 0; d < VGPU_MAX_DEVICES; ++d)
            if (g_green_state[d] == 1 && g_green_ctx[d] == *pctx) *pctx = g_primary_ctx[d];
    return rc;
}
EXPORT CUresult cuCtxPopCurrent(CUcontext *pctx) { return cuCtxPopCurrent_v2(pctx); }
EXPORT CUresult cuCtxPushCurrent(CUcontext ctx) { return cuCtxPushCurrent_v2(ctx); }

/* ------------------------------------------------------------------ residency under oversubscription
 * (VGPU_OVERSUB_BACKEND=vmm)  What the reference gets from UVM page faults — an application whose allocations add up
 * to more than the HBM it can have — without page faults: every allocation above 2 MiB is a range of the
 * virtual-device-memory manager, and WHICH ranges are backed by HBM is decided at the calls that can touch them.
 *   - While everything fits, allocations are backed at once and no call pays anything ("unpaged": one shared lock).
 *   - From the first allocation that does not fit, the process is "paged": a kernel launch scans its argument
 *     buffer (cuFuncGetParamInfo gives each parameter's size) for words that point into a managed range — into
 *     structs passed by value as well — and makes those ranges resident before the launch; the 1-D / 2-D / 3-D copy
 *     and memset calls do the same for exactly the bytes they name; a graph launch needs every range.  Room is made
 *     by spilling the least recently used ranges that the call at hand does not need (spill kernel -> tier: the
 *     named peer GPU's HBM first, pinned host RAM second), after draining the contexts of this process.
 *   - cuMemPrefetchAsync / cuMemAdvise stay the explicit controls, and are the ONLY way to page a range that is
 *     larger than the GPU itself ("windowed": a launch cannot know which part of it a kernel will read).
 * What this cannot see: pointers a kernel finds in device memory (pointer tables of batched BLAS, device-side
 * linked structures), arguments baked into a captured graph after the ranges moved, and peer / IPC access to the
 * ranges.  Such a program keeps its working set resident with prefetch hints, or stays on the managed backend.
 * Locks: g_mode_rw (shared by every unpaged call, exclusive for the switch to paged), then g_res_mu (paged calls,
 * from the residency decision to the driver call it protects), then g_alloc_mu. */
REAL(CUresult, cuCtxSynchronize, void)
static pthread_rwlock_t g_mode_rw = PTHREAD_RWLOCK_INITIALIZER;
static volatile int g_vmm_paged;
static uint64_t g_res_clock, g_res_epoch;
static CUdeviceptr g_vmm_lo = ~0ull, g_vmm_hi;  /* envelope of every managed range ever made: cheap reject for the scan */

static uint64_t env_mb(const char *name, uint64_t dflt) {
    const char *e = getenv(name);
    return (e && *e) ? strtoull(e, NULL, 0) << 20 : dflt;
}
static uint64_t vmm_reserve(void) { return env_mb("VGPU_OVERSUB_RESERVE_MB", 1ull << 30); }

static int res_begin(void) {  /* 0: no managed ranges in this process, 1: unpaged (shared), 2: paged (g_res_mu held) */
    if (!g_backend_vmm) return 0;
    if (!g_vmm_paged) {
        pthread_rwlock_rdlock(&g_mode_rw);
        if (!g_vmm_paged) return 1;
        pthread_rwlock_unlock(&g_mode_rw);
    }
    pthread_mutex_lock(&g_res_mu);
    return 2;
}
static void res_end(int m) {
    if (m == 1) pthread_rwlock_unlock(&g_mode_rw);
    else if (m == 2) pthread_mutex_unlock(&g_res_mu);
}
static void vmm_go_paged(void) {  /* waits for the unpaged calls in flight: after it every call decides residency */
    if (g_vmm_paged) return;
    pthread_rwlock_wrlock(&g_mode_rw);
    g_vmm_paged = 1;
    pthread_rwlock_unlock(&g_mode_rw);
    SHIM_INFO("oversubscribed: managed allocations no longer fit the GPU together, residency is decided per call from here on");
}

/* everything this process has queued, in every context it owns on the GPU, is done */
static void vmm_quiesce(void) {
    const int was = tl_bypass;
    tl_bypass = 1;
    if (get_cuCtxSynchronize()) real_cuCtxSynchronize();
    CUcontext cur = NULL, popped;
    if (get_cuCtxGetCurrent()) real_cuCtxGetCurrent(&cur);
    CUcontext seen[2 * VGPU_MAX_DEVICES + CTX_REG_MAX];
    int n = 0;
    for (int d = 0; d < VGPU_MAX_DEVICES; ++d) {
        if (g_green_state[d] == 1 && g_green_ctx[d]) seen[n++] = g_green_ctx[d];
        if (g_primary_ctx[d]) seen[n++] = g_primary_ctx[d];
    }
    for (int i = 0; i < g_nvdev; ++i)
        if (g_vdev[i].ctx) seen[n++] = g_vdev[i].ctx;
    for (int i = 0; i < g_ctx_reg_n; ++i) seen[n++] = g_ctx_reg[i].ctx;
    for (int i = 0; i < n; ++i) {
        int dup = seen[i] == cur;
        for (int k = 0; k < i; ++k) dup |= seen[k] == seen[i];
        if (dup || !get_cuCtxPushCurrent_v2() || !get_cuCtxPopCurrent_v2()) continue;
        if (real_cuCtxPushCurrent_v2(seen[i]) == 0) {
            real_cuCtxSynchronize();
            real_cuCtxPopCurrent_v2(&popped);
        }
    }
    tl_bypass = was;
}

static uint64_t node_chunks(const alloc_node_t *a) { return (a->len + IPCSIZE - 1) / IPCSIZE; }
static uint64_t node_resident(alloc_node_t *a) {  /* chunks backed by HBM right now */
    uint64_t miss = 0;
    g_dp.missing(a->vmem, 0, node_chunks(a), &miss);
    return node_chunks(a) > miss ? node_chunks(a) - miss : 0;
}
/* the managed range that contains p (g_alloc_mu held): binary search in the address-sorted index, since a launch looks
 * up every 8-byte word of its arguments and a framework's allocator holds hundreds of ranges */
static size_t idx_upper(CUdeviceptr p) {  /* number of ranges whose base is <= p */
    size_t lo = 0, hi = g_idx_n;
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        if (g_idx[mid]->addr <= p) lo = mid + 1; else hi = mid;
    }
    return lo;
}
static void idx_add(alloc_node_t *a) {
    if (g_idx_n == g_idx_cap) {
        const size_t cap = g_idx_cap ? g_idx_cap * 2 : 256;
        alloc_node_t **n = (alloc_node_t **)realloc(g_idx, cap * sizeof(*n));
        if (!n) abort();
        g_idx = n;
        g_idx_cap = cap;
    }
    const size_t at = idx_upper(a->addr);
    memmove(g_idx + at + 1, g_idx + at, (g_idx_n - at) * sizeof(*g_idx));
    g_idx[at] = a;
    ++g_idx_n;
}
static void idx_del(alloc_node_t *a) {
    if (!a->vmem) return;
    const size_t at = idx_upper(a->addr);
    if (at == 0 || g_idx[at - 1] != a) return;
    memmove(g_idx + at - 1, g_idx + at, (g_idx_n - at) * sizeof(*g_idx));
    --g_idx_n;
}
static alloc_node_t *vmm_find(CUdeviceptr p) {
    if (p < g_vmm_lo || p >= g_vmm_hi) return NULL;
    const size_t at = idx_upper(p);
    if (at == 0) return NULL;
    alloc_node_t *a = g_idx[at - 1];
    return p < a->addr + node_chunks(a) * IPCSIZE ? a : NULL;
}
/* least recently used range with something resident that the call at hand (epoch) does not need */
static alloc_node_t *vmm_victim(uint64_t epoch) {
    /* cuMemAdvise(SET_PREFERRED_LOCATION) ranks before age: ranges that prefer the CPU go first, ranges that prefer
     * the GPU only when nothing else is left */
    alloc_node_t *best = NULL;
    for (alloc_node_t *a = g_head; a; a = a->next)
        if (a->vmem && a->epoch != epoch &&
            (!best || a->prefer < best->prefer || (a->prefer == best->prefer && a->last_use < best->last_use)) && node_resident(a) > 0)
            best = a;
    return best;
}
static int vmm_evict_node(alloc_node_t *v) {
    vmm_quiesce();
    SHIM_INFO("oversubscribed: spilling range 0x%llx (%zu MiB, last used at tick %lu)", v->addr, v->len >> 20, (unsigned long)v->last_use);
    int rc = g_dp.evict(v->vmem, 0, node_chunks(v));
    if (rc != 0) SHIM_ERR("vmm backend evict: %s", g_dp.last_error());
    return rc;
}

static int vmm_swap_nodes(alloc_node_t *v, alloc_node_t *a, uint64_t first, uint64_t n) {
    vmm_quiesce();
    SHIM_INFO("oversubscribed: range 0x%llx (%zu MiB, last used at tick %lu) leaves while 0x%llx arrives", v->addr, v->len >> 20,
              (unsigned long)v->last_use, a->addr);
    int rc = g_dp.swap(v->vmem, a->vmem, first, n);
    if (rc != 0) SHIM_ERR("vmm backend swap: %s", g_dp.last_error());
    return rc;
}

/* the driver itself ran out of HBM (g_res_mu and g_alloc_mu held): spill one range.  -> 1 if something was spilled */
static int vmm_make_room(void) {
    alloc_node_t *v = vmm_victim(++g_res_epoch);
    if (!v) return 0;
    vmm_go_paged();
    const int was = tl_bypass;
    tl_bypass = 1;
    const int rc = vmm_evict_node(v);
    tl_bypass = was;
    return rc == 0;
}

/* after an allocation outside the managed ranges: the driver's own needs (module loading at a kernel's first launch,
 * library workspaces) come out of the same free HBM, and no launch of an unmanaged tensor would restore it */
static void vmm_keep_reserve(void) {
    const int was = tl_bypass;
    for (;;) {
        size_t fr = 0, tot = 0;
        tl_bypass = 1;
        const int have = get_cuMemGetInfo_v2() && real_cuMemGetInfo_v2(&fr, &tot) == 0;
        tl_bypass = was;
        if (!have || fr >= vmm_reserve() / 2 || !vmm_make_room()) return;
    }
}

/* Make chunks [first, first+n) of `a` resident for the call `epoch`.  g_res_mu and g_alloc_mu held.  -> 0 or a VGPU_E_* */
static int vmm_make_resident(alloc_node_t *a, uint64_t first, uint64_t n, uint64_t epoch) {
    uint64_t miss = 0;
    const int was = tl_bypass;
    tl_bypass = 1;
    a->last_use = ++g_res_clock;
    a->epoch = epoch;
    int rc = g_dp.missing(a->vmem, first, n, &miss);
    if (rc != 0 || miss == 0) {
        tl_bypass = was;
        return rc;
    }
    const uint64_t cap = env_mb("VGPU_OVERSUB_RESIDENT_MB", 0);  /* operator's bound on this vGPU's resident managed bytes */
    for (;;) {
        size_t fr = 0, tot = 0;
        if (get_cuMemGetInfo_v2()) real_cuMemGetInfo_v2(&fr, &tot);
        int tight = fr < miss * IPCSIZE + vmm_reserve();
        if (!tight && cap) {
            uint64_t res = 0;
            for (alloc_node_t *k = g_head; k; k = k->next)
                if (k->vmem) res += node_resident(k) * IPCSIZE;
            tight = res + miss * IPCSIZE > cap;
        }
        if (tight) {
            alloc_node_t *v = vmm_victim(epoch);
            /* the victim leaves while this range arrives, both directions of the link at once, into the HBM the
             * victim frees block by block; what is still missing afterwards goes round the loop again */
            if (v && g_dp.swap && v->dev == a->dev && vmm_swap_nodes(v, a, first, n) == 0) {
                if (g_dp.missing(a->vmem, first, n, &miss) != 0 || miss == 0) break;
                continue;
            }
            if (v && vmm_evict_node(v) == 0) continue;
            /* nothing of ours left to spill: the HBM is held by other tenants.  Do NOT try anyway — the manager maps
             * block by block until the driver refuses, which would eat the reserve every other allocation on the GPU
             * (ours below 2 MiB, the other tenants') depends on.  The caller waits and retries. */
            rc = -3;
            break;
        }
        rc = g_dp.ensure(a->vmem, first, n);
        if (rc == -3 /* VGPU_E_NOMEM */) {
            alloc_node_t *v = vmm_victim(epoch);
            if (v && vmm_evict_node(v) == 0) continue;
        }
        break;
    }
    if (rc != 0) SHIM_ERR("vmm backend: cannot make 0x%llx + %lu chunks resident: %s", a->addr, (unsigned long)n, g_dp.last_error());
    tl_bypass = was;
    return rc;
}

/* The peer GPU of the spill tier as a driver ordinal: VGPU_OVERSUB_PEER=<GPU UUID> (kept visible to the driver behind
 * the vGPU's own GPUs by narrow_visible_devices, hidden from the application by the virtual-device table) or, for
 * tests, a plain driver ordinal.  -1: none. */
static int vmm_peer_ordinal(void) {
    static int cached = -2;
    if (cached != -2) return cached;
    cached = -1;
    const char *pe = getenv("VGPU_OVERSUB_PEER");
    if (!pe || !*pe) return cached;
    int n = 0;
    if (!get_cuDeviceGetCount() || real_cuDeviceGetCount(&n) != 0) return cached;
    if (strncasecmp(pe, "GPU-", 4) != 0) {
        int o = atoi(pe);
        if (o >= 0 && o < n) cached = o;
    } else if (get_cuDeviceGetUuid()) {
        for (int i = 0; i < n && cached < 0; ++i) {
            unsigned char u[16];
            char txt[VGPU_UUID_LEN];
            if (real_cuDeviceGetUuid(u, i) != 0) continue;
            uuid_text(u, txt, sizeof(txt));
            if (strcasecmp(txt, pe) == 0) cached = i;
        }
    }
    if (cached < 0) SHIM_ERR("VGPU_OVERSUB_PEER=%s names no GPU this process can open: spilling to pinned host RAM only", pe);
    else g_peer_left = env_mb("VGPU_OVERSUB_PEER_MB", ~0ull);
    return cached;
}

/* cuMemAlloc of n > 2 MiB bytes on the managed route.  g_res_mu and g_alloc_mu held. */
static int vmm_alloc(int dev, size_t n, CUdeviceptr *p) {
    const uint64_t chunks = (n + IPCSIZE - 1) / IPCSIZE, bytes = chunks * IPCSIZE;
    size_t fr = 0, tot = 0;
    if (get_cuMemGetInfo_v2()) real_cuMemGetInfo_v2(&fr, &tot);
    const uint64_t reserve = vmm_reserve();
    uint64_t cap = tot > 2 * reserve ? tot - reserve : tot / 2;  /* what a single range may ever have resident */
    const uint64_t op_cap = env_mb("VGPU_OVERSUB_RESIDENT_MB", 0);
    if (op_cap && op_cap < cap) cap = op_cap;
    cap -= cap % IPCSIZE;
    const int windowed = bytes > cap;
    const int phys = phys_of(dev), peer = vmm_peer_ordinal();
    uint64_t peer_bytes = 0;
    if (peer >= 0 && peer != phys && g_peer_left >= IPCSIZE) {
        peer_bytes = bytes < g_peer_left ? bytes : g_peer_left - g_peer_left % IPCSIZE;
        if (g_peer_left != ~0ull) g_peer_left -= peer_bytes;
    }
    void *h = NULL, *base = NULL;
    tl_bypass = 1;
    int vrc = g_dp.create_tiered(phys, n, windowed ? cap : bytes, peer, peer_bytes, bytes, &h);
    if (vrc == 0) vrc = g_dp.base(h, &base);
    uint64_t res_now = 0;
    if (op_cap)
        for (alloc_node_t *k = g_head; k; k = k->next)
            if (k->vmem) res_now += node_resident(k) * IPCSIZE;
    if (vrc == 0 && !g_vmm_paged && !windowed && fr >= bytes + reserve && !(op_cap && res_now + bytes > op_cap)) {
        vrc = g_dp.ensure(h, 0, chunks);  /* it fits: back it now, as cuMemAlloc would */
        if (vrc == -3) vrc = 0;           /* ... it does not after all: page it in at its first use */
    }
    tl_bypass = 0;
    if (vrc != 0) {
        SHIM_ERR("vmm backend: %s", g_dp.last_error());
        tl_bypass = 1;
        if (h) g_dp.destroy(h);
        tl_bypass = 0;
        if (g_peer_left != ~0ull) g_peer_left += peer_bytes;
        return vrc;
    }
    *p = (CUdeviceptr)(uintptr_t)base;
    track_vmem(*p, n, dev, h);
    g_tail->windowed = windowed;
    g_tail->peer_bytes = peer_bytes;
    g_tail->last_use = ++g_res_clock;
    idx_add(g_tail);
    if (*p < g_vmm_lo) g_vmm_lo = *p;
    if (*p + bytes > g_vmm_hi) g_vmm_hi = *p + bytes;
    uint64_t miss = 0;
    tl_bypass = 1;
    g_dp.missing(h, 0, chunks, &miss);
    tl_bypass = 0;
    if (miss) vmm_go_paged();
    return 0;
}

/* ---- what a call needs: launches (argument scan), copies (exact bytes), graphs (everything) */
typedef CUresult (*paraminfo_t)(void *, size_t, size_t *, size_t *);
#define PSCAN_MAX 96
typedef struct { void *fn; int n; unsigned short size[PSCAN_MAX]; } pscan_t;
static pscan_t g_pscan[1024];  /* by kernel handle, under g_res_mu */
static const pscan_t *kernel_params(void *fn) {
    static paraminfo_t by_func, by_kernel;
    static int looked;
    if (!looked) {
        by_func = (paraminfo_t)cuda_sym("cuFuncGetParamInfo");
        by_kernel = (paraminfo_t)cuda_sym("cuKernelGetParamInfo");
        looked = 1;
    }
    pscan_t *e = &g_pscan[((uintptr_t)fn >> 4) & 1023];
    if (e->fn == fn) return e;
    e->fn = fn;
    e->n = -1;  /* unknown layout: the caller falls back to "every range" */
    int no_params = 0;
    const int was = tl_bypass;
    tl_bypass = 1;
    for (int pass = 0; pass < 2; ++pass) {  /* a CUfunction first, then the same handle as a CUkernel */
        paraminfo_t q = pass ? by_kernel : by_func;
        if (!q) continue;
        int n = 0;
        size_t off = 0, size = 0;
        CUresult r;
        while (n < PSCAN_MAX && (r = q(fn, (size_t)n, &off, &size)) == 0) e->size[n++] = (unsigned short)(size > 32764 ? 32764 : size);  /* CUDA 12.1+: up to 32 764 B of arguments */
        if (n > 0) {
            if (n < PSCAN_MAX || q(fn, (size_t)n, &off, &size) != 0) e->n = n;  /* more parameters than the table holds: "every range" */
            break;
        }
        no_params |= r == 1;  /* CUDA_ERROR_INVALID_VALUE at index 0: a kernel that takes nothing */
    }
    if (e->n < 0 && no_params) e->n = 0;
    tl_bypass = was;
    return e;
}

/* the ranges one call names.  4096 entries hold every word of the largest argument buffer CUDA allows (32 764 B): a
 * multi-tensor kernel passes a hundred pointers and more in one struct.  One call at a time fills it (g_res_mu). */
#define NEED_MAX 4096
typedef struct { alloc_node_t **a; int n; } need_t;
static alloc_node_t *g_need[NEED_MAX];
static void need_word(need_t *nd, uint64_t w, uint64_t epoch) {
    alloc_node_t *a = vmm_find((CUdeviceptr)w);
    if (!a || a->epoch == epoch) return;
    a->epoch = epoch;  /* from here on not a victim of this call */
    if (nd->n < NEED_MAX) nd->a[nd->n++] = a;
}
static int need_one(alloc_node_t *a, uint64_t epoch) {
    if (!a->windowed) return vmm_make_resident(a, 0, node_chunks(a), epoch);
    a->last_use = ++g_res_clock;
    return 0;
}
static int need_commit(need_t *nd, uint64_t epoch) {
    int rc = 0;
    for (int i = 0; i < nd->n && rc == 0; ++i) rc = need_one(nd->a[i], epoch);
    return rc;
}
static int vmm_need_all(uint64_t epoch) {  /* a graph launch: its kernels' arguments are not ours to see */
    int rc = 0;
    for (alloc_node_t *a = g_head; a; a = a->next)
        if (a->vmem) a->epoch = epoch;  /* none of them is a victim of this call */
    for (alloc_node_t *a = g_head; a && rc == 0; a = a->next)
        if (a->vmem) rc = need_one(a, epoch);
    return rc;
}
/* g_res_mu held; takes g_alloc_mu.  -> 0, or nonzero when the ranges the launch names cannot all be resident */
static int vmm_scan_launch(void *fn, void **params, void **extra) {
    need_t nd = {.a = g_need, .n = 0};
    pthread_mutex_lock(&g_alloc_mu);
    const uint64_t epoch = ++g_res_epoch;
    int rc = 0;
    const unsigned char *buf = NULL;
    size_t buf_n = 0;
    for (void **x = extra; x && *x; x += 2) {  /* CU_LAUNCH_PARAM_BUFFER_POINTER = 1, _SIZE = 2, END = 0 */
        if (*x == (void *)1) buf = (const unsigned char *)x[1];
        else if (*x == (void *)2) buf_n = *(size_t *)x[1];
    }
    if (buf && buf_n) {
        for (size_t o = 0; o + 8 <= buf_n; o += 8) {
            uint64_t w;
            memcpy(&w, buf + o, 8);
            need_word(&nd, w, epoch);
        }
    } else if (params) {
        const pscan_t *ps = kernel_params(fn);
        if (ps->n < 0) {
            rc = vmm_need_all(epoch);  /* no parameter layout for this handle: be safe */
        } else {
            for (int i = 0; i < ps->n; ++i)
                for (size_t o = 0; o + 8 <= ps->size[i]; o += 8) {
                    uint64_t w;
                    memcpy(&w, (const unsigned char *)params[i] + o, 8);
                    need_word(&nd, w, epoch);
                }
        }
    }
    if (rc == 0) rc = need_commit(&nd, epoch);
    pthread_mutex_unlock(&g_alloc_mu);
    return rc;
}
/* a pointer query (attributes, address range): the allocation that contains ptr (*len = 0: none), its chunk made
 * resident when the process is paged so that the driver describes mapped device memory.  -> the mode for vmm_query_end */
static int vmm_query_begin(CUdeviceptr ptr, CUdeviceptr *base, size_t *len) {
    *len = 0;
    if (!g_backend_vmm) return 0;
    ensure_initialized();
    const int m = res_begin();
    pthread_mutex_lock(&g_alloc_mu);
    alloc_node_t *a = vmm_find(ptr);
    if (a) {
        *base = a->addr;
        *len = a->len;
        if (m == 2) (void)vmm_make_resident(a, (ptr - a->addr) / IPCSIZE, 1, ++g_res_epoch);
    }
    pthread_mutex_unlock(&g_alloc_mu);
    return m;
}
static void vmm_query_end(int mode) { res_end(mode); }
/* the bytes [p, p+n) if they lie in a managed range (g_res_mu and g_alloc_mu held) */
static int vmm_touch(CUdeviceptr p, size_t n, uint64_t epoch) {
    alloc_node_t *a = vmm_find(p);
    if (!a) return 0;
    const uint64_t first = (p - a->addr) / IPCSIZE;
    uint64_t last = (p + (n ? n : 1) - 1 - a->addr) / IPCSIZE;
    if (last >= node_chunks(a)) last = node_chunks(a) - 1;
    return vmm_make_resident(a, first, last - first + 1, epoch);
}
static int vmm_touch2(CUdeviceptr p, CUdeviceptr q, size_t n, uint64_t epoch) {
    const int r = vmm_touch(p, n, epoch);
    return r != 0 ? r : vmm_touch(q, n, epoch);
}
/* descriptor-based copies: the whole range around a device pointer (pitches and heights are not worth the parsing) */
static int vmm_touch_whole(CUdeviceptr p, uint64_t epoch) {
    alloc_node_t *a = p ? vmm_find(p) : NULL;
    if (!a || a->windowed) return 0;
    return vmm_make_resident(a, 0, node_chunks(a), epoch);
}
static int vmm_touch_desc(const void *desc, size_t src_type, size_t src_dev, size_t dst_type, size_t dst_dev, uint64_t epoch) {
    const unsigned char *d = (const unsigned char *)desc;
    int rc = 0;
    if (!d) return 0;
    unsigned st, dt;
    CUdeviceptr sp, dp;
    memcpy(&st, d + src_type, 4);
    memcpy(&dt, d + dst_type, 4);
    memcpy(&sp, d + src_dev, 8);
    memcpy(&dp, d + dst_dev, 8);
    if (st == 2 /* CU_MEMORYTYPE_DEVICE */ || st == 4 /* UNIFIED */) rc = vmm_touch_whole(sp, epoch);
    if (rc == 0 && (dt == 2 || dt == 4)) rc = vmm_touch_whole(dp, epoch);
    return rc;
}
#define TOUCH_2D(p) vmm_touch_desc(p, 16, 32, 72, 88, ep_)    /* CUDA_MEMCPY2D: srcMemoryType, srcDevice, dstMemoryType, dstDevice */
#define TOUCH_3D(p) vmm_touch_desc(p, 32, 48, 120, 136, ep_)  /* CUDA_MEMCPY3D (and _PEER: same leading layout) */

/* cuVGPUOversubStats: {chunks spilled to the peer tier, to the host tier, restored, checksum failures, resident now,
 * managed in total, paged mode} summed over this process's managed ranges — for tests, the validator and the monitor */
typedef struct {
    uint64_t chunks_total, chunks_resident, chunks_spilled, resident_limit, tier_slots, evictions, fetches, checksum_failures;
    double evict_ms, fetch_ms;
    uint64_t evictions_peer, evictions_host, restore_retries;
    double tier_setup_ms;
} shim_vmem_stats_t;
EXPORT int cuVGPUOversubStats(uint64_t out[8]) {
    typedef int (*stats_t)(void *, shim_vmem_stats_t *);
    static stats_t st;
    memset(out, 0, 8 * sizeof(uint64_t));
    if (!g_backend_vmm) return -1;
    if (!st) st = (stats_t)real_dlsym(g_dp.lib, "vgpu_vmem_stats");
    if (!st) return -1;
    pthread_mutex_lock(&g_res_mu);
    pthread_mutex_lock(&g_alloc_mu);
    for (alloc_node_t *a = g_head; a; a = a->next)
        if (a->vmem) {
            shim_vmem_stats_t v;
            if (st(a->vmem, &v) != 0) continue;
            out[0] += v.evictions_peer;
            out[1] += v.evictions_host;
            out[2] += v.fetches;
            out[3] += v.checksum_failures;
            out[4] += v.chunks_resident;
            out[5] += v.chunks_total;
        }
    out[6] = (uint64_t)g_vmm_paged;
    pthread_mutex_unlock(&g_alloc_mu);
    pthread_mutex_unlock(&g_res_mu);
    return 0;
}

/* cuVGPUResidentBytes: how much of the managed range around `p` is backed by HBM right now (-1: not a managed range) */
EXPORT long long cuVGPUResidentBytes(CUdeviceptr p) {
    if (!g_backend_vmm) return -1;
    pthread_mutex_lock(&g_res_mu);
    pthread_mutex_lock(&g_alloc_mu);
    alloc_node_t *a = vmm_find(p);
    tl_bypass = 1;
    const long long r = a ? (long long)(node_resident(a) * IPCSIZE) : -1;
    tl_bypass = 0;
    pthread_mutex_unlock(&g_alloc_mu);
    pthread_mutex_unlock(&g_res_mu);
    return r;
}

/* cuLaunchKernel@0x37c2b: the suspend gate stays; rate_limiter@0x463b1 is gone — the SM cap is
 * spatial (green context), so there is nothing to do per launch. */
/* What is left of rate_limiter@0x463b1 at every launch once the host pid is known (pidfound): the monitor's lever
 * and its activity mark.  recent_kernel < 0 in the cache file holds every launch of the vGPU (the monitor parks
 * low-priority tasks that way), and a launch writes recent_kernel = 2 ("was active a moment ago"; the monitor counts
 * it down).  The token bucket that followed is what the green context replaces. */
static void launch_mark(void) {
    if (!g_hostpid_known || !g_region_ok) return;
    vgpu_region_t *r = region_get();
    while (*(volatile int32_t *)&r->recent_kernel < 0) sleep(1);
    if (r->recent_kernel != 2) r->recent_kernel = 2;  /* no write when it is already there: the line is shared by every process */
}

/* One body for the launch entry points and their per-thread-default-stream twins: gate, activity mark, residency of
 * what the arguments name (paged mode only), then the driver's function of the same name. */
/* HBM that other tenants of the GPU hold is not ours to take: when the ranges a call needs do not fit and nothing of
 * our own is left to spill, the call steps out of the gate (so that this process can itself be suspended meanwhile),
 * waits and tries again — the node's monitor is expected to make room (monitor.Scheduler) — for VGPU_OVERSUB_WAIT_S
 * seconds (default 60) before it fails with CUDA_ERROR_OUT_OF_MEMORY. */
static int res_wait_retry(int *tries) {
    static int max_tries = -1;
    if (max_tries < 0) {
        const char *e = getenv("VGPU_OVERSUB_WAIT_S");
        max_tries = (e && *e ? atoi(e) : 60) * 10;
    }
    if (*tries >= max_tries) return 0;
    if ((*tries)++ == 0) SHIM_WARN("waiting for HBM held by other tenants of this GPU (up to %d s)", max_tries / 10);
    usleep(100000);
    return 1;
}

#define LAUNCH_BODY(SYM, CALL, FN, PARAMS, EXTRA)                      \
    if (tl_bypass) return f ? f CALL : 3;                              \
    ensure_initialized();                                              \
    launch_mark();                                                     \
    for (int tries_ = 0;;) {                                           \
        gate_enter();                                                  \
        const int m_ = res_begin();                                    \
        CUresult rc_ = 0;                                              \
        const int r_ = m_ == 2 ? vmm_scan_launch(FN, PARAMS, EXTRA) : 0; \
        if (r_ == 0) rc_ = f ? f CALL : 3;                             \
        res_end(m_);                                                   \
        gate_exit();                                                   \
        if (r_ == -3 /* VGPU_E_NOMEM */ && res_wait_retry(&tries_)) continue; \
        return r_ != 0 ? 2 : rc_;                                      \
    }

#define LAUNCH_HOOK(name)                                                                                              \
    EXPORT CUresult name(void *fn, unsigned gx, unsigned gy, unsigned gz, unsigned bx, unsigned by, unsigned bz,       \
                         unsigned shm, void *stream, void **params, void **extra) {                                    \
        static cuLaunchKernel_t f;                                                                                     \
        if (!f) f = (cuLaunchKernel_t)cuda_sym(#name);                                                                 \
        LAUNCH_BODY(name, (fn, gx, gy, gz, bx, by, bz, shm, stream, params, extra), fn, params, extra)                 \
    }
LAUNCH_HOOK(cuLaunchKernel)
LAUNCH_HOOK(cuLaunchKernel_ptsz)

#define LAUNCH_EX_HOOK(name)                                                                      \
    EXPORT CUresult name(const void *config, void *fn, void **params, void **extra) {             \
        typedef CUresult (*ex_t)(const void *, void *, void **, void **);                         \
        static ex_t f;                                                                            \
        if (!f) f = (ex_t)cuda_sym(#name);                                                        \
        LAUNCH_BODY(name, (config, fn, params, extra), fn, params, extra)                         \
    }
LAUNCH_EX_HOOK(cuLaunchKernelEx)
LAUNCH_EX_HOOK(cuLaunchKernelEx_ptsz)

#define LAUNCH_COOP_HOOK(name)                                                                                          \
    EXPORT CUresult name(void *fn, unsigned gx, unsigned gy, unsigned gz, unsigned bx, unsigned by, unsigned bz,        \
                         unsigned shm, void *st, void **params) {                                                       \
        typedef CUresult (*co_t)(void *, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, void *, void **); \
        static co_t f;                                                                                                  \
        if (!f) f = (co_t)cuda_sym(#name);                                                                              \
        LAUNCH_BODY(name, (fn, gx, gy, gz, bx, by, bz, shm, st, params), fn, params, NULL)                              \
    }
LAUNCH_COOP_HOOK(cuLaunchCooperativeKernel)
LAUNCH_COOP_HOOK(cuLaunchCooperativeKernel_ptsz)

/* The suspend gate on the copy / memset entry points, as in the reference (cuda/memory.c runs
 * wait_status_self in every cuMemcpy / cuMemset hook): while a vGPU is suspended — and, with the VMM
 * backend, until its data is back behind its pointers — nothing may touch device memory. */
/* GATED2(name, twin suffix, parameters, arguments, TOUCH): the entry point and its per-thread-default-stream twin
 * (name##_ptds for the synchronous forms, name##_ptsz for the stream-ordered ones; the driver exports both and hands
 * the twin out when cuGetProcAddress is asked with CU_GET_PROC_ADDRESS_PER_THREAD_DEFAULT_STREAM, or when a program is
 * built with --default-stream per-thread and links it directly).  Both wait at the gate, count themselves in flight,
 * and — in paged mode — make the device bytes they name resident first (TOUCH, an expression over vmm_touch / TOUCH_2D
 * / TOUCH_3D with ep_ as the epoch; 0 = nothing to touch). */
#define GATED_ONE(fname, sym, params, args, TOUCH)                    \
    EXPORT CUresult fname params {                                    \
        typedef CUresult(*fn_t) params;                               \
        static fn_t f;                                                \
        if (!f) f = (fn_t)cuda_sym(sym);                              \
        if (tl_bypass) return f ? f args : 3;                         \
        ensure_initialized();                                         \
        for (int tries_ = 0;;) {                                      \
            gate_enter();                                             \
            const int m_ = res_begin();                               \
            CUresult rc_ = 0;                                         \
            int r_ = 0;                                               \
            if (m_ == 2) {                                            \
                pthread_mutex_lock(&g_alloc_mu);                      \
                const uint64_t ep_ = ++g_res_epoch;                   \
                (void)ep_;                                            \
                r_ = (TOUCH);                                         \
                pthread_mutex_unlock(&g_alloc_mu);                    \
            }                                                         \
            if (r_ == 0) rc_ = f ? f args : 3;                        \
            res_end(m_);                                              \
            gate_exit();                                              \
            if (r_ == -3 /* VGPU_E_NOMEM */ && res_wait_retry(&tries_)) continue; \
            return r_ != 0 ? 2 : rc_;                                 \
        }                                                             \
    }
#define GATED2(name, sfx, params, args, TOUCH)        \
    GATED_ONE(name, #name, params, args, TOUCH)       \
    GATED_ONE(name##sfx, #name #sfx, params, args, TOUCH)
#define T1(p, n) vmm_touch(p, n, ep_)
#define T2(p, q, n) vmm_touch2(p, q, n, ep_)
GATED2(cuMemcpy, _ptds, (CUdeviceptr d, CUdeviceptr s_, size_t n), (d, s_, n), T2(d, s_, n))
GATED2(cuMemcpyAsync, _ptsz, (CUdeviceptr d, CUdeviceptr s_, size_t n, void *st), (d, s_, n, st), T2(d, s_, n))
GATED2(cuMemcpyDtoH_v2, _ptds, (void *d, CUdeviceptr s_, size_t n), (d, s_, n), T1(s_, n))
GATED2(cuMemcpyDtoHAsync_v2, _ptsz, (void *d, CUdeviceptr s_, size_t n, void *st), (d, s_, n, st), T1(s_, n))
GATED2(cuMemcpyHtoD_v2, _ptds, (CUdeviceptr d, const void *s_, size_t n), (d, s_, n), T1(d, n))
GATED2(cuMemcpyHtoDAsync_v2, _ptsz, (CUdeviceptr d, const void *s_, size_t n, void *st), (d, s_, n, st), T1(d, n))
GATED2(cuMemcpyDtoD_v2, _ptds, (CUdeviceptr d, CUdeviceptr s_, size_t n), (d, s_, n), T2(d, s_, n))
GATED2(cuMemcpyDtoDAsync_v2, _ptsz, (CUdeviceptr d, CUdeviceptr s_, size_t n, void *st), (d, s_, n, st), T2(d, s_, n))
GATED2(cuMemsetD8_v2, _ptds, (CUdeviceptr d, unsigned char v, size_t n), (d, v, n), T1(d, n))
GATED2(cuMemsetD8Async, _ptsz, (CUdeviceptr d, unsigned char v, size_t n, void *st), (d, v, n, st), T1(d, n))
GATED2(cuMemsetD32_v2, _ptds, (CUdeviceptr d, unsigned v, size_t n), (d, v, n), T1(d, 4 * n))
GATED2(cuMemsetD32Async, _ptsz, (CUdeviceptr d, unsigned v, size_t n, void *st), (d, v, n, st), T1(d, 4 * n))
/* the rest of the family: descriptor-based, peer, array, 16-bit and 2-D forms, graph launches */
GATED2(cuMemcpy2D_v2, _ptds, (const void *p), (p), TOUCH_2D(p))
GATED2(cuMemcpy2DUnaligned_v2, _ptds, (const void *p), (p), TOUCH_2D(p))
GATED2(cuMemcpy2DAsync_v2, _ptsz, (const void *p, void *st), (p, st), TOUCH_2D(p))
GATED2(cuMemcpy3D_v2, _ptds, (const void *p), (p), TOUCH_3D(p))
GATED2(cuMemcpy3DAsync_v2, _ptsz, (const void *p, void *st), (p, st), TOUCH_3D(p))
GATED2(cuMemcpy3DPeer, _ptds, (const void *p), (p), TOUCH_3D(p))
GATED2(cuMemcpy3DPeerAsync, _ptsz, (const void *p, void *st), (p, st), TOUCH_3D(p))
GATED2(cuMemcpyPeer, _ptds, (CUdeviceptr d, CUcontext dc, CUdeviceptr s_, CUcontext sc, size_t n), (d, dc, s_, sc, n), T2(d, s_, n))
GATED2(cuMemcpyPeerAsync, _ptsz, (CUdeviceptr d, CUcontext dc, CUdeviceptr s_, CUcontext sc, size_t n, void *st), (d, dc, s_, sc, n, st), T2(d, s_, n))
/* device <-> CUDA array: exported and gated by the reference as well (nm -D vgpu/libvgpu.so) */
GATED2(cuMemcpyAtoD_v2, _ptds, (CUdeviceptr d, void *arr, size_t off, size_t n), (d, arr, off, n), T1(d, n))
GATED2(cuMemcpyDtoA_v2, _ptds, (void *arr, size_t off, CUdeviceptr s_, size_t n), (arr, off, s_, n), T1(s_, n))
GATED2(cuMemsetD16_v2, _ptds, (CUdeviceptr d, unsigned short v, size_t n), (d, v, n), T1(d, 2 * n))
GATED2(cuMemsetD16Async, _ptsz, (CUdeviceptr d, unsigned short v, size_t n, void *st), (d, v, n, st), T1(d, 2 * n))
GATED2(cuMemsetD2D8_v2, _ptds, (CUdeviceptr d, size_t pitch, unsigned char v, size_t w, size_t h), (d, pitch, v, w, h), T1(d, pitch * h))
GATED2(cuMemsetD2D8Async, _ptsz, (CUdeviceptr d, size_t pitch, unsigned char v, size_t w, size_t h, void *st), (d, pitch, v, w, h, st), T1(d, pitch * h))
GATED2(cuMemsetD2D16_v2, _ptds, (CUdeviceptr d, size_t pitch, unsigned short v, size_t w, size_t h), (d, pitch, v, w, h), T1(d, pitch * h))
GATED2(cuMemsetD2D16Async, _ptsz, (CUdeviceptr d, size_t pitch, unsigned short v, size_t w, size_t h, void *st), (d, pitch, v, w, h, st), T1(d, pitch * h))
GATED2(cuMemsetD2D32_v2, _ptds, (CUdeviceptr d, size_t pitch, unsigned v, size_t w, size_t h), (d, pitch, v, w, h), T1(d, pitch * h))
GATED2(cuMemsetD2D32Async, _ptsz, (CUdeviceptr d, size_t pitch, unsigned v, size_t w, size_t h, void *st), (d, pitch, v, w, h, st), T1(d, pitch * h))
GATED2(cuGraphLaunch, _ptsz, (void *graph, void *st), (graph, st), vmm_need_all(ep_))
EXPORT CUresult cuMemcpyAtoD(CUdeviceptr d, void *arr, size_t off, size_t n) { return cuMemcpyAtoD_v2(d, arr, off, n); }
EXPORT CUresult cuMemcpyDtoA(void *arr, size_t off, CUdeviceptr s_, size_t n) { return cuMemcpyDtoA_v2(arr, off, s_, n); }

/* cuMipmappedArrayCreate (a20): same allocate-then-check_oom pattern as the host allocations */
EXPORT CUresult cuMipmappedArrayCreate(void **handle, const void *desc, unsigned levels) {
    typedef CUresult (*fn_t)(void **, const void *, unsigned);
    typedef CUresult (*del_t)(void *);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMipmappedArrayCreate");
    if (tl_bypass) return f ? f(handle, desc, levels) : 3;
    ensure_initialized();
    CUresult rc = f ? f(handle, desc, levels) : 3;
    if (rc != 0) return rc;
    if (oom_check(-1, 0)) {
        del_t d = (del_t)cuda_sym("cuMipmappedArrayDestroy");
        if (d) d(*handle);
        return 2;
    }
    return 0;
}

/* ------------------------------------------------------------------ NVML (a23) */
typedef struct { unsigned long long total, free, used; } nvmlMemory_t;
typedef struct { unsigned version; unsigned long long total, reserved, free, used; } nvmlMemory_v2_t;

static int nvml_dev_index(void *handle) {
    typedef int (*idx_t)(void *, unsigned *);
    idx_t f = (idx_t)nvml_sym("nvmlDeviceGetIndex");
    unsigned i = 0;
    if (!f || f(handle, &i) != 0) return -1;
    return (int)i;
}

/* NVML's side of the map.  Frameworks count GPUs through NVML before they touch CUDA (PyTorch's device_count
 * does), so NVML's count and index -> handle are narrowed to the GPUs the map names, in physical order, built from
 * NVML's own UUIDs because cuInit may not have happened yet.  NVML keeps PHYSICAL devices: a GPU named twice is one
 * NVML device whose memory info is the sum of its vGPUs — which is also what the CUDA runtime sees (merged device,
 * above).  (The reference makes up virtual NVML handles and wraps all 247 NVML entry points to translate them.) */
static int g_nv_phys[VGPU_MAX_DEVICES], g_nv_n, g_nv_remap;
static int g_nv_col0[VGPU_MAX_DEVICES], g_nv_ncols[VGPU_MAX_DEVICES];  /* the cache-file columns of each listed GPU's vGPUs */
static pthread_once_t g_nv_once = PTHREAD_ONCE_INIT;
static void build_nvml_view(void) {
    typedef int (*count_t)(unsigned *);
    typedef int (*byidx_t)(unsigned, void **);
    typedef int (*uuid_t)(void *, char *, unsigned);
    count_t cnt = (count_t)nvml_sym("nvmlDeviceGetCount_v2");
    byidx_t byidx = (byidx_t)nvml_sym("nvmlDeviceGetHandleByIndex_v2");
    uuid_t uuid = (uuid_t)nvml_sym("nvmlDeviceGetUUID");
    char want[VGPU_MAX_DEVICES][VGPU_UUID_LEN];
    const int nwant = device_map_uuids(want, VGPU_MAX_DEVICES);
    unsigned n = 0;
    if (nwant == 0 || !cnt || !byidx || !uuid || cnt(&n) != 0) return;
    int col = 0;
    for (unsigned i = 0; i < n && i < VGPU_MAX_DEVICES; ++i) {
        void *h;
        char u[VGPU_UUID_LEN];
        if (byidx(i, &h) != 0 || uuid(h, u, sizeof(u)) != 0) continue;
        int named = 0;
        for (int k = 0; k < nwant; ++k) named += strcasecmp(want[k], u) == 0;
        if (named && g_nv_n < VGPU_MAX_DEVICES) {  /* a GPU named twice is still one NVML device, with two columns */
            g_nv_phys[g_nv_n] = (int)i;
            g_nv_col0[g_nv_n] = col;
            g_nv_ncols[g_nv_n++] = named;
            col += named;
        }
    }
    g_nv_remap = g_nv_n > 0 && g_nv_n != (int)n;
    for (int i = 0; i < g_nv_n; ++i) g_nv_remap |= g_nv_phys[i] != i || g_nv_ncols[i] != 1;
}

EXPORT int nvmlDeviceGetCount_v2(unsigned *count) {
    typedef int (*fn_t)(unsigned *);
    fn_t f = (fn_t)nvml_sym("nvmlDeviceGetCount_v2");
    int rc = f ? f(count) : 3;
    if (rc != 0) return rc;
    pthread_once(&g_nv_once, build_nvml_view);
    if (g_nv_remap) *count = (unsigned)g_nv_n;
    return 0;
}
EXPORT int nvmlDeviceGetCount(unsigned *count) { return nvmlDeviceGetCount_v2(count); }

EXPORT int nvmlDeviceGetHandleByIndex_v2(unsigned index, void **handle) {
    typedef int (*fn_t)(unsigned, void **);
    fn_t f = (fn_t)nvml_sym("nvmlDeviceGetHandleByIndex_v2");
    if (!f) return 3;
    pthread_once(&g_nv_once, build_nvml_view);
    if (g_nv_remap) {
        if (index >= (unsigned)g_nv_n) return 2;  /* NVML_ERROR_INVALID_ARGUMENT */
        index = (unsigned)g_nv_phys[index];
    }
    return f(index, handle);
}
EXPORT int nvmlDeviceGetHandleByIndex(unsigned index, void **handle) { return nvmlDeviceGetHandleByIndex_v2(index, handle); }

static int env_switch_on(const char *name, int when_unset);
EXPORT uint64_t get_gpu_memory_monitor(int dev);
/* usage / limit as NVML should show them for a physical GPU (by its NVML index, whatever the CUDA side enumerates):
 * the sums over the cache-file columns of this container's vGPUs on it; any unlimited vGPU makes it unlimited */
static void nvml_quota(int nvml_index, uint64_t *usage, uint64_t *limit) {
    pthread_once(&g_nv_once, build_nvml_view);
    int col0 = nvml_index, ncols = 1;
    if (g_nv_remap) {
        ncols = 0;  /* a GPU the map does not name: no quota of ours applies */
        for (int i = 0; i < g_nv_n; ++i)
            if (g_nv_phys[i] == nvml_index) {
                col0 = g_nv_col0[i];
                ncols = g_nv_ncols[i];
            }
    }
    int unlimited = ncols == 0;
    uint64_t monitored = 0;
    *usage = *limit = 0;
    for (int c = col0; c < col0 + ncols; ++c) {
        *usage += region_usage(c);
        *limit += region_limit(c);
        unlimited |= region_limit(c) == 0;
        monitored += get_gpu_memory_monitor(c);
    }
    if (unlimited) *limit = 0;
    /* nvmlDeviceGetMemoryInfo@0x25052 with MEMORY_OVERRIDE (read @0x2546a): what the outside monitor measured wins when it is more */
    if (env_switch_on("MEMORY_OVERRIDE", 0) && monitored > *usage) *usage = monitored;
}

EXPORT int nvmlDeviceGetMemoryInfo(void *handle, nvmlMemory_t *m) {
    ensure_initialized();
    typedef int (*fn_t)(void *, nvmlMemory_t *);
    fn_t f = (fn_t)nvml_sym("nvmlDeviceGetMemoryInfo");
    int rc = f ? f(handle, m) : 3;
    if (rc != 0 || !g_region_ok) return rc;
    int dev = nvml_dev_index(handle);
    if (dev < 0) return rc;
    uint64_t usage, limit;
    nvml_quota(dev, &usage, &limit);
    if (limit == 0) {
        m->used = usage;  /* hook.c:328-334: only `used` is overwritten when unlimited */
    } else {
        m->total = limit;
        m->free = limit - usage;
        m->used = usage;
    }
    return 0;
}

EXPORT int nvmlDeviceGetMemoryInfo_v2(void *handle, nvmlMemory_v2_t *m) {
    ensure_initialized();
    typedef int (*fn_t)(void *, nvmlMemory_v2_t *);
    fn_t f = (fn_t)nvml_sym("nvmlDeviceGetMemoryInfo_v2");
    int rc = f ? f(handle, m) : 3;
    if (rc != 0 || !g_region_ok) return rc;
    int dev = nvml_dev_index(handle);
    if (dev < 0) return rc;
    uint64_t usage, limit;
    nvml_quota(dev, &usage, &limit);
    if (limit == 0) {
        m->used = usage;
    } else {
        m->total = limit;
        m->free = limit - usage;
        m->used = usage;
    }
    return 0;
}

/* ------------------------------------------------------------------ virtual devices: the entry points
 * Everything in the driver API that takes or returns a device ordinal.  With an identity map each of
 * these is one predictable branch in front of the driver's function. */
/* the runtime API announces itself (see merge_shared) */
EXPORT CUresult cuGetExportTable(const void **table, const void *id) {
    typedef CUresult (*fn_t)(const void **, const void *);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuGetExportTable");
    if (!tl_bypass && !g_runtime_client) {
        /* the driver's own libraries fetch tables through the same (interposable) symbol — libcuda during cuInit,
         * libnvidia-ml during nvmlInit (seen on 580.x): only a caller outside the driver stack counts */
        Dl_info di;
        const int known = dladdr(__builtin_return_address(0), &di) && di.dli_fname;
        if (!known || !(strstr(di.dli_fname, "libcuda.so") || strstr(di.dli_fname, "libnvidia-"))) {
            g_runtime_client = 1;
            SHIM_INFO("cuGetExportTable from %s: CUDA-runtime client", known ? di.dli_fname : "?");
        }
    }
    return f ? f(table, id) : 3;
}

EXPORT CUresult cuDeviceGetCount(int *count) {
    cuDeviceGetCount_t f = get_cuDeviceGetCount();
    CUresult rc = f ? f(count) : 3;
    if (rc != 0 || tl_bypass) return rc;
    ensure_vdev();
    if (g_remap) *count = g_nvdev;
    return 0;
}

EXPORT CUresult cuDeviceGet(CUdevice *dev, int ordinal) {
    typedef CUresult (*fn_t)(CUdevice *, int);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuDeviceGet");
    if (!tl_bypass) {
        ensure_vdev();
        if (g_remap) {
            if (ordinal < 0 || ordinal >= g_nvdev) return 101;
            *dev = ordinal;  /* the handle IS the virtual ordinal */
            return 0;
        }
    }
    return f ? f(dev, ordinal) : 3;
}

EXPORT CUresult cuCtxGetDevice(CUdevice *dev) {
    if (tl_bypass) return get_cuCtxGetDevice()(dev);
    return cur_dev(dev);
}

/* name(params) forwards with `d` (and `d2` when the call has two devices) translated */
#define VDEV1(name, params, args, d)                                  \
    EXPORT CUresult name params {                                     \
        typedef CUresult(*fn_t) params;                               \
        static fn_t f;                                                \
        if (!f) f = (fn_t)cuda_sym(#name);                            \
        if (!tl_bypass && !to_phys(&d)) return 101;                   \
        return f ? f args : 3;                                        \
    }
#define VDEV2(name, params, args, d, d2)                              \
    EXPORT CUresult name params {                                     \
        typedef CUresult(*fn_t) params;                               \
        static fn_t f;                                                \
        if (!f) f = (fn_t)cuda_sym(#name);                            \
        if (!tl_bypass && (!to_phys(&d) || !to_phys(&d2))) return 101; \
        return f ? f args : 3;                                        \
    }
VDEV1(cuDeviceGetName, (char *name, int len, CUdevice d), (name, len, d), d)
/* CU_DEVICE_ATTRIBUTE_MULTIPROCESSOR_COUNT (16) inside an SM partition is the partition's size: code that sizes a
 * cooperative or persistent grid as occupancy x SM count must not ask for more CTAs than the green context can
 * co-schedule (CUDA_ERROR_COOPERATIVE_LAUNCH_TOO_LARGE) — the reference's temporal throttle never had that failure. */
EXPORT CUresult cuDeviceGetAttribute(int *pi, int attrib, CUdevice d) {
    typedef CUresult (*fn_t)(int *, int, CUdevice);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuDeviceGetAttribute");
    const CUdevice v = d;
    if (!tl_bypass && !to_phys(&d)) return 101;
    CUresult rc = f ? f(pi, attrib, d) : 3;
    if (rc == 0 && !tl_bypass && attrib == 16 && pi) {
        const unsigned sms = green_planned_sms(v);
        if (sms) *pi = (int)sms;
    }
    return rc;
}
VDEV1(cuDeviceGetUuid, (unsigned char *uuid, CUdevice d), (uuid, d), d)
VDEV1(cuDeviceGetUuid_v2, (unsigned char *uuid, CUdevice d), (uuid, d), d)
VDEV1(cuDeviceGetLuid, (char *luid, unsigned *mask, CUdevice d), (luid, mask, d), d)
VDEV1(cuDeviceComputeCapability, (int *major, int *minor, CUdevice d), (major, minor, d), d)
VDEV1(cuDeviceGetProperties, (void *prop, CUdevice d), (prop, d), d)
VDEV1(cuDeviceGetTexture1DLinearMaxWidth, (size_t *w, int fmt, unsigned ch, CUdevice d), (w, fmt, ch, d), d)
VDEV1(cuDeviceGetExecAffinitySupport, (int *pi, int type, CUdevice d), (pi, type, d), d)
VDEV1(cuDeviceGetDefaultMemPool, (void **pool, CUdevice d), (pool, d), d)
VDEV1(cuDeviceGetMemPool, (void **pool, CUdevice d), (pool, d), d)
VDEV1(cuDeviceSetMemPool, (CUdevice d, void *pool), (d, pool), d)
VDEV1(cuDeviceGetGraphMemAttribute, (CUdevice d, int attr, void *v), (d, attr, v), d)
VDEV1(cuDeviceSetGraphMemAttribute, (CUdevice d, int attr, void *v), (d, attr, v), d)
VDEV1(cuDeviceGraphMemTrim, (CUdevice d), (d), d)
VDEV1(cuDeviceGetDevResource, (CUdevice d, void *res, int type), (d, res, type), d)
VDEV1(cuDevicePrimaryCtxGetState, (CUdevice d, unsigned *flags, int *active), (d, flags, active), d)
VDEV1(cuDevicePrimaryCtxSetFlags_v2, (CUdevice d, unsigned flags), (d, flags), d)
VDEV1(cuDevicePrimaryCtxReset_v2, (CUdevice d), (d), d)
VDEV1(cuGreenCtxCreate, (void **g, void *desc, CUdevice d, unsigned flags), (g, desc, d, flags), d)
VDEV2(cuDeviceCanAccessPeer, (int *can, CUdevice d, CUdevice d2), (can, d, d2), d, d2)
VDEV2(cuDeviceGetP2PAttribute, (int *v, int attr, CUdevice d, CUdevice d2), (v, attr, d, d2), d, d2)
EXPORT CUresult cuDevicePrimaryCtxSetFlags(CUdevice d, unsigned flags) { return cuDevicePrimaryCtxSetFlags_v2(d, flags); }
EXPORT CUresult cuDevicePrimaryCtxReset(CUdevice d) { return cuDevicePrimaryCtxReset_v2(d); }

/* VMM / pool entry points carry the device inside a CUmemLocation {type, id}; type 1 = DEVICE */
typedef struct { int type, id; } shim_mem_location_t;
static int loc_to_phys(shim_mem_location_t *l) { return l->type != 1 || to_phys(&l->id); }

/* cuMemPrefetchAsync / cuMemAdvise: the device argument may also be CU_DEVICE_CPU (-1) */
/* With the VMM oversubscribe backend the application's "managed" allocations are explicit ranges of the data path, and
 * the prefetch calls a UVM program already makes become its paging controls: prefetch to the CPU = spill the allocation
 * to the tier and give its HBM back (only when the call covers the whole allocation: the manager moves whole blocks, and
 * unlike UVM nothing faults back in — the data is out until it is prefetched to the device again); prefetch to the
 * device = make the range resident again, checksum-verified.  Synchronous with respect to `stream`.
 * -> 1 handled (rc set), 0 not one of ours */
static int vmm_prefetch(CUdeviceptr p, size_t n, int to_device, void *stream, CUresult *rc) {
    (void)stream;
    if (!g_backend_vmm || tl_bypass) return 0;
    gate_enter();
    pthread_mutex_lock(&g_res_mu);
    pthread_mutex_lock(&g_alloc_mu);
    alloc_node_t *a = vmm_find(p);
    int handled = a != NULL, vrc = 0;
    if (a) {
        const uint64_t chunks = node_chunks(a);
        const uint64_t first = (p - a->addr) / IPCSIZE;
        uint64_t last = (p + (n ? n : 1) - 1 - a->addr) / IPCSIZE;
        if (last >= chunks) last = chunks - 1;
        if (to_device) {
            vrc = vmm_make_resident(a, first, last - first + 1, ++g_res_epoch);
        } else if ((p == a->addr && n >= a->len) || a->windowed) {
            /* out: the whole allocation, or — for a range the GPU cannot hold at once — the blocks the call names */
            vmm_quiesce();  /* what was queued may still use the data */
            tl_bypass = 1;
            vrc = g_dp.evict(a->vmem, a->windowed ? first : 0, a->windowed ? last - first + 1 : chunks);
            if (vrc != 0) SHIM_ERR("vmm backend prefetch: %s", g_dp.last_error());
            tl_bypass = 0;
            if (vrc == 0) vmm_go_paged();
        }
    }
    pthread_mutex_unlock(&g_alloc_mu);
    pthread_mutex_unlock(&g_res_mu);
    gate_exit();
    *rc = vrc == 0 ? 0 : 999;
    return handled;
}

EXPORT CUresult cuMemPrefetchAsync(CUdeviceptr p, size_t n, CUdevice d, void *stream) {
    typedef CUresult (*fn_t)(CUdeviceptr, size_t, CUdevice, void *);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemPrefetchAsync");
    CUresult rc;
    if (vmm_prefetch(p, n, d >= 0, stream, &rc)) return rc;
    if (!tl_bypass && d >= 0 && !to_phys(&d)) return 101;
    return f ? f(p, n, d, stream) : 3;
}
/* Advice about a managed range: there is no fault-driven migration to steer, but the preferred location is a
 * statement about who should keep its HBM under pressure — SET_PREFERRED_LOCATION (3) towards a device makes the range
 * the last candidate for spilling, towards the CPU the first; UNSET (4) takes it back.  Everything else is accepted
 * and ignored.  -> 1 when the pointer is ours */
static int vmm_advise(CUdeviceptr p, int advice, int to_device) {
    if (!g_backend_vmm || tl_bypass) return 0;
    pthread_mutex_lock(&g_alloc_mu);
    alloc_node_t *a = vmm_find(p);
    if (a && advice == 3) a->prefer = to_device ? 1 : -1;
    if (a && advice == 4) a->prefer = 0;
    pthread_mutex_unlock(&g_alloc_mu);
    return a != NULL;
}

EXPORT CUresult cuMemAdvise(CUdeviceptr p, size_t n, int advice, CUdevice d) {
    typedef CUresult (*fn_t)(CUdeviceptr, size_t, int, CUdevice);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemAdvise");
    if (vmm_advise(p, advice, d >= 0)) return 0;
    if (!tl_bypass && d >= 0 && !to_phys(&d)) return 101;
    return f ? f(p, n, advice, d) : 3;
}

EXPORT CUresult cuMemPrefetchAsync_v2(CUdeviceptr p, size_t n, shim_mem_location_t loc, unsigned flags, void *stream) {
    typedef CUresult (*fn_t)(CUdeviceptr, size_t, shim_mem_location_t, unsigned, void *);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemPrefetchAsync_v2");
    CUresult rc;
    if (vmm_prefetch(p, n, loc.type == 1 /* DEVICE; HOST / HOST_NUMA = out */, stream, &rc)) return rc;
    if (!tl_bypass && !loc_to_phys(&loc)) return 101;
    return f ? f(p, n, loc, flags, stream) : 3;
}
EXPORT CUresult cuMemAdvise_v2(CUdeviceptr p, size_t n, int advice, shim_mem_location_t loc) {
    typedef CUresult (*fn_t)(CUdeviceptr, size_t, int, shim_mem_location_t);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemAdvise_v2");
    if (vmm_advise(p, advice, loc.type == 1)) return 0;
    if (!tl_bypass && !loc_to_phys(&loc)) return 101;
    return f ? f(p, n, advice, loc) : 3;
}
/* newer context constructors: translated, not billed (the reference knows cuCtxCreate_v2 only) */
VDEV1(cuCtxCreate_v3, (CUcontext *pctx, void *params, int nparams, unsigned flags, CUdevice d), (pctx, params, nparams, flags, d), d)
VDEV1(cuCtxCreate_v4, (CUcontext *pctx, void *params, unsigned flags, CUdevice d), (pctx, params, flags, d), d)

/* The k-th further vGPU of a GPU shows PCI function .k (assigning_virtual_pcibusID@0x2c5c1: "0000:10:00.1") */
EXPORT CUresult cuDeviceGetPCIBusId(char *id, int len, CUdevice dev) {
    typedef CUresult (*fn_t)(char *, int, CUdevice);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuDeviceGetPCIBusId");
    CUdevice phys = dev;
    if (!tl_bypass && !to_phys(&phys)) return 101;
    CUresult rc = f ? f(id, len, phys) : 3;
    if (rc == 0 && !tl_bypass && g_remap && g_vdev[dev].dup > 0) {
        char *dot = strrchr(id, '.');
        if (dot && dot[1]) dot[1] = "0123456789abcdef"[g_vdev[dev].dup & 15];
    }
    return rc;
}
EXPORT CUresult cuDeviceGetByPCIBusId(CUdevice *dev, const char *id) {
    typedef CUresult (*fn_t)(CUdevice *, const char *);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuDeviceGetByPCIBusId");
    if (tl_bypass) return f ? f(dev, id) : 3;
    ensure_vdev();
    if (!g_remap) return f ? f(dev, id) : 3;
    char real[64];
    snprintf(real, sizeof(real), "%s", id ? id : "");
    char *dot = strrchr(real, '.');
    int k = 0;
    if (dot && dot[1]) {
        k = (int)strtol(dot + 1, NULL, 16);
        dot[1] = '0';
        dot[2] = 0;
    }
    CUdevice phys = -1;
    CUresult rc = f ? f(&phys, real) : 3;
    if (rc != 0) return rc;
    for (int i = 0; i < g_nvdev; ++i)
        if (g_vdev[i].phys == phys && g_vdev[i].dup == k) {
            *dev = i;
            return 0;
        }
    return 101;
}

typedef struct { int type, handle_types; shim_mem_location_t location; void *win32; unsigned char flags[8]; } shim_alloc_prop_t;
_Static_assert(sizeof(shim_alloc_prop_t) == 32, "CUmemAllocationProp ABI");
typedef struct { shim_mem_location_t location; int flags; } shim_access_desc_t;

/* ------------------------------------------------------------------ strict quota (opt-in, no reference counterpart)
 * The reference meters cuMemAlloc / Managed / Pitch only: a tenant that allocates through the stream-ordered
 * allocator (cudaMallocAsync, PyTorch backend:cudaMallocAsync) or the VMM API (cuMemCreate + cuMemMap, PyTorch
 * expandable_segments) is outside its quota altogether.  With VGPU_STRICT_QUOTA=1 those paths are billed as well:
 *   cuMemAllocAsync / cuMemAllocFromPoolAsync  checked like cuMemAllocManaged (-> 2 = OUT_OF_MEMORY when over), tracked in
 *                                              the allocation list, so cuMemFreeAsync and cuMemFree both give them back;
 *   cuMemCreate                                checked and billed to the vGPU of prop->location; the bytes come back
 *                                              when the handle has been released AND unmapped everywhere (physical
 *                                              memory outlives cuMemRelease while a mapping exists).
 * Off by default so that recorded traces decide exactly as the reference does. */
typedef struct { unsigned long long h; size_t n; int vdev, released, maps; } vmm_handle_t;
typedef struct { CUdeviceptr p; size_t n; unsigned long long h; } vmm_map_t;
static vmm_handle_t *g_vh;
static vmm_map_t *g_vm;
static size_t g_nvh, g_capvh, g_nvm, g_capvm;

/* room for one more element; 0 when the host is out of memory (the caller then refuses the allocation) */
static int grow(void **arr, size_t *cap, size_t n, size_t elem) {
    if (n < *cap) return 1;
    const size_t ncap = *cap ? 2 * *cap : 64;
    void *p = realloc(*arr, ncap * elem);
    if (!p) return 0;
    *arr = p;
    *cap = ncap;
    return 1;
}
static vmm_handle_t *vh_find(unsigned long long h) {
    for (size_t i = 0; i < g_nvh; ++i)
        if (g_vh[i].h == h) return &g_vh[i];
    return NULL;
}
static void vh_settle(vmm_handle_t *v) {  /* with g_alloc_mu held */
    if (!v->released || v->maps > 0) return;
    const int vdev = v->vdev;
    const size_t n = v->n;
    *v = g_vh[--g_nvh];
    region_rm_usage(col_of(vdev), n, VGPU_USE_DATA);
}

EXPORT CUresult cuMemCreate(unsigned long long *handle, size_t size, const shim_alloc_prop_t *prop, unsigned long long flags) {
    typedef CUresult (*fn_t)(unsigned long long *, size_t, const shim_alloc_prop_t *, unsigned long long);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemCreate");
    if (tl_bypass || !prop) return f ? f(handle, size, prop, flags) : 3;
    shim_alloc_prop_t p2 = *prop;
    const int vdev = prop->location.id;
    if (!loc_to_phys(&p2.location)) return 101;
    const int metered = g_strict && g_region_ok && prop->location.type == 1;
    if (!metered) return f ? f(handle, size, &p2, flags) : 3;
    ensure_initialized();
    wait_running();
    pthread_mutex_lock(&g_alloc_mu);  /* check, allocate and bill as one step, like cuMemAlloc */
    CUresult rc = 2;
    if (!oom_check(vdev, size) && grow((void **)&g_vh, &g_capvh, g_nvh, sizeof(*g_vh))) {
        rc = f ? f(handle, size, &p2, flags) : 3;
        if (rc == 0) {
            g_vh[g_nvh++] = (vmm_handle_t){*handle, size, vdev, 0, 0};
            region_add_usage(col_of(vdev), size, VGPU_USE_DATA);
        }
    }
    pthread_mutex_unlock(&g_alloc_mu);
    return rc;
}

EXPORT CUresult cuMemRelease(unsigned long long handle) {
    typedef CUresult (*fn_t)(unsigned long long);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemRelease");
    CUresult rc = f ? f(handle) : 3;
    if (rc == 0 && !tl_bypass && g_strict) {
        pthread_mutex_lock(&g_alloc_mu);
        vmm_handle_t *v = vh_find(handle);
        if (v) {
            v->released = 1;
            vh_settle(v);
        }
        pthread_mutex_unlock(&g_alloc_mu);
    }
    return rc;
}

EXPORT CUresult cuMemMap(CUdeviceptr ptr, size_t size, size_t offset, unsigned long long handle, unsigned long long flags) {
    typedef CUresult (*fn_t)(CUdeviceptr, size_t, size_t, unsigned long long, unsigned long long);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemMap");
    CUresult rc = f ? f(ptr, size, offset, handle, flags) : 3;
    if (rc == 0 && !tl_bypass && g_strict) {
        pthread_mutex_lock(&g_alloc_mu);
        vmm_handle_t *v = vh_find(handle);
        if (v && grow((void **)&g_vm, &g_capvm, g_nvm, sizeof(*g_vm))) {
            g_vm[g_nvm++] = (vmm_map_t){ptr, size, handle};
            ++v->maps;
        }
        pthread_mutex_unlock(&g_alloc_mu);
    }
    return rc;
}

EXPORT CUresult cuMemUnmap(CUdeviceptr ptr, size_t size) {
    typedef CUresult (*fn_t)(CUdeviceptr, size_t);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemUnmap");
    CUresult rc = f ? f(ptr, size) : 3;
    if (rc == 0 && !tl_bypass && g_strict) {
        pthread_mutex_lock(&g_alloc_mu);
        for (size_t i = 0; i < g_nvm;) {  /* the range covers whole mappings (driver rule) */
            if (g_vm[i].p >= ptr && g_vm[i].p + g_vm[i].n <= ptr + size) {
                vmm_handle_t *v = vh_find(g_vm[i].h);
                g_vm[i] = g_vm[--g_nvm];
                if (v) {
                    --v->maps;
                    vh_settle(v);
                }
            } else {
                ++i;
            }
        }
        pthread_mutex_unlock(&g_alloc_mu);
    }
    return rc;
}

/* stream-ordered allocations: legacy-stream and per-thread-stream flavours share the accounting */
static CUresult alloc_async(CUdeviceptr *dptr, size_t n, void *pool, void *stream, int from_pool, int ptsz) {
    typedef CUresult (*a_t)(CUdeviceptr *, size_t, void *);
    typedef CUresult (*p_t)(CUdeviceptr *, size_t, void *, void *);
    static void *fn[4];
    static const char *names[4] = {"cuMemAllocAsync", "cuMemAllocAsync_ptsz", "cuMemAllocFromPoolAsync", "cuMemAllocFromPoolAsync_ptsz"};
    const int k = 2 * from_pool + ptsz;
    if (!fn[k]) fn[k] = cuda_sym(names[k]);
    if (!fn[k]) return 3;
    if (tl_bypass || !g_strict) return from_pool ? ((p_t)fn[k])(dptr, n, pool, stream) : ((a_t)fn[k])(dptr, n, stream);
    ensure_initialized();
    pthread_once(&g_post_once, post_init);
    wait_running();
    pthread_mutex_lock(&g_alloc_mu);
    CUresult rc = 2;
    if (!oom_check(-1, n)) {
        rc = from_pool ? ((p_t)fn[k])(dptr, n, pool, stream) : ((a_t)fn[k])(dptr, n, stream);
        if (rc == 0) add_chunk_only(*dptr, n);
    }
    pthread_mutex_unlock(&g_alloc_mu);
    return rc;
}
EXPORT CUresult cuMemAllocAsync(CUdeviceptr *dptr, size_t n, void *stream) { return alloc_async(dptr, n, NULL, stream, 0, 0); }
EXPORT CUresult cuMemAllocAsync_ptsz(CUdeviceptr *dptr, size_t n, void *stream) { return alloc_async(dptr, n, NULL, stream, 0, 1); }
EXPORT CUresult cuMemAllocFromPoolAsync(CUdeviceptr *dptr, size_t n, void *pool, void *stream) { return alloc_async(dptr, n, pool, stream, 1, 0); }
EXPORT CUresult cuMemAllocFromPoolAsync_ptsz(CUdeviceptr *dptr, size_t n, void *pool, void *stream) { return alloc_async(dptr, n, pool, stream, 1, 1); }

static CUresult free_async(CUdeviceptr dptr, void *stream, int ptsz) {
    typedef CUresult (*fn_t)(CUdeviceptr, void *);
    static fn_t fn[2];
    if (!fn[ptsz]) fn[ptsz] = (fn_t)cuda_sym(ptsz ? "cuMemFreeAsync_ptsz" : "cuMemFreeAsync");
    if (!fn[ptsz]) return 3;
    if (!tl_bypass && g_backend_vmm && dptr) {  /* a managed range (made by cuMemAlloc) freed stream-ordered: the plain free */
        pthread_mutex_lock(&g_alloc_mu);
        alloc_node_t *a = g_head;
        while (a && a->addr != dptr) a = a->next;
        const int managed = a && a->vmem;
        pthread_mutex_unlock(&g_alloc_mu);
        if (managed) return cuMemFree_v2(dptr);
    }
    if (!tl_bypass && g_strict && dptr) {
        pthread_mutex_lock(&g_alloc_mu);
        alloc_node_t *a = g_head;
        while (a && a->addr != dptr) a = a->next;
        if (a) {
            if (a->prev) a->prev->next = a->next; else g_head = a->next;
            if (a->next) a->next->prev = a->prev; else g_tail = a->prev;
            --g_nalloc;
            region_rm_usage(col_of(a->dev), a->len, VGPU_USE_DATA);
            free(a);
        }
        pthread_mutex_unlock(&g_alloc_mu);
    }
    return fn[ptsz](dptr, stream);
}
EXPORT CUresult cuMemFreeAsync(CUdeviceptr dptr, void *stream) { return free_async(dptr, stream, 0); }
EXPORT CUresult cuMemFreeAsync_ptsz(CUdeviceptr dptr, void *stream) { return free_async(dptr, stream, 1); }

EXPORT CUresult cuMemGetAllocationGranularity(size_t *g, const shim_alloc_prop_t *prop, int option) {
    typedef CUresult (*fn_t)(size_t *, const shim_alloc_prop_t *, int);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemGetAllocationGranularity");
    shim_alloc_prop_t p2;
    if (!tl_bypass && prop) {
        p2 = *prop;
        if (!loc_to_phys(&p2.location)) return 101;
        prop = &p2;
    }
    return f ? f(g, prop, option) : 3;
}
EXPORT CUresult cuMemSetAccess(CUdeviceptr ptr, size_t size, const shim_access_desc_t *desc, size_t count) {
    typedef CUresult (*fn_t)(CUdeviceptr, size_t, const shim_access_desc_t *, size_t);
    static fn_t f;
    if (!f) f = (fn_t)cuda_sym("cuMemSetAccess");
    shim_access_desc_t d2[VGPU_MAX_DEVICES];
    ensure_vdev();
    if (!tl_bypass && g_remap && desc && count <= VGPU_MAX_DEVICES) {
        for (size_t i = 0; i < count; ++i) {
            d2[i] = desc[i];
            if (!loc_to_phys(&d2[i].location)) return 101;
        }
        desc = d2;
    }
    return f ? f(ptr, size, desc, count) : 3;
}

/* ------------------------------------------------------------------ the monitor's entry points
 * An outside monitor (the "vgpu monitor" that VGPU_MONITOR_MODE exists for) loads libvgpu.so next to a container's
 * cache file and drives it through plain exported helpers: read usage / limits, report what NVML says a process really
 * uses, resize a quota, suspend and resume the vGPU's processes.  Same names, arguments and answers as the
 * reference's (multiprocess_memory_limit.c; pinned against the reference binary in tests/test_monitor_abi.py):
 *   set_gpu_device_memory_monitor@0x42e29   first slot whose hostpid matches: monitorused[dev] = bytes; with
 *       MEMORY_OVERRIDE (set to anything but "0" / "false") the slot's total BECOMES that figure (offset = bytes - total),
 *       so allocations the hooks never saw count too; and unless ACTIVE_OOM_KILLER is "0" / "false", a process seen
 *       above 1.1 x its limit gets every process of the vGPU killed (SIGKILL; not with an unlimited vGPU, not with a
 *       GPU named twice in the map).  Always answers 1.
 *   suspend_all@0x45ddf / resume_all@0x45eea   SIGUSR2 / SIGUSR1 to every process in the file. */
static int env_switch_on(const char *name, int when_unset) {
    const char *v = getenv(name);
    if (!v) return when_unset;
    return !(strcmp(v, "0") == 0 || strcmp(v, "false") == 0);
}

EXPORT uint64_t get_gpu_memory_usage(int dev) { ensure_initialized(); return region_usage(dev); }
EXPORT uint64_t get_current_device_memory_usage(int dev) { return get_gpu_memory_usage(dev); }
EXPORT uint64_t get_current_device_memory_limit(int dev) { ensure_initialized(); return region_limit(dev); }
EXPORT int set_current_device_memory_limit(int dev, uint64_t bytes) {
    ensure_initialized();
    if (dev >= 0 && dev < VGPU_MAX_DEVICES) region_get()->limit[dev] = bytes;
    return 0;
}
EXPORT int get_current_device_sm_limit(int dev) { ensure_initialized(); return (int)region_sm_limit(dev); }
EXPORT int set_current_device_sm_limit_scale(int dev, int scale) {
    /* @0x457a2: multiplies (30 x 50 = 1500), and only the first call per cache file does anything (sm_init_flag) */
    ensure_initialized();
    vgpu_region_t *r = region_get();
    if (r->sm_init_flag == 1 || dev < 0 || dev >= VGPU_MAX_DEVICES) return 0;
    r->sm_limit[dev] *= (uint64_t)scale;
    r->sm_init_flag = 1;
    return 0;
}
EXPORT int get_current_priority(void) { ensure_initialized(); return region_get()->priority; }
EXPORT int get_recent_kernel(void) { ensure_initialized(); return region_get()->recent_kernel; }
EXPORT int set_recent_kernel(int v) { ensure_initialized(); region_get()->recent_kernel = v; return 0; }
EXPORT int get_utilization_switch(void) { ensure_initialized(); return region_get()->utilization_switch; }

EXPORT uint64_t get_gpu_memory_monitor(int dev) {
    ensure_initialized();
    if (dev < 0 || dev >= VGPU_MAX_DEVICES) return 0;
    vgpu_region_t *r = region_get();
    uint64_t t = 0;
    region_lock();
    for (int i = 0; i < r->proc_num; ++i) t += r->procs[i].monitorused[dev];
    region_unlock();
    return t;
}
EXPORT uint64_t get_current_device_memory_monitor(int dev) { return get_gpu_memory_monitor(dev); }

EXPORT void *find_proc_by_hostpid(int hostpid) {
    ensure_initialized();
    vgpu_region_t *r = region_get();
    for (int i = 0; i < r->proc_num; ++i)
        if (r->procs[i].hostpid == hostpid) return &r->procs[i];
    return NULL;
}

EXPORT int set_gpu_device_memory_monitor(int hostpid, int dev, uint64_t bytes) {
    ensure_initialized();
    if (dev < 0 || dev >= VGPU_MAX_DEVICES) return 1;
    vgpu_region_t *r = region_get();
    int kill_all = 0;
    region_lock();
    for (int i = 0; i < r->proc_num; ++i) {
        vgpu_proc_slot_t *s = &r->procs[i];
        if (s->hostpid != hostpid) continue;
        s->monitorused[dev] = bytes;
        if (env_switch_on("MEMORY_OVERRIDE", 0)) {  /* what NVML sees replaces what the hooks counted */
            s->used[dev].offset = bytes - s->used[dev].total;
            s->used[dev].total = bytes;
        }
        const uint64_t limit = r->limit[dev];
        if (!g_map_has_duplicates && limit != 0 && (double)bytes > (double)limit * 1.1 && env_switch_on("ACTIVE_OOM_KILLER", 1)) {
            SHIM_ERR("hostpid %d uses %lu of %lu on device %d: killing the vGPU's processes", hostpid, (unsigned long)bytes,
                     (unsigned long)limit, dev);
            kill_all = 1;
        }
        break;
    }
    if (kill_all)  /* active_oom_killer@0x4297d */
        for (int i = 0; i < r->proc_num; ++i) kill(r->procs[i].pid, SIGKILL);
    region_unlock();
    return 1;
}

static int signal_all(int sig) {
    ensure_initialized();
    vgpu_region_t *r = region_get();
    int n = 0;
    region_lock();
    for (int i = 0; i < r->proc_num; ++i, ++n) kill(r->procs[i].pid, sig);
    region_unlock();
    return n;
}
EXPORT int suspend_all(void) { return signal_all(SIGUSR2); }
EXPORT int resume_all(void) { return signal_all(SIGUSR1); }

EXPORT int set_current_gpu_status(int status) {
    ensure_initialized();
    region_set_my_status(status);
    return status;
}
EXPORT int wait_status_self(int status) {
    ensure_initialized();
    while (region_my_status() != status) usleep(1000);
    return 1;
}
EXPORT int wait_status_all(int status) {
    ensure_initialized();
    vgpu_region_t *r = region_get();
    for (;;) {
        int all = 1;
        region_lock();
        for (int i = 0; i < r->proc_num; ++i) all &= r->procs[i].status == status;
        region_unlock();
        if (all) return 1;
        usleep(1000);
    }
}

/* ------------------------------------------------------------------ dispatch: dlsym / cuGetProcAddress */
typedef struct { const char *name; void *fn; } hook_t;
EXPORT CUresult cuGetProcAddress(const char *symbol, void **pfn, int ver, unsigned long long flags);
EXPORT CUresult cuGetProcAddress_v2(const char *symbol, void **pfn, int ver, unsigned long long flags, int *status);

static const hook_t g_hooks[] = {
    {"cuInit", (void *)cuInit},
    {"cuMemAlloc", (void *)cuMemAlloc_v2}, {"cuMemAlloc_v2", (void *)cuMemAlloc_v2},
    {"cuMemAllocManaged", (void *)cuMemAllocManaged},
    {"cuMemAllocPitch", (void *)cuMemAllocPitch_v2}, {"cuMemAllocPitch_v2", (void *)cuMemAllocPitch_v2},
    {"cuMemFree", (void *)cuMemFree_v2}, {"cuMemFree_v2", (void *)cuMemFree_v2},
    {"cuMemGetInfo", (void *)cuMemGetInfo_v2}, {"cuMemGetInfo_v2", (void *)cuMemGetInfo_v2},
    {"cuDeviceTotalMem", (void *)cuDeviceTotalMem_v2}, {"cuDeviceTotalMem_v2", (void *)cuDeviceTotalMem_v2},
    {"cuMemAllocHost", (void *)cuMemAllocHost_v2}, {"cuMemAllocHost_v2", (void *)cuMemAllocHost_v2},
    {"cuMemHostAlloc", (void *)cuMemHostAlloc},
    {"cuMemHostRegister", (void *)cuMemHostRegister_v2}, {"cuMemHostRegister_v2", (void *)cuMemHostRegister_v2},
    {"cuDevicePrimaryCtxRetain", (void *)cuDevicePrimaryCtxRetain},
    {"cuDevicePrimaryCtxRelease", (void *)cuDevicePrimaryCtxRelease_v2}, {"cuDevicePrimaryCtxRelease_v2", (void *)cuDevicePrimaryCtxRelease_v2},
    {"cuCtxCreate", (void *)cuCtxCreate_v2}, {"cuCtxCreate_v2", (void *)cuCtxCreate_v2},
    {"cuCtxSetCurrent", (void *)cuCtxSetCurrent},
    {"cuCtxGetCurrent", (void *)cuCtxGetCurrent},
    {"cuCtxPopCurrent", (void *)cuCtxPopCurrent_v2}, {"cuCtxPopCurrent_v2", (void *)cuCtxPopCurrent_v2},
    {"cuCtxPushCurrent", (void *)cuCtxPushCurrent_v2}, {"cuCtxPushCurrent_v2", (void *)cuCtxPushCurrent_v2},
    {"cuPointerGetAttributes", (void *)cuPointerGetAttributes}, {"cuPointerGetAttribute", (void *)cuPointerGetAttribute},
    {"cuMemGetAddressRange", (void *)cuMemGetAddressRange_v2}, {"cuMemGetAddressRange_v2", (void *)cuMemGetAddressRange_v2},
    {"cuMipmappedArrayCreate", (void *)cuMipmappedArrayCreate},
    {"cuMemcpy", (void *)cuMemcpy}, {"cuMemcpyAsync", (void *)cuMemcpyAsync},
    {"cuMemcpyDtoH", (void *)cuMemcpyDtoH_v2}, {"cuMemcpyDtoH_v2", (void *)cuMemcpyDtoH_v2},
    {"cuMemcpyDtoHAsync", (void *)cuMemcpyDtoHAsync_v2}, {"cuMemcpyDtoHAsync_v2", (void *)cuMemcpyDtoHAsync_v2},
    {"cuMemcpyHtoD", (void *)cuMemcpyHtoD_v2}, {"cuMemcpyHtoD_v2", (void *)cuMemcpyHtoD_v2},
    {"cuMemcpyHtoDAsync", (void *)cuMemcpyHtoDAsync_v2}, {"cuMemcpyHtoDAsync_v2", (void *)cuMemcpyHtoDAsync_v2},
    {"cuMemcpyDtoD", (void *)cuMemcpyDtoD_v2}, {"cuMemcpyDtoD_v2", (void *)cuMemcpyDtoD_v2},
    {"cuMemcpyDtoDAsync", (void *)cuMemcpyDtoDAsync_v2}, {"cuMemcpyDtoDAsync_v2", (void *)cuMemcpyDtoDAsync_v2},
    {"cuMemsetD8", (void *)cuMemsetD8_v2}, {"cuMemsetD8_v2", (void *)cuMemsetD8_v2}, {"cuMemsetD8Async", (void *)cuMemsetD8Async},
    {"cuMemsetD32", (void *)cuMemsetD32_v2}, {"cuMemsetD32_v2", (void *)cuMemsetD32_v2}, {"cuMemsetD32Async", (void *)cuMemsetD32Async},
    {"cuMemcpy2D", (void *)cuMemcpy2D_v2}, {"cuMemcpy2D_v2", (void *)cuMemcpy2D_v2},
    {"cuMemcpy2DUnaligned", (void *)cuMemcpy2DUnaligned_v2}, {"cuMemcpy2DUnaligned_v2", (void *)cuMemcpy2DUnaligned_v2},
    {"cuMemcpy2DAsync", (void *)cuMemcpy2DAsync_v2}, {"cuMemcpy2DAsync_v2", (void *)cuMemcpy2DAsync_v2},
    {"cuMemcpy3D", (void *)cuMemcpy3D_v2}, {"cuMemcpy3D_v2", (void *)cuMemcpy3D_v2},
    {"cuMemcpy3DAsync", (void *)cuMemcpy3DAsync_v2}, {"cuMemcpy3DAsync_v2", (void *)cuMemcpy3DAsync_v2},
    {"cuMemcpy3DPeer", (void *)cuMemcpy3DPeer}, {"cuMemcpy3DPeerAsync", (void *)cuMemcpy3DPeerAsync},
    {"cuMemcpyPeer", (void *)cuMemcpyPeer}, {"cuMemcpyPeerAsync", (void *)cuMemcpyPeerAsync},
    {"cuMemsetD16", (void *)cuMemsetD16_v2}, {"cuMemsetD16_v2", (void *)cuMemsetD16_v2}, {"cuMemsetD16Async", (void *)cuMemsetD16Async},
    {"cuMemsetD2D8", (void *)cuMemsetD2D8_v2}, {"cuMemsetD2D8_v2", (void *)cuMemsetD2D8_v2}, {"cuMemsetD2D8Async", (void *)cuMemsetD2D8Async},
    {"cuMemsetD2D16", (void *)cuMemsetD2D16_v2}, {"cuMemsetD2D16_v2", (void *)cuMemsetD2D16_v2}, {"cuMemsetD2D16Async", (void *)cuMemsetD2D16Async},
    {"cuMemsetD2D32", (void *)cuMemsetD2D32_v2}, {"cuMemsetD2D32_v2", (void *)cuMemsetD2D32_v2}, {"cuMemsetD2D32Async", (void *)cuMemsetD2D32Async},
    {"cuLaunchCooperativeKernel", (void *)cuLaunchCooperativeKernel}, {"cuGraphLaunch", (void *)cuGraphLaunch},
    {"cuLaunchKernel", (void *)cuLaunchKernel},
    {"cuLaunchKernelEx", (void *)cuLaunchKernelEx},
    {"cuMemcpyAtoD", (void *)cuMemcpyAtoD_v2}, {"cuMemcpyAtoD_v2", (void *)cuMemcpyAtoD_v2},
    {"cuMemcpyDtoA", (void *)cuMemcpyDtoA_v2}, {"cuMemcpyDtoA_v2", (void *)cuMemcpyDtoA_v2},
    /* per-thread-default-stream twins (looked up as <name>_ptsz / <name>_ptds, see find_hook_for_version) */
    {"cuLaunchKernel_ptsz", (void *)cuLaunchKernel_ptsz}, {"cuLaunchKernelEx_ptsz", (void *)cuLaunchKernelEx_ptsz},
    {"cuLaunchCooperativeKernel_ptsz", (void *)cuLaunchCooperativeKernel_ptsz}, {"cuGraphLaunch_ptsz", (void *)cuGraphLaunch_ptsz},
    {"cuMemcpy_ptds", (void *)cuMemcpy_ptds}, {"cuMemcpyAsync_ptsz", (void *)cuMemcpyAsync_ptsz},
    {"cuMemcpyDtoH_ptds", (void *)cuMemcpyDtoH_v2_ptds}, {"cuMemcpyDtoH_v2_ptds", (void *)cuMemcpyDtoH_v2_ptds},
    {"cuMemcpyDtoHAsync_ptsz", (void *)cuMemcpyDtoHAsync_v2_ptsz}, {"cuMemcpyDtoHAsync_v2_ptsz", (void *)cuMemcpyDtoHAsync_v2_ptsz},
    {"cuMemcpyHtoD_ptds", (void *)cuMemcpyHtoD_v2_ptds}, {"cuMemcpyHtoD_v2_ptds", (void *)cuMemcpyHtoD_v2_ptds},
    {"cuMemcpyHtoDAsync_ptsz", (void *)cuMemcpyHtoDAsync_v2_ptsz}, {"cuMemcpyHtoDAsync_v2_ptsz", (void *)cuMemcpyHtoDAsync_v2_ptsz},
    {"cuMemcpyDtoD_ptds", (void *)cuMemcpyDtoD_v2_ptds}, {"cuMemcpyDtoD_v2_ptds", (void *)cuMemcpyDtoD_v2_ptds},
    {"cuMemcpyDtoDAsync_ptsz", (void *)cuMemcpyDtoDAsync_v2_ptsz}, {"cuMemcpyDtoDAsync_v2_ptsz", (void *)cuMemcpyDtoDAsync_v2_ptsz},
    {"cuMemcpyAtoD_ptds", (void *)cuMemcpyAtoD_v2_ptds}, {"cuMemcpyAtoD_v2_ptds", (void *)cuMemcpyAtoD_v2_ptds},
    {"cuMemcpyDtoA_ptds", (void *)cuMemcpyDtoA_v2_ptds}, {"cuMemcpyDtoA_v2_ptds", (void *)cuMemcpyDtoA_v2_ptds},
    {"cuMemsetD8_ptds", (void *)cuMemsetD8_v2_ptds}, {"cuMemsetD8_v2_ptds", (void *)cuMemsetD8_v2_ptds}, {"cuMemsetD8Async_ptsz", (void *)cuMemsetD8Async_ptsz},
    {"cuMemsetD16_ptds", (void *)cuMemsetD16_v2_ptds}, {"cuMemsetD16_v2_ptds", (void *)cuMemsetD16_v2_ptds}, {"cuMemsetD16Async_ptsz", (void *)cuMemsetD16Async_ptsz},
    {"cuMemsetD32_ptds", (void *)cuMemsetD32_v2_ptds}, {"cuMemsetD32_v2_ptds", (void *)cuMemsetD32_v2_ptds}, {"cuMemsetD32Async_ptsz", (void *)cuMemsetD32Async_ptsz},
    {"cuMemsetD2D8_ptds", (void *)cuMemsetD2D8_v2_ptds}, {"cuMemsetD2D8_v2_ptds", (void *)cuMemsetD2D8_v2_ptds}, {"cuMemsetD2D8Async_ptsz", (void *)cuMemsetD2D8Async_ptsz},
    {"cuMemsetD2D16_ptds", (void *)cuMemsetD2D16_v2_ptds}, {"cuMemsetD2D16_v2_ptds", (void *)cuMemsetD2D16_v2_ptds}, {"cuMemsetD2D16Async_ptsz", (void *)cuMemsetD2D16Async_ptsz},
    {"cuMemsetD2D32_ptds", (void *)cuMemsetD2D32_v2_ptds}, {"cuMemsetD2D32_v2_ptds", (void *)cuMemsetD2D32_v2_ptds}, {"cuMemsetD2D32Async_ptsz", (void *)cuMemsetD2D32Async_ptsz},
    {"cuMemcpy2D_ptds", (void *)cuMemcpy2D_v2_ptds}, {"cuMemcpy2D_v2_ptds", (void *)cuMemcpy2D_v2_ptds},
    {"cuMemcpy2DUnaligned_ptds", (void *)cuMemcpy2DUnaligned_v2_ptds}, {"cuMemcpy2DUnaligned_v2_ptds", (void *)cuMemcpy2DUnaligned_v2_ptds},
    {"cuMemcpy2DAsync_ptsz", (void *)cuMemcpy2DAsync_v2_ptsz}, {"cuMemcpy2DAsync_v2_ptsz", (void *)cuMemcpy2DAsync_v2_ptsz},
    {"cuMemcpy3D_ptds", (void *)cuMemcpy3D_v2_ptds}, {"cuMemcpy3D_v2_ptds", (void *)cuMemcpy3D_v2_ptds},
    {"cuMemcpy3DAsync_ptsz", (void *)cuMemcpy3DAsync_v2_ptsz}, {"cuMemcpy3DAsync_v2_ptsz", (void *)cuMemcpy3DAsync_v2_ptsz},
    {"cuMemcpy3DPeer_ptds", (void *)cuMemcpy3DPeer_ptds}, {"cuMemcpy3DPeerAsync_ptsz", (void *)cuMemcpy3DPeerAsync_ptsz},
    {"cuMemcpyPeer_ptds", (void *)cuMemcpyPeer_ptds}, {"cuMemcpyPeerAsync_ptsz", (void *)cuMemcpyPeerAsync_ptsz},
    {"cuGetExportTable", (void *)cuGetExportTable},
    {"cuDeviceGetCount", (void *)cuDeviceGetCount}, {"cuDeviceGet", (void *)cuDeviceGet}, {"cuCtxGetDevice", (void *)cuCtxGetDevice},
    {"cuDeviceGetName", (void *)cuDeviceGetName}, {"cuDeviceGetAttribute", (void *)cuDeviceGetAttribute},
    {"cuDeviceGetUuid", (void *)cuDeviceGetUuid}, {"cuDeviceGetUuid_v2", (void *)cuDeviceGetUuid_v2},
    {"cuDeviceGetLuid", (void *)cuDeviceGetLuid}, {"cuDeviceComputeCapability", (void *)cuDeviceComputeCapability},
    {"cuDeviceGetProperties", (void *)cuDeviceGetProperties},
    {"cuDeviceGetTexture1DLinearMaxWidth", (void *)cuDeviceGetTexture1DLinearMaxWidth},
    {"cuDeviceGetExecAffinitySupport", (void *)cuDeviceGetExecAffinitySupport},
    {"cuDeviceGetDefaultMemPool", (void *)cuDeviceGetDefaultMemPool}, {"cuDeviceGetMemPool", (void *)cuDeviceGetMemPool},
    {"cuDeviceSetMemPool", (void *)cuDeviceSetMemPool},
    {"cuDeviceGetGraphMemAttribute", (void *)cuDeviceGetGraphMemAttribute},
    {"cuDeviceSetGraphMemAttribute", (void *)cuDeviceSetGraphMemAttribute}, {"cuDeviceGraphMemTrim", (void *)cuDeviceGraphMemTrim},
    {"cuDeviceGetDevResource", (void *)cuDeviceGetDevResource}, {"cuGreenCtxCreate", (void *)cuGreenCtxCreate},
    {"cuDevicePrimaryCtxGetState", (void *)cuDevicePrimaryCtxGetState},
    {"cuDevicePrimaryCtxSetFlags", (void *)cuDevicePrimaryCtxSetFlags_v2}, {"cuDevicePrimaryCtxSetFlags_v2", (void *)cuDevicePrimaryCtxSetFlags_v2},
    {"cuDevicePrimaryCtxReset", (void *)cuDevicePrimaryCtxReset_v2}, {"cuDevicePrimaryCtxReset_v2", (void *)cuDevicePrimaryCtxReset_v2},
    {"cuDeviceCanAccessPeer", (void *)cuDeviceCanAccessPeer}, {"cuDeviceGetP2PAttribute", (void *)cuDeviceGetP2PAttribute},
    {"cuMemPrefetchAsync", (void *)cuMemPrefetchAsync}, {"cuMemAdvise", (void *)cuMemAdvise},
    {"cuMemPrefetchAsync_v2", (void *)cuMemPrefetchAsync_v2}, {"cuMemAdvise_v2", (void *)cuMemAdvise_v2},
    {"cuCtxCreate_v3", (void *)cuCtxCreate_v3}, {"cuCtxCreate_v4", (void *)cuCtxCreate_v4},
    {"cuDeviceGetPCIBusId", (void *)cuDeviceGetPCIBusId}, {"cuDeviceGetByPCIBusId", (void *)cuDeviceGetByPCIBusId},
    {"cuMemCreate", (void *)cuMemCreate}, {"cuMemGetAllocationGranularity", (void *)cuMemGetAllocationGranularity},
    {"cuMemRelease", (void *)cuMemRelease}, {"cuMemMap", (void *)cuMemMap}, {"cuMemUnmap", (void *)cuMemUnmap},
    {"cuMemAllocAsync", (void *)cuMemAllocAsync}, {"cuMemAllocFromPoolAsync", (void *)cuMemAllocFromPoolAsync},
    {"cuMemFreeAsync", (void *)cuMemFreeAsync},
    {"cuMemAllocAsync_ptsz", (void *)cuMemAllocAsync_ptsz}, {"cuMemAllocFromPoolAsync_ptsz", (void *)cuMemAllocFromPoolAsync_ptsz},
    {"cuMemFreeAsync_ptsz", (void *)cuMemFreeAsync_ptsz},
    {"cuMemSetAccess", (void *)cuMemSetAccess},
    {"cuGetProcAddress", (void *)cuGetProcAddress}, {"cuGetProcAddress_v2", (void *)cuGetProcAddress_v2},
    {"nvmlDeviceGetCount", (void *)nvmlDeviceGetCount_v2}, {"nvmlDeviceGetCount_v2", (void *)nvmlDeviceGetCount_v2},
    {"nvmlDeviceGetHandleByIndex", (void *)nvmlDeviceGetHandleByIndex_v2}, {"nvmlDeviceGetHandleByIndex_v2", (void *)nvmlDeviceGetHandleByIndex_v2},
    {"nvmlDeviceGetMemoryInfo", (void *)nvmlDeviceGetMemoryInfo},
    {"nvmlDeviceGetMemoryInfo_v2", (void *)nvmlDeviceGetMemoryInfo_v2},
};

static void *find_hook(const char *name) {
    for (size_t i = 0; i < sizeof(g_hooks) / sizeof(g_hooks[0]); ++i)
        if (strcmp(g_hooks[i].name, name) == 0) return g_hooks[i].fn;
    return NULL;
}

/* cuGetProcAddress(symbol, cudaVersion): an unsuffixed name means the newest ABI that version knows, and for a
 * few of the hooked names that ABI has a different signature than the hook registered under the bare name */
static void *find_hook_for_version(const char *name, int ver, unsigned long long flags) {
    if (flags & 2ull) {  /* per-thread default stream: the stream-ordered allocator hooks have their own _ptsz twins */
        static const char *const ptsz[] = {"cuMemAllocAsync", "cuMemAllocFromPoolAsync", "cuMemFreeAsync"};
        for (size_t i = 0; i < sizeof(ptsz) / sizeof(ptsz[0]); ++i)
            if (strcmp(name, ptsz[i]) == 0) {
                char twin[64];
                snprintf(twin, sizeof(twin), "%s_ptsz", name);
                return find_hook(twin);
            }
    }
    static const struct { const char *name; int since; const char *use; } newer[] = {
        {"cuGetProcAddress", 12000, "cuGetProcAddress_v2"},
        {"cuDeviceGetUuid", 11040, "cuDeviceGetUuid_v2"},
        {"cuCtxCreate", 12050, "cuCtxCreate_v4"},
        {"cuCtxCreate", 11040, "cuCtxCreate_v3"},
        {"cuMemPrefetchAsync", 12020, "cuMemPrefetchAsync_v2"},
        {"cuMemAdvise", 12020, "cuMemAdvise_v2"},
    };
    for (size_t i = 0; i < sizeof(newer) / sizeof(newer[0]); ++i)
        if (ver >= newer[i].since && strcmp(name, newer[i].name) == 0) return find_hook(newer[i].use);
    return find_hook(name);
}

/* Launch / copy / memset / graph-launch entry points have per-thread-default-stream twins (_ptsz / _ptds) that the
 * driver selects from the flags of cuGetProcAddress.  A caller that asks for the per-thread flavour gets OUR twin of
 * the same name, which gates and forwards to the driver's twin (its stream semantics are the driver's). */
static void *find_ptds_twin(const char *symbol, unsigned long long flags) {
    if (!(flags & 2ull /* CU_GET_PROC_ADDRESS_PER_THREAD_DEFAULT_STREAM */)) return NULL;
    char twin[96];
    snprintf(twin, sizeof(twin), "%s_ptsz", symbol);
    void *h = find_hook(twin);
    if (!h) {
        snprintf(twin, sizeof(twin), "%s_ptds", symbol);
        h = find_hook(twin);
    }
    return h;
}
/* the few per-thread forms we do not wrap go to the driver untouched (never to a legacy-stream hook) */
static int wants_ptds(const char *symbol, unsigned long long flags) {
    return (flags & 2ull) &&
           (strncmp(symbol, "cuMemcpy", 8) == 0 || strncmp(symbol, "cuMemset", 8) == 0 || strncmp(symbol, "cuLaunch", 8) == 0 ||
            strncmp(symbol, "cuMemPrefetch", 13) == 0 || strncmp(symbol, "cuGraphLaunch", 13) == 0);
}

/* dlsym@0x12bb6: every cu, nvml lookup that names a hooked symbol lands in the shim */
EXPORT void *dlsym(void *handle, const char *symbol) {
    find_real_dlsym();
    if ((symbol[0] == 'c' && symbol[1] == 'u') || strncmp(symbol, "nvml", 4) == 0) {
        void *h = find_hook(symbol);
        if (h) return h;
    }
    return real_dlsym(handle, symbol);
}

EXPORT CUresult cuGetProcAddress(const char *symbol, void **pfn, int ver, unsigned long long flags) {
    ensure_initialized();
    void *h = symbol ? find_ptds_twin(symbol, flags) : NULL;
    if (!h) h = (symbol && !wants_ptds(symbol, flags)) ? find_hook_for_version(symbol, ver, flags) : NULL;
    if (shim_log_level() > 4) fprintf(stderr, "[4pdvGPU Trace]: cuGetProcAddress(%s, ver=%d) -> %s\n", symbol, ver, h ? "hook" : "driver");
    if (h) {
        *pfn = h;
        return 0;
    }
    cuGetProcAddress_t f = get_cuGetProcAddress();
    return f ? f(symbol, pfn, ver, flags) : 500;
}

EXPORT CUresult cuGetProcAddress_v2(const char *symbol, void **pfn, int ver, unsigned long long flags, int *status) {
    ensure_initialized();
    void *h = symbol ? find_ptds_twin(symbol, flags) : NULL;
    if (!h) h = (symbol && !wants_ptds(symbol, flags)) ? find_hook_for_version(symbol, ver, flags) : NULL;
    if (shim_log_level() > 4) fprintf(stderr, "[4pdvGPU Trace]: cuGetProcAddress_v2(%s, ver=%d) -> %s\n", symbol, ver, h ? "hook" : "driver");
    if (h) {
        *pfn = h;
        if (status) *status = 0;  /* CU_GET_PROC_ADDRESS_SUCCESS */
        return 0;
    }
    cuGetProcAddress_v2_t f = get_cuGetProcAddress_v2();
    return f ? f(symbol, pfn, ver, flags, status) : 500;
}

/* Suspend / resume of a whole vGPU (the reference's dormant suspend_all@0x45ddf / resume_all
 * @0x45eea only flip the status word; here they also move the data).  Suspend: drain the GPU,
 * spill every VMM-backed allocation to the host tier and release its HBM.  Resume: map fresh
 * HBM behind the same addresses, restore, verify checksums.  Returns chunks-level errors as
 * CUDA_ERROR_UNKNOWN (999). */
static CUresult move_all(int out) {
    CUresult rc = 0;
    pthread_mutex_lock(&g_res_mu);
    pthread_mutex_lock(&g_alloc_mu);
    tl_bypass = 1;
    if (out) vmm_quiesce();  /* every context this process owns, not only the one current here */
    for (alloc_node_t *a = g_head; a; a = a->next)
        if (a->vmem && !(a->windowed && !out)) {  /* a windowed range comes back by its owner's prefetch hints */
            const uint64_t chunks = (a->len + IPCSIZE - 1) / IPCSIZE;
            int vrc = out ? g_dp.evict(a->vmem, 0, chunks) : vmm_make_resident(a, 0, chunks, ++g_res_epoch);
            tl_bypass = 1;
            if (vrc != 0) {
                SHIM_ERR("vmm backend %s: %s", out ? "suspend" : "resume", g_dp.last_error());
                rc = 999;
            }
        }
    tl_bypass = 0;
    pthread_mutex_unlock(&g_alloc_mu);
    pthread_mutex_unlock(&g_res_mu);
    return rc;
}
static uint64_t managed_bytes(void) {
    uint64_t n = 0;
    pthread_mutex_lock(&g_alloc_mu);
    for (alloc_node_t *a = g_head; a; a = a->next)
        if (a->vmem && !a->windowed) n += node_chunks(a) * IPCSIZE;
    pthread_mutex_unlock(&g_alloc_mu);
    return n;
}
/* the in-process form: the caller's own threads are its business, but the gate still closes around the move */
static CUresult move_all_gated(int out) {
    if (!out) return move_all(0);
    g_data_out = 1;
    while (g_inflight > 0) usleep(100);
    CUresult rc = move_all(1);
    g_data_out = 0;
    return rc;
}
EXPORT CUresult cuVGPUSuspend(void) { return g_backend_vmm ? move_all_gated(1) : 801 /* NOT_SUPPORTED */; }
EXPORT CUresult cuVGPUResume(void) { return g_backend_vmm ? move_all_gated(0) : 801; }

/* Signal-driven form of the same (the reference's protocol: an external monitor sends SIGUSR2 to
 * suspend a vGPU process and SIGUSR1 to resume it, sig_swap_stub / sig_restore_stub).  The
 * handlers only flip the status word and post a semaphore; this thread does the work.  The gate closes
 * the moment the status turns 2 (g_data_out); the mover then waits until no call that passed the gate
 * earlier is still on its way into the driver (g_inflight == 0) before it drains the GPU and unmaps
 * anything, and opens the gate again only when the data is back. */
#include <semaphore.h>
static sem_t g_mover_sem;
static volatile int g_mover_target = 1;
static void mover_wake(int status) {
    g_mover_target = status;
    if (status == 2) g_data_out = 1;  /* close the gate before anything moves */
    sem_post(&g_mover_sem);
}
static void *mover_main(void *arg) {
    (void)arg;
    int bound = 0, out = 0;
    for (;;) {
        while (sem_wait(&g_mover_sem) != 0) {}
        if (!bound) {  /* work in the application's primary context of device 0 */
            CUcontext ctx = NULL;
            cuDevicePrimaryCtxRetain_t r = get_cuDevicePrimaryCtxRetain();
            cuCtxSetCurrent_t set = get_cuCtxSetCurrent();
            if (r && set && r(&ctx, phys_of(0)) == 0 && set(ctx) == 0) bound = 1;
        }
        const int target = g_mover_target;
        if (target == 2 && !out) {
            while (g_inflight > 0) usleep(100);  /* calls that were past the gate have reached the driver and returned */
            struct timespec t0, t1;
            clock_gettime(CLOCK_MONOTONIC, &t0);
            move_all(1);
            clock_gettime(CLOCK_MONOTONIC, &t1);
            SHIM_INFO("suspend: %lu MiB of managed ranges out in %.1f ms", (unsigned long)(managed_bytes() >> 20),
                      (t1.tv_sec - t0.tv_sec) * 1e3 + (t1.tv_nsec - t0.tv_nsec) / 1e6);
            out = 1;
        } else if (target == 1 && out) {
            /* The gate opens only when every range is back.  HBM that the monitor counted as free may still be on
             * its way out of another tenant: keep trying (the gate stays closed, the application waits) unless we are
             * told to suspend again meanwhile — then whatever came back goes out again. */
            int warned = 0;
            struct timespec t0, t1;
            clock_gettime(CLOCK_MONOTONIC, &t0);
            while (move_all(0) != 0 && g_mover_target == 1) {
                if (!warned++) SHIM_WARN("resume: not all ranges fit yet, waiting for HBM");
                usleep(200000);
                clock_gettime(CLOCK_MONOTONIC, &t0);  /* time the attempt that succeeds */
            }
            clock_gettime(CLOCK_MONOTONIC, &t1);
            if (g_mover_target == 1) {
                SHIM_INFO("resume: %lu MiB of managed ranges back in %.1f ms", (unsigned long)(managed_bytes() >> 20),
                          (t1.tv_sec - t0.tv_sec) * 1e3 + (t1.tv_nsec - t0.tv_nsec) / 1e6);
                out = 0;
                g_data_out = 0;
            } else {
                move_all(1);
            }
        } else if (target == 1) {
            g_data_out = 0;
        }
    }
    return NULL;
}
static void start_mover(void) {
    pthread_t th;
    sem_init(&g_mover_sem, 0, 0);
    if (pthread_create(&th, NULL, mover_main, NULL) == 0) {
        pthread_detach(th);
        region_status_hook = mover_wake;
    }
}

/* The reference also exports the raw allocation route that sits UNDER its quota: cuMemoryAllocate
 * @0x32146 switches on the allocation mode — 0: cuMemAllocManaged(ATTACH_GLOBAL), 2: cuMemAlloc, 1: a VMM branch whose
 * helper (getallochandle) is an undefined symbol in the shipped binary — stores the size in *allocated and returns
 * the driver's code; it neither checks nor bills (add_chunk@0x40b73 does that around it).  cuMemoryFree@0x3849b is
 * its counterpart.  Same here; mode 1 is the manager's range when the VMM backend is loaded.  cuMapMemory@0x31b28 /
 * cuUnmapMemory@0x37f8f cannot run in the reference (they end in the undefined helper): NOT_SUPPORTED. */
EXPORT CUresult cuMemoryAllocate(CUdeviceptr *dptr, size_t bytesize, size_t *allocated, void *data) {
    (void)data;
    ensure_initialized();
    pthread_once(&g_post_once, post_init);
    if (!dptr) return 1;
    if (allocated) *allocated = bytesize;
    if (g_allocmode == 0 && g_backend_vmm && bytesize > IPCSIZE) {
        int dev = 0;
        if (cur_dev(&dev) != 0) return 201;
        pthread_mutex_lock(&g_res_mu);
        pthread_mutex_lock(&g_alloc_mu);
        const int rc = vmm_alloc(dev, bytesize, dptr);
        if (rc == 0) g_tail->raw = 1;  /* tracked for paging and for cuMemoryFree, never billed */
        pthread_mutex_unlock(&g_alloc_mu);
        pthread_mutex_unlock(&g_res_mu);
        return rc == 0 ? 0 : 2;
    }
    if (g_allocmode == 0) return get_cuMemAllocManaged() ? real_cuMemAllocManaged(dptr, bytesize, 1) : 3;
    return get_cuMemAlloc_v2() ? real_cuMemAlloc_v2(dptr, bytesize) : 3;
}
EXPORT CUresult cuMemoryFree(CUdeviceptr dptr) {
    ensure_initialized();
    if (g_backend_vmm) {
        pthread_mutex_lock(&g_res_mu);
        pthread_mutex_lock(&g_alloc_mu);
        alloc_node_t *a = g_head;
        while (a && !(a->raw && a->addr == dptr)) a = a->next;
        if (a) {
            vmm_quiesce();
            tl_bypass = 1;
            const int rc = g_dp.destroy(a->vmem);
            tl_bypass = 0;
            if (g_peer_left != ~0ull) g_peer_left += a->peer_bytes;
            if (a->prev) a->prev->next = a->next; else g_head = a->next;
            if (a->next) a->next->prev = a->prev; else g_tail = a->prev;
            --g_nalloc;
            idx_del(a);
            free(a);
            pthread_mutex_unlock(&g_alloc_mu);
            pthread_mutex_unlock(&g_res_mu);
            return rc == 0 ? 0 : 999;
        }
        pthread_mutex_unlock(&g_alloc_mu);
        pthread_mutex_unlock(&g_res_mu);
    }
    return get_cuMemFree_v2() ? real_cuMemFree_v2(dptr) : 3;
}
EXPORT CUresult cuMapMemory(CUdeviceptr *dptr, size_t bytesize, void *handle, void *data) {
    (void)dptr; (void)bytesize; (void)handle; (void)data;
    SHIM_ERR("cuMapMemory: the reference's VMM branch is not callable (its helper is an undefined symbol); use cuMemoryAllocate");
    return 801;
}
EXPORT CUresult cuUnmapMemory(CUdeviceptr dptr, int flag) {
    (void)dptr; (void)flag;
    return 801;
}

/* cuVGPUViewAllocator@0x16c80: dump the tracked allocations */
EXPORT void cuVGPUViewAllocator(void) {
    pthread_mutex_lock(&g_alloc_mu);
    fprintf(stderr, "[4pdvGPU] %zu tracked allocations\n", g_nalloc);
    for (alloc_node_t *a = g_head; a; a = a->next)
        fprintf(stderr, "  dev %d  0x%llx  %zu bytes\n", a->dev, a->addr, a->len);
    pthread_mutex_unlock(&g_alloc_mu);
}
