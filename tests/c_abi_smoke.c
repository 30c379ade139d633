/* A plain-C caller of include/vgpu.h — what cgo generates bindings for.  Checks that the header is
 * valid C, that the library links from C, and the Contract A known answers of SURVEY.md §8 c.1.
 * Built and run by tests/test_capi_symbols.py (no GPU needed: only the Contract A entry points). */
#include <stdio.h>
#include <string.h>

#include "vgpu.h"

#define CHECK(cond)                                                     \
    do {                                                                \
        if (!(cond)) {                                                  \
            printf("FAIL line %d: %s (%s)\n", __LINE__, #cond, vgpu_last_error()); \
            return 1;                                                   \
        }                                                               \
    } while (0)

static const char *env(const vgpu_kv_t *e, size_t n, const char *k) {
    for (size_t i = 0; i < n; ++i)
        if (strcmp(e[i].key, k) == 0) return e[i].val;
    return NULL;
}

int main(void) {
    const char *uuid = "GPU-00000000-0000-0000-0000-000000000001";
    vgpu_desc_t all[10];
    CHECK(vgpu_abi_version() == VGPU_ABI_VERSION);
    CHECK(vgpu_split(uuid, 184320, 10, 1.0, all, 10) == 10);
    CHECK(strcmp(all[3].id, "GPU-00000000-0000-0000-0000-000000000001-3") == 0 && all[3].memory_mib == 18432);
    CHECK(vgpu_sm_limit(10, 1.0) == 10 && vgpu_sm_limit(3, 1.0) == 33 && vgpu_sm_limit(4, 1.0) == 25);

    const char *ids[2] = {"GPU-00000000-0000-0000-0000-000000000001-3", "GPU-00000000-0000-0000-0000-000000000001-7"};
    const vgpu_desc_t *req[2];
    CHECK(vgpu_lookup(all, 10, ids, 2, req) == VGPU_OK && req[0] == &all[3] && req[1] == &all[7]);
    const char *bogus[1] = {"GPU-bogus-0"};
    CHECK(vgpu_lookup(all, 10, bogus, 1, req) == VGPU_E_UNKNOWN_ID);
    CHECK(strcmp(vgpu_last_error(), "unknown device: GPU-bogus-0") == 0);
    CHECK(vgpu_lookup(all, 10, ids, 2, req) == VGPU_OK);

    vgpu_plugin_cfg_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.split = 10;
    cfg.mem_scaling = 1.6667;
    cfg.cores_scaling = 1.0;
    strcpy(cfg.pcibusfile, "/usr/local/vgpu/pciinfo.vgpu");
    vgpu_kv_t envs[16];
    vgpu_mount_t mounts[16];
    size_t ne = 16, nm = 16;
    const char *pu[1] = {uuid}, *pi[1] = {"0"};
    CHECK(vgpu_allocate_env(&cfg, req, 2, pu, pi, 1, "01234567-89ab-4def-8123-456789abcdef", envs, &ne, mounts, &nm) == VGPU_OK);
    CHECK(strcmp(env(envs, ne, "NVIDIA_VISIBLE_DEVICES"), uuid) == 0);
    CHECK(strcmp(env(envs, ne, "CUDA_DEVICE_MEMORY_LIMIT_1"), "18432m") == 0);
    CHECK(strcmp(env(envs, ne, "NVIDIA_DEVICE_MAP"), "0:GPU-00000000-0000-0000-0000-000000000001 1:GPU-00000000-0000-0000-0000-000000000001") == 0);
    CHECK(strcmp(env(envs, ne, "CUDA_DEVICE_MEMORY_SHARED_CACHE"), "/tmp/01234567-89ab-4def-8123-456789abcdef.cache") == 0);
    CHECK(strcmp(env(envs, ne, "CUDA_OVERSUBSCRIBE"), "true") == 0);
    CHECK(nm == 5 && strcmp(mounts[1].container_path, "/etc/ld.so.preload") == 0 && mounts[1].read_only == 1);
    CHECK(strcmp(mounts[3].host_path, "/usr/local/vgpu/vgpuvalidator") == 0);
    ne = 2;  /* caller array too small: reports what is needed */
    CHECK(vgpu_allocate_env(&cfg, req, 2, pu, pi, 1, NULL, envs, &ne, mounts, &nm) == VGPU_E_RANGE && ne == 7);
    /* VGPU_MONITOR_MODE form (server.go:494-501): one more mount, in front, and the cache file inside it */
    ne = 16; nm = 16;
    CHECK(vgpu_allocate_env_monitor(&cfg, req, 2, pu, pi, 1, "01234567-89ab-4def-8123-456789abcdef", "train-0_worker", envs, &ne, mounts, &nm) == VGPU_OK);
    CHECK(nm == 6 && strcmp(mounts[0].container_path, "/train-0_worker") == 0 && mounts[0].read_only == 0);
    CHECK(strcmp(mounts[0].host_path, "/usr/local/vgpu/shared/train-0_worker") == 0 && strcmp(mounts[2].container_path, "/etc/ld.so.preload") == 0);
    CHECK(strcmp(env(envs, ne, "CUDA_DEVICE_MEMORY_SHARED_CACHE"), "/train-0_worker/01234567-89ab-4def-8123-456789abcdef.cache") == 0);
    CHECK(vgpu_allocate_env_monitor(&cfg, req, 2, pu, pi, 1, NULL, "a/b", envs, &ne, mounts, &nm) == VGPU_E_INVAL);
    /* the one additive variable (no reference counterpart), off unless asked for: which SM partition each vGPU owns */
    ne = 16; nm = 16;
    CHECK(vgpu_allocate_env(&cfg, req, 2, pu, pi, 1, NULL, envs, &ne, mounts, &nm) == VGPU_OK && env(envs, ne, "VGPU_SM_PARTITION_INDEX_0") == NULL);
    cfg.emit_partition_index = 1;
    ne = 16; nm = 16;
    CHECK(vgpu_allocate_env(&cfg, req, 2, pu, pi, 1, NULL, envs, &ne, mounts, &nm) == VGPU_OK && ne == 9);
    CHECK(strcmp(env(envs, ne, "VGPU_SM_PARTITION_INDEX_0"), "3") == 0 && strcmp(env(envs, ne, "VGPU_SM_PARTITION_INDEX_1"), "7") == 0);
    /* start()'s lspci pass (main.go:164-185) */
    char bus[64];
    CHECK(vgpu_pcibus_list("00:00.0 Host bridge: Intel\n1b:00.0 3D controller: NVIDIA Corporation Device 2901\n", bus, sizeof(bus)) == 1);
    CHECK(strcmp(bus, "1b:00.0\n") == 0);
    CHECK(vgpu_pcibus_list("1b:00.0 3D controller: NVIDIA Corporation Device 2901\n", bus, 4) == VGPU_E_RANGE);
    /* data-path argument checks need no GPU either */
    CHECK(vgpu_spill(VGPU_XFER_LDST, VGPU_TIER_LOCAL, NULL, NULL, NULL, NULL, 0, VGPU_CHUNK_BYTES, NULL, NULL) == VGPU_OK);
    CHECK(vgpu_spill_noalloc(NULL, NULL, NULL, NULL, VGPU_NOALLOC_MAX_CHUNKS + 1, 32, NULL, NULL) == VGPU_E_RANGE);
    CHECK(vgpu_policy_set_grid(7, 32) == VGPU_E_INVAL && vgpu_policy_set_grid(VGPU_TIER_HOST, 0) == VGPU_OK);
    CHECK(vgpu_policy_get(VGPU_TIER_LOCAL, 8) == VGPU_XFER_LDST && vgpu_policy_get(VGPU_TIER_LOCAL, 4) == VGPU_XFER_TMA);
    printf("c_abi_smoke ok\n");
    return 0;
}
