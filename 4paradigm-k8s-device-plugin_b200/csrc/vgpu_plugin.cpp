// Contract A helpers (include/vgpu.h): the arithmetic and string building the
// Go plugin does in Device2VDevice / VDevicesByIDs / UniqueDeviceIDs / Allocate,
// restated behind a C-ABI so server.go and vdevice.go can call them with cgo.
// Pure functions over caller-owned memory: thread-safe, nothing allocated that
// outlives the call.  Reference: vdevice.go:36-90, server.go:459-524, 565-616.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "vgpu_internal.h"

using namespace vgpu;

namespace {

bool copy_str(char *dst, size_t cap, const std::string &s) {
    if (s.size() + 1 > cap) return false;
    memcpy(dst, s.c_str(), s.size() + 1);
    return true;
}

// uuid4 text from /dev/urandom, same shape as github.com/google/uuid NewString()
// (server.go:503): 8-4-4-4-12 lower-case hex, version nibble 4, variant 10xx.
int random_uuid4(char out[37]) {
    unsigned char b[16];
    FILE *f = fopen("/dev/urandom", "rb");
    if (!f || fread(b, 1, 16, f) != 16) {
        if (f) fclose(f);
        return fail(VGPU_E_INVAL, "cannot read /dev/urandom");
    }
    fclose(f);
    b[6] = (b[6] & 0x0f) | 0x40;
    b[8] = (b[8] & 0x3f) | 0x80;
    snprintf(out, 37, "%02x%02x%02x%02x-%02x%02x-%02x%02x-%02x%02x-%02x%02x%02x%02x%02x%02x", b[0], b[1],
             b[2], b[3], b[4], b[5], b[6], b[7], b[8], b[9], b[10], b[11], b[12], b[13], b[14], b[15]);
    return VGPU_OK;
}

}  // namespace

namespace {

// ---- best-effort policy (besteffort_policy.go) -----------------------------------------------
// A device is its index into the score matrix; kPad stands for the reference's nil padding.
constexpr int kPad = -1;
using Set = std::vector<int>;
using Partition = std::vector<Set>;

struct BestEffort {
    const int *score;
    size_t n;

    int pair(int a, int b) const { return (a == kPad || b == kPad || a == b) ? 0 : score[(size_t)a * n + (size_t)b]; }
    int set_score(const Set &s) const {
        int t = 0;
        for (size_t i = 0; i < s.size(); ++i)
            for (size_t j = i + 1; j < s.size(); ++j) t += pair(s[i], s[j]);
        return t;
    }
    int partition_score(const Partition &p) const {
        int t = 0;
        for (const Set &s : p) t += set_score(s);
        return t;
    }
    static int padding(const Set &s) {
        int c = 0;
        for (int d : s) c += d == kPad;
        return c;
    }
    static bool contains(const Set &s, int d) {
        for (int x : s)
            if (x == d) return true;
        return false;
    }
    static bool contains_all(const Set &s, const Set &sub) {
        for (int d : sub)
            if (!contains(s, d)) return false;
        return true;
    }

    // iterateGPUSets: all combinations of `size` positions, lexicographic
    template <typename F>
    static void sets(const Set &devs, int size, F &&cb) {
        if (size <= 0 || (size_t)size > devs.size()) return;
        std::vector<size_t> idx((size_t)size);
        for (size_t i = 0; i < (size_t)size; ++i) idx[i] = i;
        Set cur((size_t)size);
        for (;;) {
            for (size_t i = 0; i < (size_t)size; ++i) cur[i] = devs[idx[i]];
            cb(cur);
            int lvl = size - 1;
            while (lvl >= 0 && idx[(size_t)lvl] == devs.size() - (size_t)(size - lvl)) --lvl;
            if (lvl < 0) return;
            ++idx[(size_t)lvl];
            for (size_t i = (size_t)lvl + 1; i < (size_t)size; ++i) idx[i] = idx[i - 1] + 1;
        }
    }

    // iterateGPUPartitions
    template <typename F>
    void partitions(Set devices, int size, F &&cb) const {
        if (size <= 0 || (size_t)size > devices.size()) return;
        if (size == 1) {
            for (int d : devices) cb(Partition{Set{d}});
            return;
        }
        while (devices.size() % (size_t)size != 0) devices.push_back(kPad);
        const int pad = padding(devices);
        Partition accum;
        recurse(devices, size, pad, accum, cb);
    }
    template <typename F>
    void recurse(const Set &devices, int size, int pad, Partition &accum, F &&cb) const {
        if ((size_t)size == devices.size()) {
            accum.push_back(devices);
            cb(accum);
            accum.pop_back();
            return;
        }
        const Set rest(devices.begin() + 1, devices.end());
        sets(rest, size - 1, [&](const Set &tail) {
            Set set;
            set.push_back(devices[0]);
            set.insert(set.end(), tail.begin(), tail.end());
            const int p = padding(set);
            if (!(p == 0 || p == pad)) return;
            Set remaining;
            for (int d : devices)
                if (!contains(set, d)) remaining.push_back(d);
            accum.push_back(set);
            recurse(remaining, size, pad, accum, cb);
            accum.pop_back();
        });
    }
};

}  // namespace

extern "C" {

int vgpu_best_effort(const int *available, size_t n_available, const int *required, size_t n_required, int size,
                     const int *pair_score, size_t n_dev, int *out, size_t out_cap) {
    if ((!available && n_available) || (!required && n_required) || !pair_score || !out) return fail(VGPU_E_INVAL, "null argument");
    for (size_t i = 0; i < n_available; ++i)
        if (available[i] < 0 || (size_t)available[i] >= n_dev) return fail(VGPU_E_INVAL, "device index %d outside the score matrix", available[i]);
    for (size_t i = 0; i < n_required; ++i)
        if (required[i] < 0 || (size_t)required[i] >= n_dev) return fail(VGPU_E_INVAL, "device index %d outside the score matrix", required[i]);
    if (size <= 0 || n_available < (size_t)size || n_required > (size_t)size) return 0;  // besteffort_policy.go:35-45
    const BestEffort be{pair_score, n_dev};
    const Set avail(available, available + n_available), req(required, required + n_required);
    Partition best;
    bool have = false;
    int best_score = 0;
    be.partitions(avail, size, [&](const Partition &cand) {
        bool ok = false;  // gpuPartitionContainsSetWithAll: a padding-free set holding every required device
        for (const Set &s : cand) ok = ok || (BestEffort::contains_all(s, req) && BestEffort::padding(s) == 0);
        if (!ok) return;
        const int sc = be.partition_score(cand);
        if (sc > best_score || !have) {
            best = cand;
            best_score = sc;
            have = true;
        }
    });
    const Set *best_set = nullptr;
    int set_score = 0;
    for (const Set &s : best) {
        if (!BestEffort::contains_all(s, req)) continue;
        const int sc = be.set_score(s);
        if (!best_set || sc > set_score) {
            best_set = &s;
            set_score = sc;
        }
    }
    if (!best_set) return 0;
    if (best_set->size() > out_cap) return fail(VGPU_E_RANGE, "need room for %zu devices", best_set->size());
    for (size_t i = 0; i < best_set->size(); ++i) out[i] = (*best_set)[i];
    return (int)best_set->size();
}

int vgpu_split(const char *uuid, uint64_t total_mib, unsigned split, double mem_scaling, vgpu_desc_t *out,
               size_t out_cap) {
    if (!uuid || !out) return fail(VGPU_E_INVAL, "null argument");
    if (split < 1) return fail(VGPU_E_INVAL, "invalid --device-split-count option: %u", split);  // main.go:151-153
    if (!(mem_scaling > 0)) return fail(VGPU_E_INVAL, "invalid --device-memory-scaling option: %g", mem_scaling);  // main.go:154-156
    const std::string u(uuid);
    const bool mig = u.find("MIG") != std::string::npos;  // vdevice.go:40
    const unsigned n = mig ? 1 : split;
    if (out_cap < n) return fail(VGPU_E_RANGE, "need room for %u descriptors", n);
    // vdevice.go:49 — float64 multiply, divide, truncate.
    const uint64_t mem = mig ? 0 : (uint64_t)((double)total_mib * mem_scaling / (double)split);
    for (unsigned i = 0; i < n; ++i) {
        vgpu_desc_t &d = out[i];
        memset(&d, 0, sizeof(d));
        if (!copy_str(d.id, sizeof(d.id), u + "-" + std::to_string(i)) || !copy_str(d.uuid, sizeof(d.uuid), u))
            return fail(VGPU_E_RANGE, "uuid too long");
        d.memory_mib = mem;
        d.index = i;
        d.healthy = 1;
    }
    return (int)n;
}

int vgpu_sm_limit(unsigned split, double cores_scaling) {
    if (split < 1) return fail(VGPU_E_INVAL, "invalid --device-split-count option: %u", split);
    if (!(cores_scaling > 0)) return fail(VGPU_E_INVAL, "invalid --device-core-scaling option: %g", cores_scaling);  // main.go:157-159
    return (int)(100 * cores_scaling / (double)split);  // server.go:492
}

int vgpu_lookup(const vgpu_desc_t *all, size_t n_all, const char *const *ids, size_t n_ids,
                const vgpu_desc_t **out) {
    if ((!all && n_all) || (!ids && n_ids) || (!out && n_ids)) return fail(VGPU_E_INVAL, "null argument");
    for (size_t i = 0; i < n_ids; ++i) {
        const vgpu_desc_t *hit = nullptr;
        for (size_t j = 0; j < n_all && !hit; ++j)
            if (strcmp(all[j].id, ids[i]) == 0) hit = &all[j];
        if (!hit) return fail(VGPU_E_UNKNOWN_ID, "unknown device: %s", ids[i]);  // vdevice.go:72
        out[i] = hit;
    }
    return VGPU_OK;
}

// start() in main.go:164-185: the PCI bus ids of the node's NVIDIA devices, one per line, as the plugin writes them
// to $PCIBUSFILE (mounted read-only at /usr/local/vgpu/pciinfo.vgpu in every vGPU pod, server.go:515-517).  A line of
// `lspci` output counts when it is longer than one byte and contains "NVIDIA"; its id is everything before the first
// space.  Returns the number of ids, or VGPU_E_RANGE when `out` is too small.
int vgpu_pcibus_list(const char *lspci_out, char *out, size_t out_cap) {
    if (!lspci_out || !out || out_cap == 0) return fail(VGPU_E_INVAL, "null argument");
    std::string res;
    int n = 0;
    const char *p = lspci_out;
    while (*p) {
        const char *nl = strchr(p, '\n');
        const std::string line = nl ? std::string(p, nl) : std::string(p);
        if (line.size() > 1 && line.find("NVIDIA") != std::string::npos) {
            res += line.substr(0, line.find(' ')) + "\n";
            ++n;
        }
        if (!nl) break;
        p = nl + 1;
    }
    if (res.size() + 1 > out_cap) return fail(VGPU_E_RANGE, "need %zu bytes", res.size() + 1);
    memcpy(out, res.c_str(), res.size() + 1);
    return n;
}

int vgpu_allocate_env(const vgpu_plugin_cfg_t *cfg, const vgpu_desc_t *const *req, size_t n_req,
                      const char *const *phys_uuids, const char *const *phys_index, size_t n_phys,
                      const char *cache_uuid, vgpu_kv_t *envs, size_t *n_envs, vgpu_mount_t *mounts,
                      size_t *n_mounts) {
    return vgpu_allocate_env_monitor(cfg, req, n_req, phys_uuids, phys_index, n_phys, cache_uuid, NULL, envs, n_envs,
                                     mounts, n_mounts);
}

int vgpu_allocate_env_monitor(const vgpu_plugin_cfg_t *cfg, const vgpu_desc_t *const *req, size_t n_req,
                              const char *const *phys_uuids, const char *const *phys_index, size_t n_phys,
                              const char *cache_uuid, const char *shared_name, vgpu_kv_t *envs, size_t *n_envs,
                              vgpu_mount_t *mounts, size_t *n_mounts) {
    if (shared_name && (!*shared_name || strchr(shared_name, '/')))
        return fail(VGPU_E_INVAL, "shared_name must be a non-empty name without '/'");
    if (!cfg || (!req && n_req) || !envs || !n_envs || !mounts || !n_mounts)
        return fail(VGPU_E_INVAL, "null argument");
    const int sm = vgpu_sm_limit(cfg->split, cfg->cores_scaling);
    if (sm < 0) return sm;

    // UniqueDeviceIDs (vdevice.go:78-90): physical UUIDs, first-seen order.
    std::vector<std::string> uuids;
    for (size_t i = 0; i < n_req; ++i) {
        bool seen = false;
        for (const auto &u : uuids) seen = seen || u == req[i]->uuid;
        if (!seen) uuids.emplace_back(req[i]->uuid);
    }
    // deviceIDsFromUUIDs (server.go:565-581)
    std::vector<std::string> device_ids;
    if (cfg->id_strategy == 0) {
        device_ids = uuids;
    } else if (cfg->id_strategy == 1) {
        for (size_t d = 0; d < n_phys; ++d)  // cachedDevices order, not request order
            for (const auto &u : uuids)
                if (u == phys_uuids[d]) device_ids.emplace_back(phys_index[d]);
    } else {
        return fail(VGPU_E_INVAL, "invalid --device-id-strategy option: %d", cfg->id_strategy);
    }

    std::vector<std::pair<std::string, std::string>> kv;
    std::vector<vgpu_mount_t> mv;
    auto join = [](const std::vector<std::string> &v, const char *sep) {
        std::string s;
        for (size_t i = 0; i < v.size(); ++i) s += (i ? sep : "") + v[i];
        return s;
    };
    auto add_mount = [&](const std::string &cpath, const std::string &hpath, bool ro) -> bool {
        vgpu_mount_t m;
        memset(&m, 0, sizeof(m));
        if (!copy_str(m.container_path, sizeof(m.container_path), cpath) ||
            !copy_str(m.host_path, sizeof(m.host_path), hpath))
            return false;
        m.read_only = ro ? 1 : 0;
        mv.push_back(m);
        return true;
    };

    const char *root = "/var/run/nvidia-container-devices";  // server.go:58
    if (cfg->list_strategy == 0) {                            // envvar, server.go:469-471
        kv.emplace_back("NVIDIA_VISIBLE_DEVICES", join(device_ids, ","));
    } else if (cfg->list_strategy == 1) {                     // volume-mounts, server.go:472-475, 604-616
        kv.emplace_back("NVIDIA_VISIBLE_DEVICES", root);
        for (const auto &id : device_ids)
            if (!add_mount(std::string(root) + "/" + id, "/dev/null", false)) return fail(VGPU_E_RANGE, "path too long");
    } else {
        return fail(VGPU_E_INVAL, "invalid --device-list-strategy option: %d", cfg->list_strategy);
    }

    std::vector<std::string> map_envs;
    for (size_t i = 0; i < n_req; ++i) {  // server.go:487-491
        kv.emplace_back("CUDA_DEVICE_MEMORY_LIMIT_" + std::to_string(i), std::to_string(req[i]->memory_mib) + "m");
        map_envs.push_back(std::to_string(i) + ":" + req[i]->uuid);
    }
    kv.emplace_back("CUDA_DEVICE_SM_LIMIT", std::to_string(sm));        // server.go:492
    kv.emplace_back("NVIDIA_DEVICE_MAP", join(map_envs, " "));          // server.go:493
    char u4[37];
    if (cache_uuid) {
        if (strlen(cache_uuid) != 36) return fail(VGPU_E_INVAL, "cache_uuid must be 36 characters");
        memcpy(u4, cache_uuid, 37);
    } else if (int rc = random_uuid4(u4)) {
        return rc;
    }
    if (shared_name) {  // VGPU_MONITOR_MODE, server.go:494-501: the cache file lives in a directory the host can read
        const std::string dir = std::string("/") + shared_name;
        if (!add_mount(dir, std::string("/usr/local/vgpu/shared/") + shared_name, false)) return fail(VGPU_E_RANGE, "path too long");
        kv.emplace_back("CUDA_DEVICE_MEMORY_SHARED_CACHE", dir + "/" + u4 + ".cache");
    } else {
        kv.emplace_back("CUDA_DEVICE_MEMORY_SHARED_CACHE", std::string("/tmp/") + u4 + ".cache");  // server.go:503
    }
    if (cfg->mem_scaling > 1) kv.emplace_back("CUDA_OVERSUBSCRIBE", "true");                  // server.go:505-507
    // Additive and off by default (Contract A stays byte-identical to server.go:486-507): tell the pod WHICH of the
    // disjoint SM partitions each of its vGPUs owns — the i of "<UUID>-<i>" — so that the containers sharing a GPU
    // never land on the same SMs.  One variable per requested vGPU, numbered like CUDA_DEVICE_MEMORY_LIMIT_<n>.
    if (cfg->emit_partition_index)
        for (size_t i = 0; i < n_req; ++i)
            kv.emplace_back("VGPU_SM_PARTITION_INDEX_" + std::to_string(i), std::to_string(req[i]->index));

    // server.go:511-522, in this order, all read-only
    if (!add_mount("/usr/local/vgpu/libvgpu.so", "/usr/local/vgpu/libvgpu.so", true) ||
        !add_mount("/etc/ld.so.preload", "/usr/local/vgpu/ld.so.preload", true) ||
        !add_mount("/usr/local/vgpu/pciinfo.vgpu", cfg->pcibusfile, true) ||
        !add_mount("/usr/bin/vgpuvalidator", "/usr/local/vgpu/vgpuvalidator", true) ||
        !add_mount("/vgpu", "/usr/local/vgpu/license", true))
        return fail(VGPU_E_RANGE, "path too long");

    if (kv.size() > *n_envs || mv.size() > *n_mounts) {
        *n_envs = kv.size();
        *n_mounts = mv.size();
        return fail(VGPU_E_RANGE, "need %zu envs and %zu mounts", kv.size(), mv.size());
    }
    for (size_t i = 0; i < kv.size(); ++i) {
        memset(&envs[i], 0, sizeof(envs[i]));
        if (!copy_str(envs[i].key, sizeof(envs[i].key), kv[i].first) ||
            !copy_str(envs[i].val, sizeof(envs[i].val), kv[i].second))
            return fail(VGPU_E_RANGE, "env %s too long", kv[i].first.c_str());
    }
    for (size_t i = 0; i < mv.size(); ++i) mounts[i] = mv[i];
    *n_envs = kv.size();
    *n_mounts = mv.size();
    return VGPU_OK;
}

}  // extern "C"
