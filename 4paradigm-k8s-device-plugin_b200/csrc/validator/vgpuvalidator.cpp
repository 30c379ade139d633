// vgpuvalidator — GPU memory validator mounted at /usr/bin/vgpuvalidator in every vGPU pod
// (server.go:518-519).  The reference binary of this name is a dormant license tool
// (vgpuvalidator@0x4022e0 main; nothing calls libvgpu's check_validator); this one tests the
// vGPU's HBM slice: walking-ones and random patterns, fill kernel then verify kernel, bit-exact,
// and reports GB/s.  Inside a pod it runs under the preload shim, so the size is bounded by the
// vGPU quota and the kernels run in the vGPU's SM partition.
//
//   vgpuvalidator --memtest [--gib G | --bytes N] [--patterns walking,random] [--random-passes K]
//                 [--device i] [--seed S]
//   vgpuvalidator --roundtrip [--chunks N]      spill -> erase -> restore self-test (host tier)
//   vgpuvalidator --smids                       the SMs this vGPU's kernels run on (its green-context partition)
// Its kernels use no thread-block clusters, so under the shim it asks for the SM-granular partition split
// (VGPU_SM_PARTITION_GRANULE=fine unless the pod sets the variable itself): 10 % of 148 SMs = 14 instead of 8.
// exit code: 0 clean, 1 mismatches found, 2 usage or CUDA error.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "vgpu.h"

static int usage() {
    fprintf(stderr,
            "usage: vgpuvalidator --memtest [--gib G|--bytes N] [--patterns walking,random] [--random-passes K] [--device i] [--seed S]\n"
            "       vgpuvalidator --roundtrip [--chunks N] [--device i]\n"
            "       vgpuvalidator --smids [--device i]\n");
    return 2;
}

int main(int argc, char **argv) {
    bool memtest = false, roundtrip = false, smids = false;
    setenv("VGPU_SM_PARTITION_GRANULE", "fine", 0);  // read by the shim when it forms this process's partition
    uint64_t bytes = 1ull << 30, random_passes = 4, seed = 0, chunks = 64;
    unsigned patterns = 3;
    int device = 0;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&](const char *what) -> const char * {
            if (i + 1 >= argc) {
                fprintf(stderr, "%s needs a value\n", what);
                exit(2);
            }
            return argv[++i];
        };
        if (a == "--memtest") memtest = true;
        else if (a == "--roundtrip") roundtrip = true;
        else if (a == "--smids") smids = true;
        else if (a == "--gib") bytes = (uint64_t)(atof(next("--gib")) * (double)(1ull << 30));
        else if (a == "--bytes") bytes = strtoull(next("--bytes"), nullptr, 0);
        else if (a == "--random-passes") random_passes = strtoull(next("--random-passes"), nullptr, 0);
        else if (a == "--seed") seed = strtoull(next("--seed"), nullptr, 0);
        else if (a == "--device") device = atoi(next("--device"));
        else if (a == "--chunks") chunks = strtoull(next("--chunks"), nullptr, 0);
        else if (a == "--patterns") {
            std::string p = next("--patterns");
            patterns = 0;
            if (p.find("walking") != std::string::npos) patterns |= 1u << VGPU_PAT_WALKING_ONES;
            if (p.find("random") != std::string::npos) patterns |= 1u << VGPU_PAT_RANDOM;
        } else return usage();
    }
    if ((int)memtest + (int)roundtrip + (int)smids != 1) return usage();
    if (smids) {
        void *one = nullptr;  // any allocation brings up the device's context (the shim binds the partition there)
        uint64_t mask[4] = {0, 0, 0, 0};
        unsigned count = 0;
        if (vgpu_slab_alloc(VGPU_TIER_LOCAL, device, -1, 4096, &one) || vgpu_sm_probe(nullptr, mask, &count)) {
            fprintf(stderr, "vgpuvalidator: %s\n", vgpu_last_error());
            return 2;
        }
        vgpu_slab_free(VGPU_TIER_LOCAL, device, -1, one);
        printf("{\"device\": %d, \"sm_count\": %u, \"sm_ids\": [", device, count);
        bool first = true;
        for (unsigned sm = 0; sm < 256; ++sm)
            if (mask[sm >> 6] >> (sm & 63) & 1) {
                printf("%s%u", first ? "" : ", ", sm);
                first = false;
            }
        printf("]}\n");
        return 0;
    }
    if (memtest) {
        bytes &= ~7ull;
        vgpu_memtest_result_t r;
        memset(&r, 0, sizeof(r));
        int rc = vgpu_memtest(device, bytes, patterns, random_passes, seed, &r);
        if (rc != VGPU_OK && rc != VGPU_E_MISMATCH) {
            fprintf(stderr, "vgpuvalidator: %s\n", vgpu_last_error());
            return 2;
        }
        printf("{\"device\": %d, \"bytes\": %llu, \"passes\": %llu, \"mismatches\": %llu, \"first_bad_word\": %lld, "
               "\"bad_bits\": \"0x%llx\", \"first_bad_pattern\": %d, \"first_bad_pass\": %d, "
               "\"fill_ms\": %.3f, \"verify_ms\": %.3f, \"GBps\": %.1f}\n",
               device, (unsigned long long)r.bytes, (unsigned long long)r.passes, (unsigned long long)r.mismatches,
               r.mismatches ? (long long)r.first_bad_word : -1LL, (unsigned long long)r.bad_bits, r.first_bad_pattern,
               r.first_bad_pass, r.fill_ms, r.verify_ms, r.gbps);
        return r.mismatches ? 1 : 0;
    }
    // round trip through the host tier with the host-buffer entry points
    const size_t cb = VGPU_CHUNK_BYTES;
    void *hbm = nullptr, *host = nullptr;
    if (vgpu_slab_alloc(VGPU_TIER_LOCAL, device, -1, chunks * cb, &hbm) || vgpu_slab_alloc(VGPU_TIER_HOST, device, -1, chunks * cb, &host)) {
        fprintf(stderr, "vgpuvalidator: %s\n", vgpu_last_error());
        return 2;
    }
    vgpu_cksum_t *ck = (vgpu_cksum_t *)calloc(chunks, sizeof(*ck));
    vgpu_memtest_dev_result_t *d_res = nullptr;
    uint64_t bad = 0;
    int rc = vgpu_memtest_fill(hbm, chunks * cb, VGPU_PAT_RANDOM, 7, seed, nullptr);
    if (!rc) rc = vgpu_spill_to_host(VGPU_XFER_AUTO, device, hbm, host, chunks, cb, ck);
    if (!rc) rc = vgpu_memtest_fill(hbm, chunks * cb, VGPU_PAT_WALKING_ONES, 0, 0, nullptr);  // erase
    if (!rc) rc = vgpu_restore_from_host(VGPU_XFER_AUTO, device, host, hbm, chunks, cb, ck, &bad);
    vgpu_memtest_result_t r;
    memset(&r, 0, sizeof(r));
    (void)d_res;
    if (rc != VGPU_OK) {
        fprintf(stderr, "vgpuvalidator: %s\n", vgpu_last_error());
        return rc == VGPU_E_MISMATCH ? 1 : 2;
    }
    printf("{\"device\": %d, \"chunks\": %llu, \"bad_chunks\": %llu, \"round_trip\": \"bit-exact per (X,S) checksum\"}\n", device,
           (unsigned long long)chunks, (unsigned long long)bad);
    vgpu_slab_free(VGPU_TIER_HOST, device, -1, host);
    vgpu_slab_free(VGPU_TIER_LOCAL, device, -1, hbm);
    free(ck);
    return bad ? 1 : 0;
}
