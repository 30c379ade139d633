/*
 * vgpu.h — C-ABI of the B200-native vGPU data path and plugin helpers.
 *
 * This is the boundary a maintainer of gpues/4paradigm-k8s-device-plugin binds
 * from Go with cgo.  Everything is `extern "C"`, plain
 * pointers and sizes; no C++ or torch types cross it.  Citations are to the
 * reference tree (`/root/reference/...`) or, for the prebuilt shim, to
 * `vgpu/libvgpu.so@<vaddr> <symbol>` as in SURVEY.md §8.
 *
 * Conventions
 *   - return 0 on success; negative values are VGPU_E_*; vgpu_last_error()
 *     returns a thread-local message for the last failure on this thread.
 *   - "device pointer" means any address the launching GPU can dereference:
 *     its own HBM, a P2P-enabled peer's HBM, or mapped pinned host memory.
 *   - `stream` is a cudaStream_t / CUstream passed as void* (NULL = default).
 *   - data-path calls are asynchronous on `stream` unless stated otherwise.
 *   - there is NO CPU fallback anywhere behind this header.
 */
#ifndef VGPU_H_
#define VGPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VGPU_ABI_VERSION 2

/* ---- error codes ------------------------------------------------------- */
#define VGPU_OK            0
#define VGPU_E_INVAL      (-1)  /* bad argument (alignment, size, NULL) */
#define VGPU_E_CUDA       (-2)  /* a CUDA runtime/driver call failed */
#define VGPU_E_NOMEM      (-3)  /* allocation failed / capacity exceeded */
#define VGPU_E_UNKNOWN_ID (-4)  /* "unknown device: <id>" (vdevice.go:72) */
#define VGPU_E_RANGE      (-5)  /* caller array too small */
#define VGPU_E_MISMATCH   (-6)  /* checksum / pattern mismatch detected */
#define VGPU_E_UNSUPPORTED (-7)

int         vgpu_abi_version(void);
const char *vgpu_last_error(void);

/* ======================================================================== *
 *  Contract A — what Allocate/ListAndWatch hand to kubelet
 * ======================================================================== */

#define VGPU_ID_LEN    128
#define VGPU_KEY_LEN    64
#define VGPU_VAL_LEN  1024
#define VGPU_PATH_LEN  256

/* One virtual device.  Replaces `type VDevice` (vdevice.go:29-33). */
typedef struct vgpu_desc {
    char     id[VGPU_ID_LEN];    /* "<UUID>-<i>"            vdevice.go:52 */
    char     uuid[VGPU_ID_LEN];  /* physical GPU UUID       nvidia.go:150 */
    uint64_t memory_mib;         /* per-vGPU quota in MiB   vdevice.go:49 */
    uint32_t index;              /* i within the physical GPU */
    uint32_t healthy;            /* 1 = "Healthy"           nvidia.go:151 */
} vgpu_desc_t;

typedef struct vgpu_kv    { char key[VGPU_KEY_LEN]; char val[VGPU_VAL_LEN]; } vgpu_kv_t;
typedef struct vgpu_mount { char container_path[VGPU_PATH_LEN]; char host_path[VGPU_PATH_LEN]; int32_t read_only; } vgpu_mount_t;

/* Plugin flags that shape Contract A (main.go:98-118, 70-96). */
typedef struct vgpu_plugin_cfg {
    uint32_t split;              /* --device-split-count    main.go:98-103  */
    double   mem_scaling;        /* --device-memory-scaling main.go:105-110 */
    double   cores_scaling;      /* --device-cores-scaling  main.go:112-117 */
    int32_t  id_strategy;        /* 0 uuid, 1 index         main.go:84-89   */
    int32_t  list_strategy;      /* 0 envvar, 1 volume-mounts main.go:77-82 */
    char     pcibusfile[VGPU_PATH_LEN]; /* $PCIBUSFILE      server.go:516   */
    uint32_t emit_partition_index; /* no reference counterpart; 0 (default) keeps the response identical to
                                      server.go:486-507.  1 adds VGPU_SM_PARTITION_INDEX_<n>=<i of "<UUID>-<i>"> per
                                      requested vGPU, which the shim uses to pick its green-context partition, so
                                      the vGPUs of one GPU get disjoint SMs deterministically. */
} vgpu_plugin_cfg_t;

/* Replaces Device2VDevice (vdevice.go:36-58) for one physical GPU.
 * Writes `split` descriptors (1 with memory 0 if uuid contains "MIG",
 * vdevice.go:40-46).  Returns the number written or VGPU_E_*. */
int vgpu_split(const char *uuid, uint64_t total_mib, unsigned split,
               double mem_scaling, vgpu_desc_t *out, size_t out_cap);

/* CUDA_DEVICE_SM_LIMIT value: int(100*coresScaling/float64(split)), server.go:492. */
int vgpu_sm_limit(unsigned split, double cores_scaling);

/* Replaces VDevicesByIDs (vdevice.go:61-75): order-preserving lookup.
 * On an unknown id returns VGPU_E_UNKNOWN_ID and vgpu_last_error() is
 * exactly "unknown device: <id>". */
int vgpu_lookup(const vgpu_desc_t *all, size_t n_all, const char *const *ids,
                size_t n_ids, const vgpu_desc_t **out);

/* Replaces gpuallocator.NewBestEffortPolicy().Allocate (vendor/github.com/NVIDIA/go-gpuallocator/
 * gpuallocator/besteffort_policy.go:34-90): among all ways of dividing `available` into sets of
 * `size`, take the division with the highest total pair score that has a padding-free set containing
 * all of `required`, and return that division's best set containing them.  Devices are named by their
 * index into the n_dev x n_dev symmetric `pair_score` matrix (calculateGPUPairScore, :296-352: 100 per
 * NVLink, 10..60 for PCIe proximity).  Ties keep the first candidate in the reference's enumeration
 * order.  Returns the number of devices written to `out` (0 = no allocation) or VGPU_E_*.
 * (The reference's result for partitions of 6 or more sets depends on Go slice aliasing inside
 * iterateGPUPartitions; this restatement is exact up to 5 sets, i.e. for every case on <= 10 GPUs.) */
int vgpu_best_effort(const int *available, size_t n_available, const int *required, size_t n_required,
                     int size, const int *pair_score, size_t n_dev, int *out, size_t out_cap);

/* Replaces the per-container body of Allocate (server.go:459-524) for the
 * non-MIG, non-monitor, non-legacy path: env map (insertion order is not
 * significant in the reference — Go map — we emit a fixed order) and the
 * mount list in the reference's order.  `phys_uuids`/`phys_index` describe
 * cachedDevices for the index strategy (server.go:565-581).
 * `cache_uuid` is the 36-char uuid4 string for CUDA_DEVICE_MEMORY_SHARED_CACHE
 * (server.go:503); NULL = generate one from /dev/urandom. */
int vgpu_allocate_env(const vgpu_plugin_cfg_t *cfg,
                      const vgpu_desc_t *const *req, size_t n_req,
                      const char *const *phys_uuids, const char *const *phys_index, size_t n_phys,
                      const char *cache_uuid,
                      vgpu_kv_t *envs, size_t *n_envs,      /* in: capacity, out: count */
                      vgpu_mount_t *mounts, size_t *n_mounts);

/* The same with VGPU_MONITOR_MODE set (server.go:365-406 find the pending pod, 494-501 build the response):
 * `shared_name` is the reference's `<pod name>_<container name>`; the response then carries, BEFORE the five fixed
 * mounts, a read-write mount of /usr/local/vgpu/shared/<shared_name> at /<shared_name>, and the cache file moves
 * from /tmp into that directory, where a host-side monitor reads it.  NULL = vgpu_allocate_env.  Creating the host
 * directory (os.MkdirAll, :496) and finding the pod stay with the caller. */
int vgpu_allocate_env_monitor(const vgpu_plugin_cfg_t *cfg,
                              const vgpu_desc_t *const *req, size_t n_req,
                              const char *const *phys_uuids, const char *const *phys_index, size_t n_phys,
                              const char *cache_uuid, const char *shared_name,
                              vgpu_kv_t *envs, size_t *n_envs,
                              vgpu_mount_t *mounts, size_t *n_mounts);

/* Replaces the lspci pass of start() (main.go:164-185): `lspci_out` is the output of `lspci`; `out` receives the PCI
 * bus ids of the lines that name an NVIDIA device, one per line — the content of $PCIBUSFILE, which Allocate mounts at
 * /usr/local/vgpu/pciinfo.vgpu (server.go:515-517).  Returns the number of ids or VGPU_E_*. */
int vgpu_pcibus_list(const char *lspci_out, char *out, size_t out_cap);

/* ======================================================================== *
 *  Virtual-device-memory data path: 2 MiB chunk spill / restore
 *  (new mechanism behind CUDA_OVERSUBSCRIBE, server.go:505-507; the reference
 *  delegates to cuMemAllocManaged, libvgpu.so@0x32146 cuMemoryAllocate.)
 * ======================================================================== */

#define VGPU_CHUNK_BYTES 2097152u   /* IPCSIZE, libvgpu.so .data@0x620c8 */

/* kernel variant selector */
#define VGPU_XFER_AUTO 0   /* policy table: measured best for (tier, batch) */
#define VGPU_XFER_LDST 1   /* 256-bit vectorised coalesced loads/stores */
#define VGPU_XFER_TMA  2   /* cp.async.bulk global->smem->global ring */
#define VGPU_XFER_CE   3   /* copy engine (cudaMemcpyAsync) + checksum kernel */
#define VGPU_XFER_PULL 4   /* peer-tier spill only: run the bulk-copy kernel on the GPU that owns `dst`, so NVLink
                              carries reads (788 GB/s one way on B200) instead of writes (716 GB/s) */

/* tier of the far side of a transfer (used by the AUTO policy and stats) */
#define VGPU_TIER_LOCAL 0  /* same-GPU HBM */
#define VGPU_TIER_PEER  1  /* peer GPU HBM over NVLink P2P */
#define VGPU_TIER_HOST  2  /* pinned host RAM over PCIe */

/* Per-chunk checksum (X,S): X = xor of the chunk's u64 words,
 * S = sum of w[i]*(2i+1) mod 2^64, i = word index within the chunk. */
typedef struct vgpu_cksum { uint64_t x, s; } vgpu_cksum_t;

/* Result block of a restore (device-accessible memory, written by the kernel). */
typedef struct vgpu_restore_result {
    uint64_t bad_chunks;        /* chunks whose (X,S) != expected */
    uint64_t first_bad_inv;     /* ~(lowest bad chunk number); 0 = none */
} vgpu_restore_result_t;

/* Checksum only: cksum[c] = (X,S) of chunk c of `src`.  */
int vgpu_checksum(const void *src, const uint32_t *src_slot, size_t nchunks,
                  size_t chunk_bytes, vgpu_cksum_t *cksum, void *stream);

/* Spill: for c in [0,nchunks): copy chunk src_slot[c] of slab `src` to chunk
 * dst_slot[c] of slab `dst` (NULL slot table = identity) and write its
 * checksum to cksum[c].  One fused kernel; `tier` names where `dst` lives. */
int vgpu_spill(int variant, int tier, const void *src, void *dst,
               const uint32_t *src_slot, const uint32_t *dst_slot,
               size_t nchunks, size_t chunk_bytes,
               vgpu_cksum_t *cksum, void *stream);

/* Restore: same movement in the other direction; the checksum of what
 * arrived is compared with expect[c] inside the kernel and mismatches are
 * accumulated into *result (device-accessible; caller zeroes it).  `cksum`
 * may be NULL.  `tier` names where `src` lives. */
int vgpu_restore(int variant, int tier, const void *src, void *dst,
                 const uint32_t *src_slot, const uint32_t *dst_slot,
                 size_t nchunks, size_t chunk_bytes,
                 const vgpu_cksum_t *expect, vgpu_cksum_t *cksum,
                 vgpu_restore_result_t *result, void *stream);

/* Spill for callers that cannot allocate device memory: the eviction path of a vGPU
 * whose HBM slice is already full, and code running inside the preload shim's own
 * allocation hooks (an allocation there would be charged to the vGPU being relieved, or
 * simply fail).  Same result as vgpu_spill(VGPU_XFER_LDST, ...), but the per-chunk scratch
 * is a fixed 128 KiB array inside the library, so: at most VGPU_NOALLOC_MAX_CHUNKS chunks
 * per call (VGPU_E_RANGE beyond that); calls that overlap on one GPU race on the shared
 * array — serialise them; the first call on each GPU resolves the array's address and
 * should be made at start-up, before any stream capture.  Allocates nothing.
 * No counterpart in the reference (it has no spill path of its own at all). */
#define VGPU_NOALLOC_MAX_CHUNKS 4096
int vgpu_spill_noalloc(const void *src, void *dst, const uint32_t *src_slot,
                       const uint32_t *dst_slot, size_t nchunks, size_t chunk_bytes,
                       vgpu_cksum_t *cksum, void *stream);

/* Host-buffer forms (synchronous): `host` is caller memory.  If it is not
 * already pinned it is pinned for the duration of the call.  The checksums
 * land in host memory.  These are the calls bench.py's `e2e` leg times. */
int vgpu_spill_to_host(int variant, int device, const void *d_src, void *host,
                       size_t nchunks, size_t chunk_bytes, vgpu_cksum_t *h_cksum);
int vgpu_restore_from_host(int variant, int device, const void *host, void *d_dst,
                           size_t nchunks, size_t chunk_bytes,
                           const vgpu_cksum_t *h_expect, uint64_t *bad_chunks);

/* Spill-tier slabs (helpers for vgpuvalidator / bench / the vmem manager). */
int vgpu_slab_alloc(int tier, int device, int peer_device, size_t bytes, void **out);
int vgpu_slab_free(int tier, int device, int peer_device, void *p);

/* AUTO policy: which variant wins for (tier, nchunks).  vgpu_calibrate
 * measures the variants on scratch slabs of `device` (and `peer_device` for
 * VGPU_TIER_PEER) and fills the table; vgpu_policy_get reads it. */
int vgpu_calibrate(int device, int tier, int peer_device, size_t max_chunks);
int vgpu_policy_get(int tier, size_t nchunks);
int vgpu_policy_set(int tier, size_t min_chunks, int variant);
/* CTAs per transfer launch for a tier (0 = the built-in default).  The link tiers default to a grid that just covers
 * the link's bandwidth-delay product, so that a spill running in the primary context does not park link-bound CTAs on
 * the SMs of every tenant's partition. */
int vgpu_policy_set_grid(int tier, unsigned ctas);

/* Copy-engine rate of the link behind `tier` (cudaMemcpyAsync of `bytes` from `device`'s HBM out to a
 * slab of the tier and back in, best of 3, GB/s): the measured practical peak that the kernel rates
 * are quoted against beside the nominal 900 GB/s NVLink / 64 GB/s PCIe figures.  Synchronous. */
int vgpu_link_rate(int device, int tier, int peer_device, size_t bytes, double *out_gbps, double *in_gbps);

/* ======================================================================== *
 *  Virtual device memory manager: a vGPU address range larger than its resident
 *  HBM.  Replaces the reference's allocmode-0 route (cuMemoryAllocate@0x32146 ->
 *  cuMemAllocManaged, i.e. UVM demand paging) and finishes its dead allocmode-1
 *  VMM branch (@0x3225f-0x322a0, cuMapMemory@0x31b28 / cuUnmapMemory@0x37f8f,
 *  whose helpers are undefined symbols in the shipped binary): the range is
 *  reserved with cuMemAddressReserve, backed per 2 MiB chunk with
 *  cuMemCreate/cuMemMap in blocks of up to 512 chunks (1 GiB; VMM driver calls cost
 *  ~0.1-0.3 ms each whatever their size), and chunks move to/from the spill tier
 *  with the kernels above while the application's pointers stay valid.
 *  Residency is tracked per block: chunk ranges are rounded out to blocks.
 * ======================================================================== */
typedef struct vgpu_vmem_stats {
    uint64_t chunks_total;      /* size of the virtual range in chunks */
    uint64_t chunks_resident;   /* backed by HBM right now */
    uint64_t chunks_spilled;    /* living in the tier */
    uint64_t resident_limit;    /* max chunks that may be resident (the vGPU's HBM quota) */
    uint64_t tier_slots;        /* capacity of the tier in chunks */
    uint64_t evictions, fetches;          /* chunks moved out / in so far */
    uint64_t checksum_failures;           /* chunks whose (X,S) did not match on restore */
    double   evict_ms, fetch_ms;          /* wall time spent in evict / fetch incl. map/unmap */
    uint64_t evictions_peer, evictions_host; /* chunks sent to each tier */
    uint64_t restore_retries;             /* blocks restored a second time (copy engine) after a checksum failure */
    double   tier_setup_ms;               /* part of evict_ms spent on set-up rather than transfer: a context on the peer and
                                             its slab (once), pinning host memory by the block when the process-wide pool
                                             of pinned pieces has none to hand out */
} vgpu_vmem_stats_t;

/* Create a range of `virtual_bytes` (rounded up to 2 MiB) on `device`, of which at most
 * `resident_bytes` are backed by HBM at any time; the rest lives in a tier of `tier_bytes` (VGPU_TIER_PEER: a slab on
 * `peer_device`, allocated at the first eviction; VGPU_TIER_HOST: pinned host memory taken block by block as blocks
 * leave, from a pool shared by all ranges of the process — VGPU_VMEM_HOST_POOL_MB). */
int vgpu_vmem_create(int device, uint64_t virtual_bytes, uint64_t resident_bytes, int tier,
                     int peer_device, uint64_t tier_bytes, void **handle);
/* Two tiers in the order the north star asks for: victims go to `peer_device`'s HBM (over NVLink P2P)
 * while it has room, then to pinned host RAM.  Either size may be 0. */
int vgpu_vmem_create_tiered(int device, uint64_t virtual_bytes, uint64_t resident_bytes, int peer_device,
                            uint64_t peer_bytes, uint64_t host_bytes, void **handle);
int vgpu_vmem_base(void *handle, void **base);
/* Make chunks [first, first+n) resident: evicts least-recently-used chunks if the resident
 * limit would be exceeded, maps fresh memory for never-touched chunks, restores (and
 * checksum-verifies) spilled ones.  Synchronous.  VGPU_E_MISMATCH if a restore fails its
 * checksum, VGPU_E_NOMEM if n exceeds the resident limit or the tier is full. */
int vgpu_vmem_ensure(void *handle, uint64_t first, uint64_t n);
/* A block whose restore fails its checksum is restored once more through the copy engine; if it still fails it is
 * unmapped again and STAYS in the tier (slot, checksums and state untouched), so nothing is lost by the failed call
 * and a later vgpu_vmem_ensure can succeed.  The other blocks of the call become resident as usual. */
/* Where chunk `chunk` lives right now: *tier = VGPU_TIER_LOCAL and its address in the range while resident, the tier
 * kind and its address inside that tier (the peer's slab, or the block's pinned host piece) while spilled, *ptr = NULL
 * if never touched.  For tools and tests. */
int vgpu_vmem_locate(void *handle, uint64_t chunk, int *tier, void **ptr);
/* Chunks of [first, first+n) that are not resident now (what an ensure of the range would have to map or restore,
 * counted in whole blocks): lets a caller that shares the GPU's HBM between several ranges make room first. */
int vgpu_vmem_missing(void *handle, uint64_t first, uint64_t n, uint64_t *chunks);
/* One range leaves while another arrives: every resident block of `out` is spilled, and of chunks [first, first+n) of
 * `in` as many missing blocks are brought in as the memory freed by that allows — in lock step, one block each way per
 * step, so that a full-duplex link (PCIe, NVLink) carries both transfers at once instead of one after the other as with
 * vgpu_vmem_evict(out) followed by vgpu_vmem_ensure(in).  The call never holds more HBM than before it.  Both ranges
 * on one device.  Synchronous.  VGPU_OK also when part of [first, first+n) is still missing (`out` was the smaller
 * one): ask vgpu_vmem_missing.  Errors and per-block guarantees as for evict / ensure.  (No counterpart in the
 * reference: UVM's eviction and fault-in are the driver's, /root/reference/vgpu/libvgpu.so@0x32146.) */
int vgpu_vmem_swap(void *out, void *in, uint64_t first, uint64_t n);
/* Push chunks [first, first+n) that are resident out to the tier and release their HBM. */
int vgpu_vmem_evict(void *handle, uint64_t first, uint64_t n);
int vgpu_vmem_stats(void *handle, vgpu_vmem_stats_t *out);
int vgpu_vmem_destroy(void *handle);

/* ======================================================================== *
 *  vgpuvalidator memtest (new capability under the reference's mount name
 *  /usr/bin/vgpuvalidator, server.go:518-519; the reference binary is a
 *  license tool, vgpuvalidator@0x4022e0 main — no memtest exists there).
 * ======================================================================== */

#define VGPU_PAT_WALKING_ONES 0  /* word i = 1ull << ((i+pass)&63) */
#define VGPU_PAT_RANDOM       1  /* word i = splitmix64((GOLDEN^pass^seed)+i) */

typedef struct vgpu_memtest_dev_result {   /* device-side accumulator */
    uint64_t mismatches;        /* mismatching u64 words */
    uint64_t first_bad_inv;     /* ~(lowest mismatching word index); 0 = none */
    uint64_t bad_bits;          /* OR of got^expected */
} vgpu_memtest_dev_result_t;

typedef struct vgpu_memtest_result {
    uint64_t bytes;             /* buffer bytes per pass */
    uint64_t passes;            /* (fill,verify) pairs executed */
    uint64_t mismatches;
    uint64_t first_bad_word;    /* UINT64_MAX if none */
    uint64_t bad_bits;
    int32_t  first_bad_pattern; /* -1 if none */
    int32_t  first_bad_pass;
    double   fill_ms;           /* summed device time (CUDA events) */
    double   verify_ms;
    double   gbps;              /* 2*bytes*passes / (fill_ms+verify_ms) */
} vgpu_memtest_result_t;

/* Asynchronous single kernels on a caller buffer (16-byte aligned, bytes%16==0). */
int vgpu_memtest_fill(void *buf, uint64_t bytes, int pattern, uint64_t pass,
                      uint64_t seed, void *stream);
int vgpu_memtest_verify(const void *buf, uint64_t bytes, int pattern, uint64_t pass,
                        uint64_t seed, vgpu_memtest_dev_result_t *d_result, void *stream);

/* Synchronous: run `npasses` passes of `pattern` over a caller device buffer
 * on the current device, timing with CUDA events; fills *out (host). */
int vgpu_memtest_run(void *buf, uint64_t bytes, int pattern, uint64_t first_pass,
                     uint64_t npasses, uint64_t seed, void *stream,
                     vgpu_memtest_result_t *out);

/* Synchronous, self-allocating: what `vgpuvalidator --memtest` calls.
 * `patterns` is a bit mask (1<<VGPU_PAT_*); walking-ones runs 64 passes,
 * random runs `random_passes`. */
int vgpu_memtest(int device, uint64_t bytes, unsigned patterns, uint64_t random_passes,
                 uint64_t seed, vgpu_memtest_result_t *out);

/* ======================================================================== *
 *  SM partitioning (replaces rate_limiter, libvgpu.so@0x463b1, with Blackwell
 *  green contexts).  Implemented over the driver API.
 * ======================================================================== */

typedef struct vgpu_sm_partition {
    void    *green_ctx;   /* CUgreenCtx */
    void    *ctx;         /* CUcontext from cuCtxFromGreenCtx */
    void    *stream;      /* CUstream created on the green context */
    uint32_t sm_count;    /* SMs actually granted */
    uint32_t sm_total;    /* SMs on the device */
} vgpu_sm_partition_t;

/* Split `device` into `n` partitions of floor(total*percent/100) SMs each
 * (rounded by the driver to its granule).  percent = CUDA_DEVICE_SM_LIMIT. */
int vgpu_sm_partition_create(int device, unsigned percent, unsigned n,
                             vgpu_sm_partition_t *out);
/* Same with flags.  VGPU_SM_FINE asks the driver for SM-granular groups
 * (CU_DEV_SM_RESOURCE_SPLIT_IGNORE_SM_COSCHEDULING): 148 SMs / 10 vGPUs gives 10 x 14 instead of
 * 10 x 8, at the price of cluster co-scheduling guarantees inside a partition. */
#define VGPU_SM_FINE 1u
int vgpu_sm_partition_create_ex(int device, unsigned percent, unsigned n, unsigned flags,
                                vgpu_sm_partition_t *out);
int vgpu_sm_partition_destroy(vgpu_sm_partition_t *parts, unsigned n);
/* The SMs that kernels launched on `stream` in the calling thread's current context actually run on: bit b of
 * mask[i] = SM 64*i+b (read from %smid by a probe kernel).  Inside a pod (under the preload shim, where the current
 * context is the vGPU's green context) this is the vGPU's partition; `vgpuvalidator --smids` prints it.  Synchronous. */
int vgpu_sm_probe(void *stream, uint64_t mask[4], unsigned *count);

#ifdef __cplusplus
}
#endif
#endif /* VGPU_H_ */
