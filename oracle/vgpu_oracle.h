/*
 * oracle/vgpu_oracle.h — CPU restatement of the vGPU data-path math.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this.  The product
 * (libvgpu_dp.so) never links, loads or calls it.
 *
 * PARITY UNPINNED BY THE REFERENCE for everything in this file: the reference
 * has no spill kernel, no checksum and no memtest (SURVEY.md §0 facts 1-3:
 * "virtual device memory" is cuMemAllocManaged, libvgpu.so@0x32146;
 * vgpuvalidator@0x4022e0 is a license tool).  These functions restate the math
 * defined in SURVEY.md §8(d) — the only reference constant is the 2 MiB chunk
 * (IPCSIZE, libvgpu.so .data@0x620c8).  They are pinned by the hand-computed
 * known answers in tests/test_oracle.py.
 */
#ifndef VGPU_ORACLE_H_
#define VGPU_ORACLE_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORACLE_PAT_WALKING_ONES 0
#define ORACLE_PAT_RANDOM 1

uint64_t oracle_splitmix64(uint64_t x);
/* expected word i of a pass; SURVEY.md §8(d) config 2 */
uint64_t oracle_pattern_word(int pattern, uint64_t i, uint64_t pass, uint64_t seed);
/* words [first, first+n) of the buffer */
void oracle_fill(uint64_t *buf, uint64_t first, uint64_t n, int pattern, uint64_t pass, uint64_t seed);
/* returns mismatch count; *first_bad = lowest bad global word index (UINT64_MAX if none); *bad_bits = OR of diffs */
uint64_t oracle_verify(const uint64_t *buf, uint64_t first, uint64_t n, int pattern, uint64_t pass, uint64_t seed,
                       uint64_t *first_bad, uint64_t *bad_bits);
/* (X,S) of one chunk; SURVEY.md §8(d) config 4 */
void oracle_checksum(const void *chunk, size_t bytes, uint64_t *x, uint64_t *s);
/* spill/restore of nchunks chunks: memcpy + checksum per chunk.  slot tables may be NULL. */
void oracle_spill(const void *src, void *dst, const uint32_t *src_slot, const uint32_t *dst_slot, size_t nchunks,
                  size_t chunk_bytes, uint64_t *cksum_xs /* [2*nchunks] */);

/* ---- timed multi-thread drivers for the CPU baseline ------------------------ */
/* fill+verify `passes` passes over buf with `threads` threads; returns seconds; *mismatches out */
double oracle_memtest_mt(uint64_t *buf, uint64_t nwords, int pattern, uint64_t first_pass, uint64_t passes,
                         uint64_t seed, int threads, uint64_t *mismatches);
/* spill nchunks chunks src->dst with `threads` threads (one chunk stream per thread); returns seconds */
double oracle_spill_mt(const void *src, void *dst, size_t nchunks, size_t chunk_bytes, uint64_t *cksum_xs, int threads);

#ifdef __cplusplus
}
#endif
#endif
