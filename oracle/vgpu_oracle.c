/* See vgpu_oracle.h: test infrastructure only; parity unpinned by the reference. */
#define _GNU_SOURCE
#include "vgpu_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define GOLDEN 0x9E3779B97F4A7C15ull

uint64_t oracle_splitmix64(uint64_t x) {
    uint64_t z = x + GOLDEN;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

uint64_t oracle_pattern_word(int pattern, uint64_t i, uint64_t pass, uint64_t seed) {
    if (pattern == ORACLE_PAT_WALKING_ONES) return 1ull << ((i + pass) & 63);
    return oracle_splitmix64((GOLDEN ^ pass ^ seed) + i);
}

void oracle_fill(uint64_t *buf, uint64_t first, uint64_t n, int pattern, uint64_t pass, uint64_t seed) {
    for (uint64_t k = 0; k < n; ++k) buf[first + k] = oracle_pattern_word(pattern, first + k, pass, seed);
}

uint64_t oracle_verify(const uint64_t *buf, uint64_t first, uint64_t n, int pattern, uint64_t pass, uint64_t seed,
                       uint64_t *first_bad, uint64_t *bad_bits) {
    uint64_t bad = 0, bits = 0, fb = UINT64_MAX;
    for (uint64_t k = 0; k < n; ++k) {
        const uint64_t d = buf[first + k] ^ oracle_pattern_word(pattern, first + k, pass, seed);
        if (d) {
            if (!bad) fb = first + k;
            ++bad;
            bits |= d;
        }
    }
    if (first_bad) *first_bad = fb;
    if (bad_bits) *bad_bits = bits;
    return bad;
}

void oracle_checksum(const void *chunk, size_t bytes, uint64_t *x, uint64_t *s) {
    const uint64_t *w = (const uint64_t *)chunk;
    const size_t n = bytes / 8;
    uint64_t X = 0, S = 0;
    for (size_t i = 0; i < n; ++i) {
        X ^= w[i];
        S += w[i] * (uint64_t)(2 * i + 1);
    }
    *x = X;
    *s = S;
}

void oracle_spill(const void *src, void *dst, const uint32_t *src_slot, const uint32_t *dst_slot, size_t nchunks,
                  size_t chunk_bytes, uint64_t *cksum_xs) {
    for (size_t c = 0; c < nchunks; ++c) {
        const unsigned char *s = (const unsigned char *)src + (size_t)(src_slot ? src_slot[c] : c) * chunk_bytes;
        unsigned char *d = (unsigned char *)dst + (size_t)(dst_slot ? dst_slot[c] : c) * chunk_bytes;
        memcpy(d, s, chunk_bytes);
        oracle_checksum(d, chunk_bytes, &cksum_xs[2 * c], &cksum_xs[2 * c + 1]);
    }
}

/* ---- multi-thread drivers ---------------------------------------------------- */
static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

typedef struct {
    uint64_t *buf;
    uint64_t first, n;
    int pattern;
    uint64_t first_pass, passes, seed, bad;
} mt_job_t;

static void *mt_worker(void *arg) {
    mt_job_t *j = (mt_job_t *)arg;
    for (uint64_t p = j->first_pass; p < j->first_pass + j->passes; ++p) {
        oracle_fill(j->buf, j->first, j->n, j->pattern, p, j->seed);
        j->bad += oracle_verify(j->buf, j->first, j->n, j->pattern, p, j->seed, NULL, NULL);
    }
    return NULL;
}

double oracle_memtest_mt(uint64_t *buf, uint64_t nwords, int pattern, uint64_t first_pass, uint64_t passes,
                         uint64_t seed, int threads, uint64_t *mismatches) {
    if (threads < 1) threads = 1;
    pthread_t *th = (pthread_t *)calloc((size_t)threads, sizeof(*th));
    mt_job_t *jobs = (mt_job_t *)calloc((size_t)threads, sizeof(*jobs));
    const uint64_t per = (nwords + (uint64_t)threads - 1) / (uint64_t)threads;
    const double t0 = now_s();
    for (int t = 0; t < threads; ++t) {
        const uint64_t first = per * (uint64_t)t;
        const uint64_t n = first >= nwords ? 0 : (nwords - first < per ? nwords - first : per);
        jobs[t] = (mt_job_t){buf, first, n, pattern, first_pass, passes, seed, 0};
        pthread_create(&th[t], NULL, mt_worker, &jobs[t]);
    }
    uint64_t bad = 0;
    for (int t = 0; t < threads; ++t) {
        pthread_join(th[t], NULL);
        bad += jobs[t].bad;
    }
    const double t1 = now_s();
    if (mismatches) *mismatches = bad;
    free(th);
    free(jobs);
    return t1 - t0;
}

typedef struct {
    const unsigned char *src;
    unsigned char *dst;
    size_t c0, c1, chunk_bytes;
    uint64_t *xs;
} sp_job_t;

static void *sp_worker(void *arg) {
    sp_job_t *j = (sp_job_t *)arg;
    if (j->c1 > j->c0)
        oracle_spill(j->src + j->c0 * j->chunk_bytes, j->dst + j->c0 * j->chunk_bytes, NULL, NULL, j->c1 - j->c0,
                     j->chunk_bytes, j->xs + 2 * j->c0);
    return NULL;
}

double oracle_spill_mt(const void *src, void *dst, size_t nchunks, size_t chunk_bytes, uint64_t *cksum_xs, int threads) {
    if (threads < 1) threads = 1;
    pthread_t *th = (pthread_t *)calloc((size_t)threads, sizeof(*th));
    sp_job_t *jobs = (sp_job_t *)calloc((size_t)threads, sizeof(*jobs));
    const size_t per = (nchunks + (size_t)threads - 1) / (size_t)threads;
    const double t0 = now_s();
    for (int t = 0; t < threads; ++t) {
        size_t c0 = per * (size_t)t, c1 = c0 + per;
        if (c0 > nchunks) c0 = nchunks;
        if (c1 > nchunks) c1 = nchunks;
        jobs[t] = (sp_job_t){(const unsigned char *)src, (unsigned char *)dst, c0, c1, chunk_bytes, cksum_xs};
        pthread_create(&th[t], NULL, sp_worker, &jobs[t]);
    }
    for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
    const double t1 = now_s();
    free(th);
    free(jobs);
    return t1 - t0;
}
