/* synth_gen.c — threaded generator of the synthetic benchmark inputs.
 *
 * Byte-identical with the numpy definition in vlg_matching_b200/synth.py (english_text / dna_text), which
 * stays the specification: everything derives from splitmix64 used as a counter-based generator, so blocks
 * can be produced independently.  The numpy version makes 16 MB/s; the 8 GiB text of BASELINE config 4
 * needs this one.  Stands in for the reference's collections (benchmark/gapped-matching/test_case.config:27-53),
 * which are not redistributable.  Input generation only: nothing here matches anything.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline uint64_t draw(uint64_t seed, uint64_t i) /* draw i of stream `seed` (synth.splitmix64) */
{
    uint64_t z = seed + (i + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

/* ---------------------------------------------------------------- english_text */
typedef struct {
    uint8_t* out;
    uint64_t n, seed;
    const uint8_t* blob;   /* vocabulary bytes */
    const int64_t* off;    /* [vocab + 1] */
    const uint64_t* cum;   /* [vocab] cumulative Zipf weights */
    uint32_t vocab;
    uint64_t block_words;
    /* per round */
    uint64_t b0;           /* first block of the round */
    int nblk;
    uint64_t* blk_len;     /* [nblk] bytes each block produces */
    uint64_t* blk_start;   /* [nblk] where it goes */
    uint32_t** wid;        /* [nblk] word ids of the block (kept between the two passes) */
    int pass;
} eng_job;

typedef struct {
    eng_job* job;
    int tid, nthreads;
} eng_arg;

static inline uint32_t upper_bound_u64(const uint64_t* a, uint32_t n, uint64_t x) /* first idx with a[idx] > x */
{
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if (a[mid] <= x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

static void* eng_worker(void* p)
{
    eng_arg* a = (eng_arg*)p;
    eng_job* j = a->job;
    const uint64_t total = j->cum[j->vocab - 1];
    static const uint8_t punct[4] = {'.', ',', ',', ';'};
    for (int b = a->tid; b < j->nblk; b += a->nthreads) {
        const uint64_t w0 = (j->b0 + (uint64_t)b) * j->block_words;
        uint32_t* wid = j->wid[b];
        if (j->pass == 0) {
            uint64_t len = 0;
            for (uint64_t w = 0; w < j->block_words; w++) {
                const uint64_t r0 = draw(j->seed, 2 * (w0 + w)), r1 = draw(j->seed, 2 * (w0 + w) + 1);
                const uint32_t id = upper_bound_u64(j->cum, j->vocab, r0 % total);
                const uint32_t pr = (uint32_t)(r1 % 64);
                wid[w] = id | (pr < 4 ? (0x80000000u | (pr << 29)) : 0u);
                len += (uint64_t)(j->off[id + 1] - j->off[id]) + (pr < 4) + 1;
            }
            j->blk_len[b] = len;
        } else {
            uint64_t at = j->blk_start[b];
            if (at >= j->n) continue;
            for (uint64_t w = 0; w < j->block_words && at < j->n; w++) {
                const uint32_t v = wid[w], id = v & 0x1FFFFFFFu;
                const uint64_t wl = (uint64_t)(j->off[id + 1] - j->off[id]);
                uint8_t tmp[32];
                uint64_t l = wl;
                memcpy(tmp, j->blob + j->off[id], wl);
                if (v & 0x80000000u) tmp[l++] = punct[(v >> 29) & 3u];
                tmp[l++] = 0x20;
                const uint64_t take = at + l <= j->n ? l : j->n - at;
                memcpy(j->out + at, tmp, take);
                at += take;
            }
        }
    }
    return NULL;
}

static void run_threads(void* (*fn)(void*), void* job, int nthreads)
{
    pthread_t th[256];
    eng_arg args[256];
    if (nthreads > 256) nthreads = 256;
    for (int t = 0; t < nthreads; t++) {
        args[t].job = (eng_job*)job;
        args[t].tid = t;
        args[t].nthreads = nthreads;
        pthread_create(&th[t], NULL, fn, &args[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
}

/* returns 0, or -1 on allocation failure / words longer than 30 bytes */
int vlgs_english(uint8_t* out, uint64_t n, uint64_t seed, const uint8_t* blob, const int64_t* off, const uint64_t* cum,
                 uint32_t vocab, uint64_t block_words, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    for (uint32_t i = 0; i < vocab; i++)
        if (off[i + 1] - off[i] > 30) return -1;
    eng_job j;
    memset(&j, 0, sizeof j);
    j.out = out; j.n = n; j.seed = seed; j.blob = blob; j.off = off; j.cum = cum; j.vocab = vocab;
    j.block_words = block_words;
    j.nblk = nthreads;
    j.blk_len = (uint64_t*)calloc((size_t)j.nblk, sizeof(uint64_t));
    j.blk_start = (uint64_t*)calloc((size_t)j.nblk, sizeof(uint64_t));
    j.wid = (uint32_t**)calloc((size_t)j.nblk, sizeof(uint32_t*));
    int ok = j.blk_len && j.blk_start && j.wid;
    for (int b = 0; ok && b < j.nblk; b++) ok = (j.wid[b] = (uint32_t*)malloc(block_words * sizeof(uint32_t))) != NULL;
    uint64_t filled = 0;
    while (ok && filled < n) {
        j.pass = 0;
        run_threads(eng_worker, &j, nthreads);
        for (int b = 0; b < j.nblk; b++) {
            j.blk_start[b] = filled;
            filled += j.blk_len[b];
        }
        j.pass = 1;
        run_threads(eng_worker, &j, nthreads);
        j.b0 += (uint64_t)j.nblk;
    }
    if (j.wid) for (int b = 0; b < j.nblk; b++) free(j.wid[b]);
    free(j.wid); free(j.blk_len); free(j.blk_start);
    return ok ? 0 : -1;
}

/* ---------------------------------------------------------------- dna_text */
typedef struct {
    uint8_t* out;
    uint64_t n, seed;
} dna_job;

static void* dna_worker(void* p)
{
    eng_arg* a = (eng_arg*)p;
    dna_job* j = (dna_job*)a->job;
    static const uint8_t acgt[4] = {'A', 'C', 'G', 'T'};
    const uint64_t words = (j->n + 31) / 32;
    const uint64_t per = (words + (uint64_t)a->nthreads - 1) / (uint64_t)a->nthreads;
    uint64_t w0 = per * (uint64_t)a->tid, w1 = w0 + per < words ? w0 + per : words;
    for (uint64_t w = w0; w < w1; w++) {
        uint64_t r = draw(j->seed, w);
        const uint64_t base = w * 32;
        const uint64_t cnt = base + 32 <= j->n ? 32 : j->n - base;
        for (uint64_t k = 0; k < cnt; k++) j->out[base + k] = acgt[(r >> (2 * k)) & 3u];
    }
    return NULL;
}

int vlgs_dna(uint8_t* out, uint64_t n, uint64_t seed, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    dna_job j = {out, n, seed};
    run_threads(dna_worker, &j, nthreads);
    return 0;
}
