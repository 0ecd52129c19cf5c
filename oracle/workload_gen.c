/*
 * workload_gen.c -- host-side twin of fedrann_b200/synthetic_device.py (TEST / BENCH INFRASTRUCTURE,
 * never linked by fedrann_b200/).
 *
 * bench.py's reference arm (`--impl reference`) and its cpu_baseline leg must not touch the GPU
 * or the product library, yet they need the SAME workload the native arm builds on the device:
 * the genome, the reads of a sample, and the k-mer library counted over ALL reads of the
 * configuration (7.4 Gbases at the chr1 scale).  The device generator derives every base from
 * counter hashes of global positions (splitmix64), so the same bytes can be produced here with
 * plain loops on the host threads.  tests/test_host_cpu.py checks byte equality against the
 * torch implementation run on the CPU device.
 *
 * k-mer counting + sampling follows the definition the harness uses for N1 (DESIGN.md "N1";
 * replaces jellyfish count -C | dump -L | awk sampler of fedrann/count_kmers.py:76-123):
 * canonical k-mers (min of k-mer and reverse complement, A0 C1 G2 T3), windows inside one read,
 * kept iff fr_sample_hash(canonical, seed) < fraction * 2^32.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline uint64_t sm64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

/* ---------------------------------------------------------------- genome */
typedef struct { uint8_t *out; int64_t a, b; uint64_t key; } GenomeJob;

static void *genome_worker(void *p) {
    GenomeJob *j = (GenomeJob *)p;
    for (int64_t i = j->a; i < j->b; ++i) j->out[i] = (uint8_t)((sm64((uint64_t)i ^ j->key) >> 40) & 3);
    return NULL;
}

/* synthetic_device.make_genome: codes 0..3 */
void wl_genome(uint8_t *out, int64_t n, uint64_t seed, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    GenomeJob jobs[256];
    const uint64_t key = seed * 0x2545F4914F6CDD1Dull + 0x1234567ull;
    for (int t = 0; t < threads; ++t) {
        jobs[t].out = out; jobs[t].key = key;
        jobs[t].a = n * t / threads; jobs[t].b = n * (t + 1) / threads;
        pthread_create(&th[t], NULL, genome_worker, &jobs[t]);
    }
    for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
}

/* ---------------------------------------------------------------- reads */
typedef struct {
    const uint8_t *genome;
    const int64_t *lens, *starts, *pre;   /* pre-indel length, genome start, global pre-indel offset */
    const int8_t *strands;
    uint64_t key;                         /* seed * 0x9E3779B97F4A7C15 + 0xABCDEF */
    uint32_t thr_sub, thr_ins, thr_del;   /* p * 2^24; 0 = off */
} ReadModel;

/* Emits the bases (codes 0..3) of read r through `emit(ctx, code)`; returns the emitted count.
 * Mirrors synthetic_device.generate_reads base for base. */
#define WL_FOR_EACH_BASE(M, r, EMIT)                                                              \
    do {                                                                                          \
        const int64_t L_ = (M)->lens[r], st_ = (M)->starts[r], pre_ = (M)->pre[r];                \
        const int rev_ = (M)->strands[r] > 0;                                                     \
        for (int64_t j_ = 0; j_ < L_; ++j_) {                                                     \
            uint32_t code_ = (M)->genome[rev_ ? st_ + L_ - 1 - j_ : st_ + j_];                    \
            if (rev_) code_ = 3u - code_;                                                         \
            const uint64_t h_ = sm64((uint64_t)(pre_ + j_) ^ (M)->key);                           \
            if ((M)->thr_sub && (uint32_t)(h_ & 0xFFFFFF) < (M)->thr_sub)                         \
                code_ = (code_ + 1u + (uint32_t)(((h_ >> 24) & 0xFFFF) % 3)) & 3u;                \
            int rep_ = 1;                                                                         \
            uint64_t h2_ = 0;                                                                     \
            if ((M)->thr_ins || (M)->thr_del) {                                                   \
                h2_ = sm64(h_);                                                                   \
                if ((M)->thr_del && (uint32_t)(h2_ & 0xFFFFFF) < (M)->thr_del) rep_ -= 1;         \
                if ((M)->thr_ins && (uint32_t)((h2_ >> 24) & 0xFFFFFF) < (M)->thr_ins) rep_ += 1; \
            }                                                                                     \
            if (rep_ >= 1) { const uint32_t c_ = code_; EMIT(c_); }                               \
            if (rep_ == 2) { const uint32_t c_ = (uint32_t)((h2_ >> 50) & 3); EMIT(c_); }         \
        }                                                                                         \
    } while (0)

typedef struct {
    ReadModel m;
    int64_t r0, r1;
    int64_t *out_lens;       /* pass 1: post-indel lengths of reads [r0, r1) */
    const int64_t *out_offs; /* pass 2: offsets (relative to r0) */
    uint8_t *out_bases;      /* pass 2: ASCII */
    int t, T;
} ReadsJob;

static void *reads_len_worker(void *p) {
    ReadsJob *j = (ReadsJob *)p;
    for (int64_t r = j->r0 + j->t; r < j->r1; r += j->T) {
        int64_t n = 0;
#define EMIT_COUNT(c) (void)(c); ++n
        WL_FOR_EACH_BASE(&j->m, r, EMIT_COUNT);
        j->out_lens[r - j->r0] = n;
    }
    return NULL;
}

static void *reads_fill_worker(void *p) {
    ReadsJob *j = (ReadsJob *)p;
    static const char ascii[4] = {'A', 'C', 'G', 'T'};
    for (int64_t r = j->r0 + j->t; r < j->r1; r += j->T) {
        uint8_t *o = j->out_bases + j->out_offs[r - j->r0];
#define EMIT_STORE(c) *o++ = (uint8_t)ascii[c]
        WL_FOR_EACH_BASE(&j->m, r, EMIT_STORE);
    }
    return NULL;
}

static void model_init(ReadModel *m, const uint8_t *genome, const int64_t *lens, const int64_t *starts,
                       const int8_t *strands, const int64_t *pre, uint64_t seed, uint32_t thr_sub, uint32_t thr_ins,
                       uint32_t thr_del) {
    m->genome = genome; m->lens = lens; m->starts = starts; m->strands = strands; m->pre = pre;
    m->key = seed * 0x9E3779B97F4A7C15ull + 0xABCDEFull;
    m->thr_sub = thr_sub; m->thr_ins = thr_ins; m->thr_del = thr_del;
}

static void run_jobs(void *(*fn)(void *), ReadsJob *proto, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    ReadsJob jobs[256];
    for (int t = 0; t < threads; ++t) {
        jobs[t] = *proto; jobs[t].t = t; jobs[t].T = threads;
        pthread_create(&th[t], NULL, fn, &jobs[t]);
    }
    for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
}

/* post-indel lengths of reads [r0, r1) */
void wl_read_lengths(const uint8_t *genome, const int64_t *lens, const int64_t *starts, const int8_t *strands,
                     const int64_t *pre, uint64_t seed, uint32_t thr_sub, uint32_t thr_ins, uint32_t thr_del,
                     int64_t r0, int64_t r1, int64_t *out_lens, int threads) {
    ReadsJob j;
    memset(&j, 0, sizeof j);
    model_init(&j.m, genome, lens, starts, strands, pre, seed, thr_sub, thr_ins, thr_del);
    j.r0 = r0; j.r1 = r1; j.out_lens = out_lens;
    run_jobs(reads_len_worker, &j, threads);
}

/* ASCII bases of reads [r0, r1) at out_bases + out_offs[r - r0] */
void wl_read_bases(const uint8_t *genome, const int64_t *lens, const int64_t *starts, const int8_t *strands,
                   const int64_t *pre, uint64_t seed, uint32_t thr_sub, uint32_t thr_ins, uint32_t thr_del,
                   int64_t r0, int64_t r1, const int64_t *out_offs, uint8_t *out_bases, int threads) {
    ReadsJob j;
    memset(&j, 0, sizeof j);
    model_init(&j.m, genome, lens, starts, strands, pre, seed, thr_sub, thr_ins, thr_del);
    j.r0 = r0; j.r1 = r1; j.out_offs = out_offs; j.out_bases = out_bases;
    run_jobs(reads_fill_worker, &j, threads);
}

/* ---------------------------------------------------------------- sampled canonical k-mer counts */
typedef struct { uint64_t *keys; uint32_t *vals; uint64_t cap, n; } CountMap;   /* open addressing, key+1 stored (0 = empty) */

static void cm_init(CountMap *m, uint64_t cap) {
    m->cap = cap; m->n = 0;
    m->keys = (uint64_t *)calloc(cap, 8);
    m->vals = (uint32_t *)calloc(cap, 4);
}
static void cm_add(CountMap *m, uint64_t key, uint32_t by);
static void cm_grow(CountMap *m) {
    CountMap big;
    cm_init(&big, m->cap * 2);
    for (uint64_t i = 0; i < m->cap; ++i)
        if (m->keys[i]) cm_add(&big, m->keys[i] - 1, m->vals[i]);
    free(m->keys); free(m->vals);
    *m = big;
}
static void cm_add(CountMap *m, uint64_t key, uint32_t by) {
    if ((m->n + 1) * 10 > m->cap * 7) cm_grow(m);
    uint64_t i = sm64(key) & (m->cap - 1);
    const uint64_t k1 = key + 1;
    while (m->keys[i] && m->keys[i] != k1) i = (i + 1) & (m->cap - 1);
    if (!m->keys[i]) { m->keys[i] = k1; m->n++; }
    m->vals[i] += by;
}

/* fr_sample_hash of csrc/common.cuh / synthetic.sample_keep */
static inline uint32_t sample_hash(uint64_t c, uint32_t seedmix) {
    uint32_t x = (uint32_t)c ^ ((uint32_t)(c >> 32) * 0x9E3779B1u) ^ seedmix;
    x ^= x >> 16; x *= 0x85EBCA6Bu;
    x ^= x >> 13; x *= 0xC2B2AE35u;
    x ^= x >> 16;
    return x;
}

typedef struct {
    ReadModel m;
    int64_t r0, r1;
    int k, t, T, all_pass;
    uint32_t seedmix, thr32;
    CountMap map;
} CountJob;

static void *count_worker(void *p) {
    CountJob *j = (CountJob *)p;
    const int k = j->k;
    const uint64_t mask = k < 32 ? ((1ull << (2 * k)) - 1) : ~0ull;
    const int rsh = 2 * (k - 1);
    cm_init(&j->map, 1u << 16);
    for (int64_t r = j->r0 + j->t; r < j->r1; r += j->T) {
        uint64_t fw = 0, rc = 0;
        int have = 0;
#define EMIT_ROLL(c)                                                             \
        fw = ((fw << 2) | (uint64_t)(c)) & mask;                                 \
        rc = (rc >> 2) | ((uint64_t)(3u - (c)) << rsh);                          \
        if (++have >= k) {                                                       \
            const uint64_t can = fw < rc ? fw : rc;                              \
            if (j->all_pass || sample_hash(can, j->seedmix) < j->thr32) cm_add(&j->map, can, 1); \
        }
        WL_FOR_EACH_BASE(&j->m, r, EMIT_ROLL);
    }
    return NULL;
}

static CountJob *g_jobs = NULL;
static int g_njobs = 0;

/* Counts the sampled canonical k-mers of reads [r0, r1).  Returns the total number of (code, count)
 * entries over all threads' tables (a code may appear once per thread); fetch them with wl_count_fetch. */
int64_t wl_count_sampled(const uint8_t *genome, const int64_t *lens, const int64_t *starts, const int8_t *strands,
                         const int64_t *pre, uint64_t seed, uint32_t thr_sub, uint32_t thr_ins, uint32_t thr_del,
                         int64_t r0, int64_t r1, int k, uint32_t seedmix, uint32_t thr32, int all_pass, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    if (g_jobs) {
        for (int t = 0; t < g_njobs; ++t) { free(g_jobs[t].map.keys); free(g_jobs[t].map.vals); }
        free(g_jobs);
    }
    g_jobs = (CountJob *)calloc(threads, sizeof(CountJob));
    g_njobs = threads;
    pthread_t th[256];
    for (int t = 0; t < threads; ++t) {
        CountJob *j = &g_jobs[t];
        model_init(&j->m, genome, lens, starts, strands, pre, seed, thr_sub, thr_ins, thr_del);
        j->r0 = r0; j->r1 = r1; j->k = k; j->t = t; j->T = threads;
        j->seedmix = seedmix; j->thr32 = thr32; j->all_pass = all_pass;
        pthread_create(&th[t], NULL, count_worker, j);
    }
    int64_t total = 0;
    for (int t = 0; t < threads; ++t) { pthread_join(th[t], NULL); total += (int64_t)g_jobs[t].map.n; }
    return total;
}

void wl_count_fetch(uint64_t *codes, uint32_t *counts) {
    int64_t o = 0;
    for (int t = 0; t < g_njobs; ++t) {
        CountMap *m = &g_jobs[t].map;
        for (uint64_t i = 0; i < m->cap; ++i)
            if (m->keys[i]) { codes[o] = m->keys[i] - 1; counts[o] = m->vals[i]; ++o; }
        free(m->keys); free(m->vals);
        m->keys = NULL; m->vals = NULL; m->n = 0; m->cap = 0;
    }
    free(g_jobs);
    g_jobs = NULL;
    g_njobs = 0;
}
