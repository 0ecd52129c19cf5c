/*
 * fedrann_oracle.c -- CPU restatement of the FEDRANN hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under fedrann_b200/ may import, link or
 * execute this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or the
 * timed CPU baseline.
 *
 * Parity status: the extract functions are pinned against the reference binary
 * compiled from /root/reference/kmer_searcher/kmer_searcher.cpp (oracle/_ref,
 * see oracle/Makefile) on its own fixtures (kmer_searcher/test/data) and on
 * seeded synthetic reads; the projection is pinned against the reference's
 * imported fedrann/feature_extraction.py (tests/golden/make_golden.py).  The
 * kNN stage restates exact cosine kNN (the reference's pynndescent backend is
 * an absent third-party dependency, pynndescent==0.5.12): parity unpinned for
 * that stage, cross-checked against sklearn brute force instead.
 *
 * Every function cites the reference lines it follows (paths relative to
 * /root/reference).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#define ORC_EMPTY UINT64_MAX

/* ---- a minimal u64 -> u64 map (role of robin_hood::unordered_map in
 * kmer_searcher.cpp:259; results never depend on hash order) ---- */
typedef struct {
    uint64_t *keys;
    uint64_t *vals;
    uint64_t cap; /* power of two */
    uint64_t n;
} orc_map;

static uint64_t orc_mix(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
    x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
    x ^= x >> 33; return x;
}

orc_map *orc_map_new(uint64_t expected) {
    orc_map *m = (orc_map *)malloc(sizeof(orc_map));
    uint64_t cap = 16;
    while (cap < expected * 2 + 2) cap <<= 1;
    m->cap = cap; m->n = 0;
    m->keys = (uint64_t *)malloc(cap * sizeof(uint64_t));
    m->vals = (uint64_t *)malloc(cap * sizeof(uint64_t));
    memset(m->keys, 0xff, cap * sizeof(uint64_t));
    return m;
}

void orc_map_free(orc_map *m) {
    if (!m) return;
    free(m->keys); free(m->vals); free(m);
}

static int orc_map_find(const orc_map *m, uint64_t key, uint64_t *val) {
    uint64_t i = orc_mix(key) & (m->cap - 1);
    while (m->keys[i] != ORC_EMPTY) {
        if (m->keys[i] == key) { *val = m->vals[i]; return 1; }
        i = (i + 1) & (m->cap - 1);
    }
    return 0;
}

static void orc_map_put(orc_map *m, uint64_t key, uint64_t val) {
    uint64_t i = orc_mix(key) & (m->cap - 1);
    while (m->keys[i] != ORC_EMPTY) i = (i + 1) & (m->cap - 1);
    m->keys[i] = key; m->vals[i] = val; m->n++;
}

uint64_t orc_map_size(const orc_map *m) { return m->n; }

/* kmer_searcher.cpp:138-151 -- 2 bits per base, A0 C1 G2 T3, case-insensitive,
 * anything else -> UINT64_MAX. */
uint64_t orc_kmer_to_int(const char *s, int len) {
    uint64_t v = 0;
    for (int i = 0; i < len; ++i) {
        v <<= 2;
        switch (s[i]) {
            case 'A': case 'a': break;
            case 'C': case 'c': v |= 1; break;
            case 'G': case 'g': v |= 2; break;
            case 'T': case 't': v |= 3; break;
            default: return UINT64_MAX;
        }
    }
    return v;
}

static int orc_is_space(char c) {
    return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f';
}

/* kmer_searcher.cpp:259-273 -- whitespace-separated tokens; tokens whose size
 * != K are skipped, tokens with a non-ACGT base are skipped, duplicates are
 * skipped; none of the skipped tokens consumes an index.  `codes_out` (room
 * for as many tokens as the text can hold) receives index -> code.  Returns a
 * map code -> index; *n_out = number of unique k-mers. */
orc_map *orc_library_load(const char *text, uint64_t text_len, int K,
                          uint64_t *codes_out, uint64_t *n_out) {
    uint64_t ntok = 0, i = 0;
    while (i < text_len) {
        while (i < text_len && orc_is_space(text[i])) ++i;
        if (i >= text_len) break;
        ++ntok;
        while (i < text_len && !orc_is_space(text[i])) ++i;
    }
    orc_map *m = orc_map_new(ntok);
    uint64_t index = 0;
    i = 0;
    while (i < text_len) {
        while (i < text_len && orc_is_space(text[i])) ++i;
        if (i >= text_len) break;
        uint64_t s = i;
        while (i < text_len && !orc_is_space(text[i])) ++i;
        if ((int64_t)(i - s) != (int64_t)K) continue;
        uint64_t code = orc_kmer_to_int(text + s, K);
        uint64_t dummy;
        if (code != UINT64_MAX && !orc_map_find(m, code, &dummy)) {
            orc_map_put(m, code, index);
            if (codes_out) codes_out[index] = code;
            ++index;
        }
    }
    *n_out = index;
    return m;
}

static int orc_cmp_u64(const void *a, const void *b) {
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

/* kmer_searcher.cpp:308-351 -- one read: literal replay of the consumer loop
 * (init over the first min(K,len) bases with ONE lookup at its end, then one
 * lookup per further base; a non-ACGT base sets the register to UINT64_MAX and
 * it is NOT reset afterwards).  Writes the read's hit set to `out` (capacity
 * >= max(1, len)), canonicalised ascending (the reference's in-row order is
 * robin_hood iteration order, kmer_searcher.cpp:94, i.e. unspecified).
 * Returns the number of distinct indices. */
uint64_t orc_extract_read(const orc_map *lib, const char *seq, uint64_t len, int K,
                          uint64_t *out) {
    uint64_t n = 0;
    uint64_t cur = 0;
    const uint64_t mask = (1ULL << (2 * K)) - 1;
    uint64_t idx;
    uint64_t i;
    for (i = 0; i < (uint64_t)K && i < len; ++i) {
        cur = (cur << 2) & mask;
        switch (seq[i]) {
            case 'A': case 'a': break;
            case 'C': case 'c': cur |= 1; break;
            case 'G': case 'g': cur |= 2; break;
            case 'T': case 't': cur |= 3; break;
            default: cur = UINT64_MAX; break;
        }
    }
    if (cur != UINT64_MAX && orc_map_find(lib, cur, &idx)) out[n++] = idx;
    for (i = (uint64_t)K; i < len; ++i) {
        cur = (cur << 2) & mask;
        switch (seq[i]) {
            case 'A': case 'a': break;
            case 'C': case 'c': cur |= 1; break;
            case 'G': case 'g': cur |= 2; break;
            case 'T': case 't': cur |= 3; break;
            default: cur = UINT64_MAX; break;
        }
        if (cur != UINT64_MAX && orc_map_find(lib, cur, &idx)) out[n++] = idx;
    }
    if (n > 1) {
        qsort(out, n, sizeof(uint64_t), orc_cmp_u64);
        uint64_t w = 1;
        for (uint64_t r = 1; r < n; ++r)
            if (out[r] != out[w - 1]) out[w++] = out[r];
        n = w;
    }
    return n;
}

/* kmer_searcher.cpp:295-354 + 56-86 -- all reads of a batch.  `bases` is the
 * concatenation of the read sequences, read r = bases[offsets[r]..offsets[r+1]).
 * Rows come out in INPUT order with indices ascending (the canonical form of
 * the reference's unordered output).  Two-call protocol:
 * indices == NULL -> only indptr / docfreq are filled.  docfreq[i] = number of
 * reads containing library k-mer i (kmer_searcher.cpp:327-328, 345-346). */
void orc_extract(const orc_map *lib, const char *bases, const int64_t *offsets,
                 int64_t n_reads, int K, int64_t *indptr, int32_t *indices,
                 uint64_t *docfreq) {
    uint64_t maxlen = 1;
    for (int64_t r = 0; r < n_reads; ++r) {
        uint64_t l = (uint64_t)(offsets[r + 1] - offsets[r]);
        if (l > maxlen) maxlen = l;
    }
    uint64_t *tmp = (uint64_t *)malloc(maxlen * sizeof(uint64_t));
    indptr[0] = 0;
    for (int64_t r = 0; r < n_reads; ++r) {
        uint64_t l = (uint64_t)(offsets[r + 1] - offsets[r]);
        uint64_t n = orc_extract_read(lib, bases + offsets[r], l, K, tmp);
        if (indices)
            for (uint64_t j = 0; j < n; ++j) indices[indptr[r] + (int64_t)j] = (int32_t)tmp[j];
        if (docfreq && !indices)
            for (uint64_t j = 0; j < n; ++j) docfreq[tmp[j]]++;
        indptr[r + 1] = indptr[r] + (int64_t)n;
    }
    free(tmp);
}

/* feature_extraction.py:136-140 (row pairing) + 204-213 (A_chunk.dot(P)).
 * For read r with ascending hit list H: row 2r = sum_{i in H} P[i,:], row
 * 2r+1 = the same over the index-swapped set {i+nk if i<nk else i-nk}, again
 * ascending.  scipy's csr_matmat visits A's columns in ascending order and
 * adds P's entries to a fp32 accumulator one at a time (A's data is 1, so
 * v*p == p exactly); we do the same, which reproduces the reference bit for
 * bit.  A read with no hits gives two all-zero rows (the
 * reference leaves garbage there, feature_extraction.py:187-195 -- stated
 * divergence).  P is CSR (F rows, D columns). */
void orc_project(const int64_t *indptr, const int32_t *indices, int64_t n_reads,
                 const int64_t *p_indptr, const int32_t *p_indices, const float *p_data,
                 int64_t nk, int64_t D, float *E) {
    for (int64_t r = 0; r < n_reads; ++r) {
        float *fwd = E + (2 * r) * D;
        float *rev = E + (2 * r + 1) * D;
        for (int64_t j = 0; j < 2 * D; ++j) fwd[j] = 0.0f;
        int64_t a = indptr[r], b = indptr[r + 1];
        for (int64_t t = a; t < b; ++t) {
            int64_t i = indices[t];
            for (int64_t e = p_indptr[i]; e < p_indptr[i + 1]; ++e) {
                volatile float s = fwd[p_indices[e]] + p_data[e];
                fwd[p_indices[e]] = s;
            }
        }
        /* swapped set, ascending: first the hits >= nk (mapped down), then the
         * hits < nk (mapped up) */
        int64_t split = a;
        while (split < b && indices[split] < nk) ++split;
        for (int pass = 0; pass < 2; ++pass) {
            int64_t lo = pass == 0 ? split : a, hi = pass == 0 ? b : split;
            for (int64_t t = lo; t < hi; ++t) {
                int64_t i = pass == 0 ? indices[t] - nk : indices[t] + nk;
                for (int64_t e = p_indptr[i]; e < p_indptr[i + 1]; ++e) {
                    volatile float s = rev[p_indices[e]] + p_data[e];
                    rev[p_indices[e]] = s;
                }
            }
        }
    }
}

/* Exact cosine kNN -- the restatement of "the reference's kNN backend in
 * exact mode" (nearest_neighbors.py:39-55 calls pynndescent, metric="cosine";
 * SURVEY.md 8(c)).  dist = 1 - x.y/(|x||y|) evaluated in fp32 on fp32
 * unit-normalised rows (dot accumulated in double, rounded once, so the
 * oracle carries no summation-order noise of its own); a zero-norm row is at
 * distance 1 from everything (including itself); full sort, ties broken by
 * lower index.  Queries are rows [q0, q1) of X; database = all N rows. */
typedef struct { float d; int32_t i; } orc_pair;

static int orc_cmp_pair(const void *a, const void *b) {
    const orc_pair *x = (const orc_pair *)a, *y = (const orc_pair *)b;
    if (x->d < y->d) return -1;
    if (x->d > y->d) return 1;
    return x->i < y->i ? -1 : (x->i > y->i ? 1 : 0);
}

void orc_normalize(const float *X, int64_t N, int64_t D, float *Xn, float *norms) {
    for (int64_t r = 0; r < N; ++r) {
        double s = 0.0;
        for (int64_t j = 0; j < D; ++j) s += (double)X[r * D + j] * (double)X[r * D + j];
        float nrm = (float)sqrt(s);
        if (norms) norms[r] = nrm;
        for (int64_t j = 0; j < D; ++j)
            Xn[r * D + j] = nrm > 0.0f ? X[r * D + j] / nrm : 0.0f;
    }
}

void orc_knn(const float *Xn, const float *norms, int64_t N, int64_t D, int64_t q0,
             int64_t q1, int k, int32_t *idx, float *dist) {
    orc_pair *row = (orc_pair *)malloc((size_t)N * sizeof(orc_pair));
    for (int64_t q = q0; q < q1; ++q) {
        const float *x = Xn + q * D;
        for (int64_t c = 0; c < N; ++c) {
            const float *y = Xn + c * D;
            float d;
            if (norms[q] == 0.0f || norms[c] == 0.0f) {
                d = 1.0f;
            } else {
                double s = 0.0;
                for (int64_t j = 0; j < D; ++j) s += (double)x[j] * (double)y[j];
                d = 1.0f - (float)s;
            }
            row[c].d = d; row[c].i = (int32_t)c;
        }
        qsort(row, (size_t)N, sizeof(orc_pair), orc_cmp_pair);
        for (int j = 0; j < k; ++j) {
            if (j < N) { idx[(q - q0) * k + j] = row[j].i; dist[(q - q0) * k + j] = row[j].d; }
            else { idx[(q - q0) * k + j] = -1; dist[(q - q0) * k + j] = INFINITY; }
        }
    }
    free(row);
}
