// Shared declarations of the fedrann_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/fedrann_b200.h"

#define FR_EMPTY_KEY 0xFFFFFFFFFFFFFFFFull
#define FR_HASH_MUL 0x9E3779B97F4A7C15ull

// Device-side view of the k-mer lookup structure (K2 builds it, K3 probes it).
//  * bloom: one-bit-per-key filter, 32-64 bits per key (a false-positive rate of 1.5-3 %), small
//    enough to stay L2-resident; a probe is ONE 32-byte L2 sector and ~5 instructions.
//  * slots: open-addressing table (linear probing, <= 70 % load).  When key and value fit 63
//    bits together (2k + ceil(log2(F+1)) <= 63: every configuration of the path) a slot is one
//    packed 64-bit word -- half the footprint, so the table stays L2-resident next to the
//    streamed reads; otherwise 16-byte slots {key lo, key hi, value, pad}.  Either way a probe
//    that passes the filter costs one 32-byte sector.
struct FrTable {
    const uint32_t *bloom;
    const uint4 *slots;        // wide mode (vbits == 0)
    const uint64_t *cslots;    // compact mode: (key << vbits) | value in 8 bytes, all-ones = empty
    int vbits;
    int lg_words;  // log2(#bloom words)
    int lg_cap;    // log2(#table slots)
    int k;
    int64_t n;     // number of features F
    // Set by fr_table_set_sampler once VERIFIED on the device: every key's canonical form c
    // satisfies fr_sample_keep(c, seed, threshold) (the library is a hash-threshold sample, as N1
    // produces it).  K3 then rejects a position whose canonical k-mer fails that test in
    // registers, without touching the filter.
    int sampled;
    uint64_t seed, threshold;
    // Core bitmap (K3a first stage): the k-mers ending at `core_w` consecutive positions all contain the
    // (k - core_w + 1)-mer ending at the FIRST of them; bit `core` is set iff some library k-mer contains that
    // core at one of the core_w offsets.  One probe (a 32-byte L2 sector, index = the core itself when
    // 2 * core_c <= core_lg, a multiplicative hash of it otherwise) rejects core_w positions at once.
    const uint32_t *cores;
    int core_w, core_c, core_lg, core_hashed;
};

struct fr_ctx {
    int device = 0;
    int sm_count = 148;
    std::string err;
    int64_t launches = 0;

    // K2
    uint32_t *d_bloom = nullptr;
    uint32_t *d_cores = nullptr;
    uint4 *d_slots = nullptr;
    uint64_t *d_cslots = nullptr;
    FrTable table{};
    bool has_table = false;

    // K3 scratch (kept between calls, grown on demand)
    int32_t *d_scratch = nullptr;   // per-read sorted unique hits, bump-allocated
    int64_t scratch_cap = 0;
    uint32_t *d_spill = nullptr;    // overflow hit lists for very long reads
    int64_t spill_cap = 0;
    int64_t *d_read_off = nullptr;  // [n_reads] offset into d_scratch
    int32_t *d_read_cnt = nullptr;  // [n_reads]
    int32_t *d_read_order = nullptr; // [n_reads] reads longest first
    int64_t reads_cap = 0;
    unsigned long long *d_counters = nullptr;  // [0] scratch used, [1] spill used, [2] flags
    unsigned long long *h_counters = nullptr;  // pinned mirror
    int64_t *d_scan_tmp = nullptr;
    int64_t scan_tmp_cap = 0;
    // state of the last fr_extract_count
    int64_t last_n_reads = -1;
    const int64_t *last_indptr = nullptr;

    // per-kernel CUDA-event timing (fr_profile_enable / fr_profile_read)
    struct KtRec { const char *name; cudaEvent_t e0, e1; };
    bool kt_on = false;
    bool kt_open = false;
    cudaStream_t kt_stream = nullptr;
    std::vector<KtRec> kt;

    // N1 k-mer count table
    uint4 *d_count_slots = nullptr;
    int64_t count_cap_bytes = 0;
    int count_lg_cap = 0;
    int count_k = 0;

    // K6/K7 scratch
    void *d_knn_ws = nullptr;
    int64_t knn_ws_cap = 0;
    cudaEvent_t knn_en_event = nullptr;   // one-shot: the fp32 rows are complete once this event has fired
    // one-shot (fr_exchange_pull): database shards that are still landing; K6 waits for flags[s] == epoch
    // before the first tile of shard s (rows [bounds[s], bounds[s+1])), starting at the local shard
    const uint32_t *knn_shard_flags = nullptr;
    int knn_n_shards = 0, knn_local_shard = 0;
    uint32_t knn_shard_epoch = 0;
    int64_t knn_shard_bounds[FR_MAX_RANKS + 1] = {0};
    void *d_knn_fb = nullptr;       // exact distances of the fallback rows (parallel fallback)
    int64_t knn_fb_cap = 0;
    int64_t knn_fallback_rows = 0;
    int knn_last_path = -1;         // 0 = tcgen05 pass + re-rank, 1 = fp32 scan (fr_knn_last_path)
    bool knn_fb_pending = false;    // h_counters[4] holds the count once knn_fb_stream has drained
    cudaStream_t knn_fb_stream = nullptr;
    void *knn_inc = nullptr;        // state of an incremental search (fr_knn_inc_*), owned by knn_tc.cu
    void (*knn_inc_free)(void *) = nullptr;
};

int fr_fail(fr_ctx *ctx, int code, const char *what);
int fr_cuda_fail(fr_ctx *ctx, cudaError_t e, const char *what);
int fr_ensure(fr_ctx *ctx, void **ptr, int64_t *cap, int64_t need_bytes, const char *what);

#define FR_CUDA(ctx, call)                                       \
    do {                                                         \
        cudaError_t _e = (call);                                 \
        if (_e != cudaSuccess) return fr_cuda_fail(ctx, _e, #call); \
    } while (0)

void fr_kt_begin(fr_ctx *ctx, const char *name, cudaStream_t st);
void fr_kt_end(fr_ctx *ctx);

#define FR_LAUNCH_CHECK(ctx, name)                               \
    do {                                                         \
        (ctx)->launches++;                                       \
        if ((ctx)->kt_open) fr_kt_end(ctx);                      \
        cudaError_t _e = cudaGetLastError();                     \
        if (_e != cudaSuccess) return fr_cuda_fail(ctx, _e, name); \
    } while (0)

__device__ __forceinline__ uint64_t fr_hash(uint64_t kmer) {
    return (kmer ^ (kmer >> 29)) * FR_HASH_MUL;
}
__device__ __forceinline__ uint32_t fr_table_slot(uint64_t h, int lg_cap) {
    return (uint32_t)(h >> (64 - lg_cap));
}
// index of a core in the core bitmap (see FrTable)
__host__ __device__ __forceinline__ uint32_t fr_core_index(uint64_t core, int lg, int hashed) {
    return hashed ? (uint32_t)((core * FR_HASH_MUL) >> (64 - lg)) : (uint32_t)core;
}
__device__ __forceinline__ uint64_t fr_splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
// N1's sampler (DESIGN.md "N1"): a canonical k-mer c is sampled iff fr_sample_keep(c, seed, threshold),
// threshold = fraction * 2^64.  A 32-bit mix (murmur3 finaliser over the folded k-mer) keeps the
// test cheap enough to run once per read position in K3's register-only pre-filter.
__host__ __device__ __forceinline__ uint32_t fr_sample_seedmix(uint64_t seed) {
    return (uint32_t)seed ^ ((uint32_t)(seed >> 32) * 0x85EBCA77u);
}
__host__ __device__ __forceinline__ uint32_t fr_sample_hash(uint32_t lo, uint32_t hi, uint32_t seedmix) {
    uint32_t x = lo ^ (hi * 0x9E3779B1u) ^ seedmix;
    x ^= x >> 16; x *= 0x85EBCA6Bu;
    x ^= x >> 13; x *= 0xC2B2AE35u;
    x ^= x >> 16;
    return x;
}
__host__ __device__ __forceinline__ bool fr_sample_keep(uint64_t c, uint64_t seed, uint64_t threshold) {
    return threshold == 0xFFFFFFFFFFFFFFFFull ||
           fr_sample_hash((uint32_t)c, (uint32_t)(c >> 32), fr_sample_seedmix(seed)) < (uint32_t)(threshold >> 32);
}
// Filter hash: 32-bit multiplicative.  k-mers of k <= 16 fit one word (hi == 0).
#define FR_BLOOM_MUL1 0x9E3779B1u
#define FR_BLOOM_MUL2 0x85EBCA77u
__device__ __forceinline__ uint32_t fr_bloom_hash(uint32_t lo, uint32_t hi) {
    return lo * FR_BLOOM_MUL1 + hi * FR_BLOOM_MUL2;
}
// word index = top lg_words bits; the tested bit is 31 - (the next 5 bits), so that the probe
// is `word << x` followed by a sign-bit test (x = h >> (27 - lg_words), shift taken mod 32)
__device__ __forceinline__ uint32_t fr_bloom_word(uint32_t h, int lg_words) { return h >> (32 - lg_words); }
__device__ __forceinline__ uint32_t fr_bloom_shift(uint32_t h, int lg_words) { return (h >> (27 - lg_words)) & 31u; }

// internal launchers (defined in the .cu files)
int fr_extract_launch(fr_ctx *ctx, const uint8_t *d_bases, int64_t n_bases, const int64_t *d_offsets,
                      int64_t n_reads, uint32_t *d_docfreq, cudaStream_t st);
int fr_scan_counts(fr_ctx *ctx, const int32_t *d_cnt, int64_t n, int64_t *d_indptr, cudaStream_t st);
int fr_gather_rows(fr_ctx *ctx, int32_t *d_indices, cudaStream_t st);
int fr_gather_rows_checked(fr_ctx *ctx, int64_t *d_indptr, int64_t n, int32_t *d_indices, int64_t cap,
                           long long *d_status, cudaStream_t st);
int fr_docfreq_launch(fr_ctx *ctx, int64_t nnz, uint32_t *d_docfreq, cudaStream_t st);
