// K2 -- k-mer library -> device lookup structure.
// Replaces the robin_hood::unordered_map built at kmer_searcher.cpp:259-273:
// feature index i (first-occurrence rank, assigned by fr_library_parse) is
// stored under its 2-bit code.  Results of K3 depend only on set membership,
// never on the hash function or insertion order.
#include "common.cuh"

#include <cmath>
#include <cstdlib>

__global__ void fr_table_insert_compact_kernel(const uint64_t *__restrict__ codes, int64_t n,
                                               unsigned long long *cslots, uint32_t *bloom, int lg_cap,
                                               int lg_words, int vbits) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t key = codes[i];
    const uint32_t bh = fr_bloom_hash((uint32_t)key, (uint32_t)(key >> 32));
    atomicOr(&bloom[fr_bloom_word(bh, lg_words)], 0x80000000u >> fr_bloom_shift(bh, lg_words));
    const uint32_t cap_mask = (uint32_t)((1ull << lg_cap) - 1);
    uint32_t slot = fr_table_slot(fr_hash(key), lg_cap);
    const unsigned long long mine = (key << vbits) | (unsigned long long)i;
    while (true) {
        unsigned long long old = atomicCAS(&cslots[slot], (unsigned long long)FR_EMPTY_KEY, mine);
        if (old == FR_EMPTY_KEY) return;
        if ((old >> vbits) == key) {                 // duplicates keep the first occurrence
            atomicMin(&cslots[slot], mine);
            return;
        }
        slot = (slot + 1) & cap_mask;
    }
}

__global__ void fr_table_insert_kernel(const uint64_t *__restrict__ codes, int64_t n, uint4 *slots,
                                       uint32_t *bloom, int lg_cap, int lg_words) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t key = codes[i];
    const uint32_t bh = fr_bloom_hash((uint32_t)key, (uint32_t)(key >> 32));
    atomicOr(&bloom[fr_bloom_word(bh, lg_words)], 0x80000000u >> fr_bloom_shift(bh, lg_words));
    const uint32_t cap_mask = (uint32_t)((1ull << lg_cap) - 1);
    uint32_t slot = fr_table_slot(fr_hash(key), lg_cap);
    while (true) {
        unsigned long long *kp = reinterpret_cast<unsigned long long *>(&slots[slot]);
        unsigned long long old = atomicCAS(kp, (unsigned long long)FR_EMPTY_KEY, (unsigned long long)key);
        if (old == FR_EMPTY_KEY || old == key) {
            atomicMin(&slots[slot].z, (uint32_t)i);  // duplicates keep the first occurrence
            return;
        }
        slot = (slot + 1) & cap_mask;
    }
}

// core bitmap: every library k-mer sets the bits of its core_w cores (the c-mers at offsets 0 .. core_w - 1)
__global__ void fr_table_cores_kernel(const uint64_t *__restrict__ codes, int64_t n, uint32_t *cores, int K, int w,
                                      int c, int lg, int hashed) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t key = codes[i];
    const uint64_t cmask = c >= 32 ? ~0ull : ((1ull << (2 * c)) - 1);
    for (int o = 0; o < w; ++o) {
        const uint64_t core = (key >> (2 * (K - c - o))) & cmask;
        const uint32_t idx = fr_core_index(core, lg, hashed);
        atomicOr(&cores[idx >> 5], 1u << (idx & 31));
    }
}

// every occupied slot's key must be a hash-threshold sample: counters[3] counts the violations
__global__ void fr_table_verify_sampler_kernel(const uint4 *__restrict__ slots, const uint64_t *__restrict__ cslots,
                                               int vbits, int64_t cap, int K, uint64_t seed, uint64_t threshold,
                                               unsigned long long *bad) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += (int64_t)gridDim.x * blockDim.x) {
        uint64_t key;
        if (vbits) {
            const uint64_t w = cslots[i];
            if (w == FR_EMPTY_KEY) continue;
            key = w >> vbits;
        } else {
            const uint4 v = slots[i];
            key = ((uint64_t)v.y << 32) | v.x;
            if (key == FR_EMPTY_KEY) continue;
        }
        uint64_t rc = 0, x = key;
        for (int j = 0; j < K; ++j) { rc = (rc << 2) | (3ull - (x & 3ull)); x >>= 2; }
        const uint64_t can = key < rc ? key : rc;
        if (!fr_sample_keep(can, seed, threshold)) atomicAdd(bad, 1ull);
    }
}

extern "C" int fr_table_set_sampler(fr_ctx *ctx, uint64_t seed, uint64_t threshold, void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (!ctx->has_table) return fr_fail(ctx, FR_ERR_STATE, "fr_table_set_sampler: no table");
    cudaStream_t st = (cudaStream_t)stream;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    ctx->table.sampled = 0;
    FR_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 3, 0, sizeof(unsigned long long), st));
    fr_kt_begin(ctx, "fr_table_verify_sampler_kernel", st);
    fr_table_verify_sampler_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(ctx->d_slots, ctx->d_cslots, ctx->table.vbits,
                                                                      1ll << ctx->table.lg_cap, ctx->table.k, seed,
                                                                      threshold, ctx->d_counters + 3);
    FR_LAUNCH_CHECK(ctx, "fr_table_verify_sampler_kernel");
    FR_CUDA(ctx, cudaMemcpyAsync(&ctx->h_counters[3], ctx->d_counters + 3, sizeof(unsigned long long),
                                 cudaMemcpyDeviceToHost, st));
    FR_CUDA(ctx, cudaStreamSynchronize(st));
    if (ctx->h_counters[3] != 0)
        return fr_fail(ctx, FR_ERR_UNVERIFIED, "fr_table_set_sampler: the library is not a sample of this sampler");
    ctx->table.sampled = 1;
    ctx->table.seed = seed;
    ctx->table.threshold = threshold;
    return FR_OK;
}

static int ceil_log2(int64_t x) {
    int l = 0;
    while ((1ll << l) < x) ++l;
    return l;
}

extern "C" int fr_table_build(fr_ctx *ctx, const uint64_t *d_codes, int64_t n, int k, void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (k < 1 || k > 31) return fr_fail(ctx, FR_ERR_ARG, "fr_table_build: k must be in [1,31]");
    if (n < 0 || n >= (1ll << 31)) return fr_fail(ctx, FR_ERR_ARG, "fr_table_build: n out of range");
    cudaStream_t st = (cudaStream_t)stream;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->d_bloom) { cudaFree(ctx->d_bloom); ctx->d_bloom = nullptr; }
    if (ctx->d_cores) { cudaFree(ctx->d_cores); ctx->d_cores = nullptr; }
    if (ctx->d_slots) { cudaFree(ctx->d_slots); ctx->d_slots = nullptr; }
    if (ctx->d_cslots) { cudaFree(ctx->d_cslots); ctx->d_cslots = nullptr; }
    ctx->has_table = false;
    int lg_cap = ceil_log2(n > 0 ? (n * 10 + 6) / 7 : 1);   // load <= 0.7
    if (lg_cap < 10) lg_cap = 10;
    int lg_words = ceil_log2(n > 0 ? n : 1);                // 32..64 filter bits per key
    if (lg_words < 8) lg_words = 8;
    if (lg_words > 27) lg_words = 27;
    if (const char *ov = getenv("FR_BLOOM_LG")) lg_words = atoi(ov);   // timing experiments only
    size_t cap = (size_t)1 << lg_cap, words = (size_t)1 << lg_words;
    int vbits = ceil_log2(n + 2);                           // values 0..n-1 never read as all-ones
    if (vbits < 1) vbits = 1;
    if (2 * k + vbits > 63 || getenv("FR_TABLE_WIDE")) vbits = 0;   // wide slots
    if (cudaMalloc(&ctx->d_bloom, words * 4) != cudaSuccess ||
        (vbits ? cudaMalloc(&ctx->d_cslots, cap * 8) : cudaMalloc(&ctx->d_slots, cap * 16)) != cudaSuccess)
        return fr_fail(ctx, FR_ERR_NOMEM, "fr_table_build: cudaMalloc failed");
    // core bitmap geometry: w positions share the (k - w + 1)-mer ending at the first of them.  A probe costs
    // about a third of what a surviving position costs downstream (queue + table probe), so minimise
    // 1/w + 3 * P(group passes), P = 1 - exp(-w n / bits); cores of <= 14 bases index the bitmap directly.
    int core_w = 1, core_c = k, core_lg = 10, core_hashed = 0;
    // bitmap of at most 2^LGMAX bits (32 MB): cores of <= 14 bases index it directly.  Smaller caps (hashed
    // 16 / 8 MB bitmaps) measured the same within noise (profiles/r02f_bench_c2_corelg*.json)
    int LGMAX = 28;
    if (const char *ov = getenv("FR_CORE_LG")) LGMAX = atoi(ov);      // timing experiments only
    if (LGMAX < 16) LGMAX = 16;
    if (LGMAX > 30) LGMAX = 30;
    {
        double best = 1e30;
        for (int w = 1; w <= 8 && w <= k; w *= 2) {
            const int c = k - w + 1;
            const int lg_eff = 2 * c <= LGMAX ? 2 * c : LGMAX;
            const double p = 1.0 - exp(-(double)w * (double)(n > 0 ? n : 1) / ldexp(1.0, lg_eff));
            const double cost = 1.0 / w + 3.0 * p;
            if (cost < best) {
                best = cost; core_w = w; core_c = c;
                core_hashed = 2 * c > LGMAX;
                core_lg = lg_eff < 10 ? 10 : lg_eff;
            }
        }
        if (const char *ov = getenv("FR_CORE_W")) {          // timing experiments only
            const int w = atoi(ov);
            if ((w == 1 || w == 2 || w == 4 || w == 8) && w <= k) {
                core_w = w; core_c = k - core_w + 1;
                core_hashed = 2 * core_c > LGMAX;
                core_lg = core_hashed ? LGMAX : (2 * core_c < 10 ? 10 : 2 * core_c);
            }
        }
    }
    const size_t core_words = ((size_t)1 << core_lg) / 32;
    if (cudaMalloc(&ctx->d_cores, core_words * 4) != cudaSuccess)
        return fr_fail(ctx, FR_ERR_NOMEM, "fr_table_build: cudaMalloc failed (core bitmap)");
    FR_CUDA(ctx, cudaMemsetAsync(ctx->d_cores, 0, core_words * 4, st));
    if (vbits) FR_CUDA(ctx, cudaMemsetAsync(ctx->d_cslots, 0xFF, cap * 8, st));
    else FR_CUDA(ctx, cudaMemsetAsync(ctx->d_slots, 0xFF, cap * 16, st));
    FR_CUDA(ctx, cudaMemsetAsync(ctx->d_bloom, 0, words * 4, st));
    if (n > 0) {
        int threads = 256;
        int64_t blocks = (n + threads - 1) / threads;
        fr_kt_begin(ctx, "fr_table_insert_kernel", st);
        if (vbits)
            fr_table_insert_compact_kernel<<<(unsigned)blocks, threads, 0, st>>>(
                d_codes, n, (unsigned long long *)ctx->d_cslots, ctx->d_bloom, lg_cap, lg_words, vbits);
        else
            fr_table_insert_kernel<<<(unsigned)blocks, threads, 0, st>>>(d_codes, n, ctx->d_slots, ctx->d_bloom,
                                                                          lg_cap, lg_words);
        FR_LAUNCH_CHECK(ctx, "fr_table_insert_kernel");
        fr_kt_begin(ctx, "fr_table_cores_kernel", st);
        fr_table_cores_kernel<<<(unsigned)blocks, threads, 0, st>>>(d_codes, n, ctx->d_cores, k, core_w, core_c, core_lg,
                                                                   core_hashed);
        FR_LAUNCH_CHECK(ctx, "fr_table_cores_kernel");
    }
    ctx->table.cores = ctx->d_cores;
    ctx->table.core_w = core_w;
    ctx->table.core_c = core_c;
    ctx->table.core_lg = core_lg;
    ctx->table.core_hashed = core_hashed;
    ctx->table.cslots = ctx->d_cslots;
    ctx->table.vbits = vbits;
    ctx->table.bloom = ctx->d_bloom;
    ctx->table.slots = ctx->d_slots;
    ctx->table.lg_cap = lg_cap;
    ctx->table.lg_words = lg_words;
    ctx->table.k = k;
    ctx->table.n = n;
    ctx->table.sampled = 0;
    ctx->has_table = true;
    return FR_OK;
}
