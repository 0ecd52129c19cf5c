// N1 -- canonical k-mer counting with a seeded hash-threshold sampler.
// Replaces the pre-step of run_kmer_searcher (fedrann/count_kmers.py:76-123):
// `jellyfish count -C` + `jellyfish dump -L <min-mult>` + the awk sampler.  The
// reference samples k-mers AFTER counting with awk's rand(); that stream cannot be
// reproduced without that awk binary and no test pins it, so
// sampling here is a deterministic function of the canonical k-mer:
//     keep(c) <=> fr_sample_keep(c, seed, fraction * 2^64)   (32-bit mix of the k-mer, common.cuh)
// which commutes with counting -- only sampled k-mers ever enter the count table.
// jellyfish semantics kept: canonical form = min(k-mer, reverse complement) on 2-bit
// codes A0 C1 G2 T3, case-insensitive, windows containing a non-ACGT base or crossing a
// read boundary are skipped.
#include "common.cuh"

#define FC_THREADS 256
#define FC_PPT 32

__device__ __forceinline__ int fc_code(uint8_t b) {
    const uint8_t x = b & 0xDF;
    if (x == 'A') return 0;
    if (x == 'C') return 1;
    if (x == 'G') return 2;
    if (x == 'T') return 3;
    return -1;
}

// slots: {key lo, key hi, count, pad}; key == FR_EMPTY_KEY marks a free slot
__global__ void __launch_bounds__(FC_THREADS)
fr_kmer_count_kernel(const uint8_t *__restrict__ bases, const int64_t *__restrict__ offsets, int64_t n_reads, int K,
                     uint64_t threshold, uint64_t seed, uint4 *slots, int lg_cap, unsigned long long *counters) {
    const uint64_t kmask = (1ull << (2 * K)) - 1;
    const uint32_t cap_mask = (uint32_t)((1ull << lg_cap) - 1);
    const int rcsh = 2 * (K - 1);
    for (int64_t r = blockIdx.x; r < n_reads; r += gridDim.x) {
        const int64_t s = offsets[r], e = offsets[r + 1];
        for (int64_t c0 = s + (int64_t)FC_PPT * threadIdx.x; c0 < e; c0 += (int64_t)FC_PPT * FC_THREADS) {
            uint64_t fw = 0, rc = 0;
            int run = 0;
            int64_t p = c0 - (K - 1);
            if (p < s) p = s;
            const int64_t stop = c0 + FC_PPT < e ? c0 + FC_PPT : e;
            for (; p < stop; ++p) {
                const int c = fc_code(__ldg(bases + p));
                if (c < 0) { run = 0; fw = 0; rc = 0; continue; }
                fw = ((fw << 2) | (uint64_t)c) & kmask;
                rc = (rc >> 2) | ((uint64_t)(3 - c) << rcsh);
                if (++run < K || p < c0) continue;
                const uint64_t can = fw < rc ? fw : rc;
                if (!fr_sample_keep(can, seed, threshold)) continue;
                uint32_t slot = fr_table_slot(fr_hash(can), lg_cap);
                int tries = 0;
                while (true) {
                    unsigned long long *kp = reinterpret_cast<unsigned long long *>(&slots[slot]);
                    unsigned long long old = *kp;
                    if (old == FR_EMPTY_KEY) old = atomicCAS(kp, (unsigned long long)FR_EMPTY_KEY, (unsigned long long)can);
                    if (old == FR_EMPTY_KEY) { atomicAdd(&counters[0], 1ull); old = can; }
                    if (old == can) { atomicAdd(&slots[slot].z, 1u); break; }
                    slot = (slot + 1) & cap_mask;
                    if (++tries > 4096) { atomicOr(&counters[1], 1ull); break; }   // table (nearly) full
                }
            }
        }
    }
}

__global__ void fr_kmer_export_kernel(const uint4 *__restrict__ slots, int64_t cap, uint32_t min_mult,
                                      uint64_t *codes_out, uint32_t *counts_out, int64_t out_cap,
                                      unsigned long long *counters) {
    const int lane = threadIdx.x & 31;
    for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~31ll; i0 < cap;
         i0 += ((int64_t)gridDim.x * blockDim.x)) {
        const int64_t i = i0 + lane;
        uint4 v = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0, 0);
        if (i < cap) v = slots[i];
        const uint64_t key = ((uint64_t)v.y << 32) | v.x;
        const bool keep = key != FR_EMPTY_KEY && v.z >= min_mult;
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, keep);
        if (!bal) continue;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&counters[2], (unsigned long long)__popc(bal));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (keep) {
            const unsigned long long o = base + __popc(bal & ((1u << lane) - 1));
            if ((int64_t)o < out_cap) { codes_out[o] = key; counts_out[o] = v.z; }
        }
    }
}

__global__ void fr_kmer_table_init_kernel(uint4 *slots, int64_t cap) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += (int64_t)gridDim.x * blockDim.x)
        slots[i] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u);
}

extern "C" int fr_kmer_count_begin(fr_ctx *ctx, int k, int64_t capacity, void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (k < 1 || k > 31) return fr_fail(ctx, FR_ERR_ARG, "fr_kmer_count_begin: k must be in [1,31]");
    cudaStream_t st = (cudaStream_t)stream;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    int lg = 10;
    while ((1ll << lg) < capacity && lg < 34) ++lg;
    int64_t capb = ctx->count_cap_bytes;
    int rc = fr_ensure(ctx, (void **)&ctx->d_count_slots, &capb, (int64_t)16 << lg, "k-mer count table");
    ctx->count_cap_bytes = capb;
    if (rc) return rc;
    ctx->count_lg_cap = lg;
    ctx->count_k = k;
    fr_kt_begin(ctx, "fr_kmer_table_init_kernel", st);
    fr_kmer_table_init_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(ctx->d_count_slots, 1ll << lg);
    FR_LAUNCH_CHECK(ctx, "fr_kmer_table_init_kernel");
    FR_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, 8 * sizeof(unsigned long long), st));
    return FR_OK;
}

extern "C" int fr_kmer_count_add(fr_ctx *ctx, const uint8_t *d_bases, const int64_t *d_offsets, int64_t n_reads,
                                 uint64_t threshold, uint64_t seed, int64_t *distinct_out, void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (!ctx->d_count_slots || ctx->count_k == 0)
        return fr_fail(ctx, FR_ERR_STATE, "fr_kmer_count_add: call fr_kmer_count_begin first");
    cudaStream_t st = (cudaStream_t)stream;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n_reads > 0) {
        int64_t grid = n_reads < (int64_t)ctx->sm_count * 32 ? n_reads : (int64_t)ctx->sm_count * 32;
        fr_kt_begin(ctx, "fr_kmer_count_kernel", st);
        fr_kmer_count_kernel<<<(unsigned)grid, FC_THREADS, 0, st>>>(d_bases, d_offsets, n_reads, ctx->count_k, threshold,
                                                                    seed, ctx->d_count_slots, ctx->count_lg_cap,
                                                                    ctx->d_counters);
        FR_LAUNCH_CHECK(ctx, "fr_kmer_count_kernel");
    }
    FR_CUDA(ctx, cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 8 * sizeof(unsigned long long),
                                 cudaMemcpyDeviceToHost, st));
    FR_CUDA(ctx, cudaStreamSynchronize(st));
    if (distinct_out) *distinct_out = (int64_t)ctx->h_counters[0];
    if (ctx->h_counters[1]) return fr_fail(ctx, FR_ERR_NOMEM, "fr_kmer_count_add: count table full (begin again with a larger capacity)");
    return FR_OK;
}

extern "C" int fr_kmer_count_export(fr_ctx *ctx, int min_multiplicity, uint64_t *d_codes_out, uint32_t *d_counts_out,
                                    int64_t cap, int64_t *n_out, void *stream) {
    if (!ctx || !n_out) return FR_ERR_ARG;
    if (!ctx->d_count_slots || ctx->count_k == 0)
        return fr_fail(ctx, FR_ERR_STATE, "fr_kmer_count_export: no count table");
    cudaStream_t st = (cudaStream_t)stream;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    FR_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 2, 0, sizeof(unsigned long long), st));
    fr_kt_begin(ctx, "fr_kmer_export_kernel", st);
    fr_kmer_export_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(ctx->d_count_slots, 1ll << ctx->count_lg_cap,
                                                             (uint32_t)(min_multiplicity < 1 ? 1 : min_multiplicity),
                                                             d_codes_out, d_counts_out, d_codes_out ? cap : 0,
                                                             ctx->d_counters);
    FR_LAUNCH_CHECK(ctx, "fr_kmer_export_kernel");
    FR_CUDA(ctx, cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 8 * sizeof(unsigned long long),
                                 cudaMemcpyDeviceToHost, st));
    FR_CUDA(ctx, cudaStreamSynchronize(st));
    *n_out = (int64_t)ctx->h_counters[2];
    return FR_OK;
}
