// K1 + K3 -- reads -> CSR feature matrix.
// Replaces the consumer loop of kmer_searcher.cpp:295-354 (roll :314-323 and
// :333-341, probe :325-326 / :343-344, per-read dedupe :327 / :345) and the
// result collection of :90-96.
//
// One CTA per read.  The read is walked in tiles of FR_TILE positions:
//   K1  the tile's ASCII is loaded with 16-byte vector loads and packed to
//       2-bit codes + a 1-bit "non-ACGT" mask in shared memory;
//   K3a every thread rolls the reference's k-mer register over 32 consecutive
//       positions (replaying the reference's state machine exactly, including
//       the UINT64_MAX quirk), tests each k-mer against the blocked Bloom
//       filter (one 32-byte L2 sector per probe) and queues the few survivors;
//   K3b the queue is drained against the open-addressing table by full warps
//       (no divergence on the rare hit path); hits go to a per-read list;
//   K3c at the end of the read the list is sorted, de-duplicated and written
//       to a bump-allocated scratch row; document frequencies are counted.
#include "common.cuh"

#define FR_THREADS 256
#define FR_PPT 32                       // positions per thread per tile
#define FR_TILE (FR_THREADS * FR_PPT)   // 8192 positions
#define FR_ROUNDS 4
#define FR_PPR (FR_PPT / FR_ROUNDS)     // positions per thread per round
#define FR_QCAP 4096                    // candidate queue (k-mers)
#define FR_HCAP 4096                    // shared-memory hit list
#define FR_NWORDS (FR_TILE / 16 + 2)    // packed words incl. 32 positions of look-behind

#define FR_FLAG_SCRATCH_OVERFLOW 1ull
#define FR_FLAG_SPILL_OVERFLOW 2ull

// ---- K1: 16 ASCII bytes -> 32 bits of codes + 16 invalid bits ----
// code = ((b>>1)^(b>>2))&3 maps A/a 0, C/c 1, G/g 2, T/t 3 (kmer_searcher.cpp:316-322)
__device__ __forceinline__ void fr_pack4(uint32_t w, uint32_t &codes8, uint32_t &inv4) {
    uint32_t t = ((w >> 1) ^ (w >> 2)) & 0x03030303u;
    codes8 = (t | (t >> 6) | (t >> 12) | (t >> 18)) & 0xFFu;
    uint32_t x = w & 0xDFDFDFDFu;  // fold case
    uint32_t ok = __vcmpeq4(x, 0x41414141u) | __vcmpeq4(x, 0x43434343u) |
                  __vcmpeq4(x, 0x47474747u) | __vcmpeq4(x, 0x54545454u);
    uint32_t bad = ~ok;
    inv4 = (bad & 1u) | ((bad >> 7) & 2u) | ((bad >> 14) & 4u) | ((bad >> 21) & 8u);
}

__device__ __forceinline__ void fr_pack16(uint4 v, uint32_t &codes, uint32_t &inv) {
    uint32_t c0, c1, c2, c3, i0, i1, i2, i3;
    fr_pack4(v.x, c0, i0);
    fr_pack4(v.y, c1, i1);
    fr_pack4(v.z, c2, i2);
    fr_pack4(v.w, c3, i3);
    codes = c0 | (c1 << 8) | (c2 << 16) | (c3 << 24);
    inv = i0 | (i1 << 4) | (i2 << 8) | (i3 << 12);
}

// ---- block-wide bitonic sort of n u32 (n padded with 0xFFFFFFFF to a power of two) ----
__device__ void fr_block_sort(uint32_t *a, int n_pow2) {
    for (int k = 2; k <= n_pow2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n_pow2; i += FR_THREADS) {
                int ixj = i ^ j;
                if (ixj > i) {
                    uint32_t x = a[i], y = a[ixj];
                    bool up = (i & k) == 0;
                    if ((x > y) == up) { a[i] = y; a[ixj] = x; }
                }
            }
            __syncthreads();
        }
    }
}

// sort + unique of a[0..n); returns the unique count (uniform across the block).
// `cap` is the usable capacity of a (>= next power of two of n).
__device__ int fr_block_sort_unique(uint32_t *a, int n, int *s_scan /*[FR_THREADS/32 + 2]*/) {
    if (n <= 1) return n;
    int p2 = 1;
    while (p2 < n) p2 <<= 1;
    for (int i = n + threadIdx.x; i < p2; i += FR_THREADS) a[i] = 0xFFFFFFFFu;
    __syncthreads();
    fr_block_sort(a, p2);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int out_base = 0;
    for (int c = 0; c < n; c += FR_THREADS) {
        int i = c + threadIdx.x;
        uint32_t v = 0, pv = 0;
        bool keep = false;
        if (i < n) {
            v = a[i];
            keep = (i == 0);
            if (i > 0) { pv = a[i - 1]; keep = (v != pv); }
        }
        unsigned bal = __ballot_sync(0xFFFFFFFFu, keep);
        int wrank = __popc(bal & ((1u << lane) - 1));
        if (lane == 0) s_scan[warp] = __popc(bal);
        __syncthreads();  // all reads of this chunk done, warp totals visible
        int woff = 0, total = 0;
        for (int w = 0; w < FR_THREADS / 32; ++w) {
            int t = s_scan[w];
            if (w < warp) woff += t;
            total += t;
        }
        if (keep) a[out_base + woff + wrank] = v;
        out_base += total;
        __syncthreads();
    }
    return out_base;
}

struct FrExtractShared {
    uint64_t queue[FR_QCAP];
    uint32_t hits[FR_HCAP];
    uint32_t codes[FR_NWORDS];
    uint32_t inv[FR_NWORDS];
    int scan[FR_THREADS / 32 + 2];
    int qcount;
    int hcount;
    long long misc;
};

struct FrHitList {
    uint32_t *ptr;  // shared memory first, a global spill slab for very long reads
    int cap;
    bool dead;      // spill pool exhausted: the host grows it and reruns
};

// K3b: drain the candidate queue against the open-addressing table.
__device__ void fr_flush(FrExtractShared &sh, FrHitList &hl, const FrTable &tb, int64_t n_lookups,
                         uint32_t *spill, int64_t spill_cap, unsigned long long *counters) {
    const int n = sh.qcount;
    int hc = sh.hcount;
    __syncthreads();  // everyone has read the counters before anyone moves them
    if (n == 0) return;
    if (hc + n > hl.cap && !hl.dead) {
        int u = fr_block_sort_unique(hl.ptr, hc, sh.scan);
        if (threadIdx.x == 0) sh.hcount = u;
        __syncthreads();
        hc = u;
        if (hc + n > hl.cap) {
            // migrate to a global slab that can hold one entry per lookup position
            long long need = 1;
            while (need < n_lookups) need <<= 1;
            if (threadIdx.x == 0) sh.misc = (long long)atomicAdd(&counters[1], (unsigned long long)need);
            __syncthreads();
            long long off = sh.misc;
            if (off + need > spill_cap) {
                if (threadIdx.x == 0) atomicOr(&counters[2], FR_FLAG_SPILL_OVERFLOW);
                hl.dead = true;
            } else {
                uint32_t *dst = spill + off;
                for (int i = threadIdx.x; i < hc; i += FR_THREADS) dst[i] = hl.ptr[i];
                hl.ptr = dst;
                hl.cap = (int)(need > 0x40000000ll ? 0x40000000ll : need);
            }
            __syncthreads();
        }
    }
    const uint32_t cap_mask = (uint32_t)((1ull << tb.lg_cap) - 1);
    for (int i = threadIdx.x; i < n; i += FR_THREADS) {
        uint64_t kmer = sh.queue[i];
        uint32_t slot = fr_table_slot(fr_hash(kmer), tb.lg_cap);
        while (true) {
            uint64_t key = __ldg(&tb.keys[slot]);
            if (key == kmer) {
                uint32_t idx = __ldg(&tb.vals[slot]);
                if (!hl.dead) {
                    int pos = atomicAdd(&sh.hcount, 1);
                    hl.ptr[pos] = idx;
                }
                break;
            }
            if (key == FR_EMPTY_KEY) break;
            slot = (slot + 1) & cap_mask;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) sh.qcount = 0;
    __syncthreads();
}

__global__ void __launch_bounds__(FR_THREADS, 4)
fr_extract_kernel(const uint8_t *__restrict__ bases, int64_t n_bases, const int64_t *__restrict__ offsets,
                  int64_t n_reads, FrTable tb, int32_t *scratch, int64_t scratch_cap, uint32_t *spill,
                  int64_t spill_cap, int64_t *read_off, int32_t *read_cnt, unsigned long long *counters) {
    extern __shared__ __align__(16) unsigned char fr_extract_smem[];
    FrExtractShared &sh = *reinterpret_cast<FrExtractShared *>(fr_extract_smem);
    const int tid = threadIdx.x, lane = tid & 31;
    const int K = tb.k;
    const uint64_t kmask = (1ull << (2 * K)) - 1;
    for (int64_t r = blockIdx.x; r < n_reads; r += gridDim.x) {
        const int64_t s = offsets[r], e = offsets[r + 1];
        const int64_t len = e - s;
        const int64_t n_lookups = len >= K ? len - K + 1 : 1;
        const int64_t lo = len >= K ? s + K - 1 : e - 1;  // first lookup position
        FrHitList hl;
        hl.ptr = sh.hits; hl.cap = FR_HCAP; hl.dead = false;
        if (tid == 0) {
            sh.hcount = 0;
            sh.qcount = 0;
            if (len == 0) { sh.queue[0] = 0; sh.qcount = 1; }  // kmer_searcher.cpp:325 with an empty read
        }
        __syncthreads();
        const int64_t abase = s & ~15ll;
        const int64_t ntiles = len > 0 ? (e - abase + FR_TILE - 1) / FR_TILE : 0;
        for (int64_t tile = 0; tile < ntiles; ++tile) {
            const int64_t tile_start = abase + tile * FR_TILE;
            // ---- K1: pack this tile (plus 32 positions of look-behind) ----
            for (int w = tid; w < FR_NWORDS; w += FR_THREADS) {
                const int64_t g = tile_start - 32 + 16ll * w;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (g + 16 > s && g < e) {
                    if (g >= 0 && g + 16 <= n_bases) {
                        v = __ldg(reinterpret_cast<const uint4 *>(bases + g));
                    } else {
                        uint32_t b[4] = {0, 0, 0, 0};
                        for (int j = 0; j < 16; ++j) {
                            int64_t p = g + j;
                            if (p >= 0 && p < n_bases) b[j >> 2] |= (uint32_t)bases[p] << (8 * (j & 3));
                        }
                        v = make_uint4(b[0], b[1], b[2], b[3]);
                    }
                }
                uint32_t c, iv;
                fr_pack16(v, c, iv);
                sh.codes[w] = c;
                sh.inv[w] = iv;
            }
            __syncthreads();
            // ---- K3a: roll + filter ----
            const int64_t cstart = tile_start + (int64_t)FR_PPT * tid;
            const bool active = cstart < e && cstart + FR_PPT > lo;
            const bool warp_active = __any_sync(0xFFFFFFFFu, active);
            uint64_t cur = 0, cw = 0;
            uint32_t ivw = 0;
            if (active) {
                const int wi = 2 * tid;
                uint64_t cww = (uint64_t)sh.codes[wi] | ((uint64_t)sh.codes[wi + 1] << 32);
                uint32_t ivv = sh.inv[wi] | (sh.inv[wi + 1] << 16);
                cw = (uint64_t)sh.codes[wi + 2] | ((uint64_t)sh.codes[wi + 3] << 32);
                ivw = sh.inv[wi + 2] | (sh.inv[wi + 3] << 16);
                const int64_t p0 = cstart - 32;
                for (int j = 32 - (K - 1); j < 32; ++j) {
                    if (p0 + j >= s) {
                        uint64_t c = (cww >> (2 * j)) & 3ull;
                        cur = ((ivv >> j) & 1u) ? FR_EMPTY_KEY : (((cur << 2) & kmask) | c);
                    }
                }
            }
#pragma unroll 1
            for (int round = 0; round < FR_ROUNDS; ++round) {
                if (warp_active) {
#pragma unroll
                    for (int jj = 0; jj < FR_PPR; ++jj) {
                        const int j = round * FR_PPR + jj;
                        const int64_t p = cstart + j;
                        bool pass = false;
                        if (active && p >= s) {
                            uint64_t c = (cw >> (2 * j)) & 3ull;
                            cur = ((ivw >> j) & 1u) ? FR_EMPTY_KEY : (((cur << 2) & kmask) | c);
                            if (p >= lo && p < e && cur != FR_EMPTY_KEY) {
                                uint64_t h = fr_hash(cur);
                                uint32_t word = __ldg(&tb.bloom[fr_bloom_word(h, tb.lg_words)]);
                                uint32_t m = fr_bloom_mask(h);
                                pass = (word & m) == m;
                            }
                        }
                        unsigned bal = __ballot_sync(0xFFFFFFFFu, pass);
                        if (bal) {
                            int base = 0;
                            const int leader = __ffs(bal) - 1;
                            if (lane == leader) base = atomicAdd(&sh.qcount, __popc(bal));
                            base = __shfl_sync(0xFFFFFFFFu, base, leader);
                            if (pass) sh.queue[base + __popc(bal & ((1u << lane) - 1))] = cur;
                        }
                    }
                }
                // barrier + uniform decision in one: the last thread to arrive sees the final count
                if (__syncthreads_or(sh.qcount > FR_QCAP - FR_THREADS * FR_PPR))
                    fr_flush(sh, hl, tb, n_lookups, spill, spill_cap, counters);
            }
        }
        // ---- K3b/K3c: final drain, sort + unique, emit ----
        fr_flush(sh, hl, tb, n_lookups, spill, spill_cap, counters);
        int u = fr_block_sort_unique(hl.ptr, sh.hcount, sh.scan);
        if (hl.dead) u = 0;
        if (tid == 0) sh.misc = (long long)atomicAdd(&counters[0], (unsigned long long)u);
        __syncthreads();
        const long long off = sh.misc;
        if (off + u <= scratch_cap) {
            for (int i = tid; i < u; i += FR_THREADS) scratch[off + i] = (int32_t)hl.ptr[i];
        } else if (tid == 0) {
            atomicOr(&counters[2], FR_FLAG_SCRATCH_OVERFLOW);
        }
        if (tid == 0) { read_off[r] = off; read_cnt[r] = u; }
        __syncthreads();
    }
}

// ---- exclusive scan of per-read counts -> indptr (single CTA, chunked) ----
__global__ void fr_scan_kernel(const int32_t *__restrict__ cnt, int64_t n, int64_t *indptr) {
    __shared__ long long s_warp[32];
    __shared__ long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { s_carry = 0; indptr[0] = 0; }
    __syncthreads();
    for (int64_t base = 0; base < n; base += blockDim.x) {
        int64_t i = base + tid;
        long long v = i < n ? (long long)cnt[i] : 0, x = v;
        for (int d = 1; d < 32; d <<= 1) {
            long long y = __shfl_up_sync(0xFFFFFFFFu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) s_warp[warp] = x;
        __syncthreads();
        if (warp == 0) {
            long long w = lane < (int)(blockDim.x >> 5) ? s_warp[lane] : 0;
            for (int d = 1; d < 32; d <<= 1) {
                long long y = __shfl_up_sync(0xFFFFFFFFu, w, d);
                if (lane >= d) w += y;
            }
            s_warp[lane] = w;  // inclusive over warps
        }
        __syncthreads();
        long long incl = x + (warp > 0 ? s_warp[warp - 1] : 0) + s_carry;
        if (i < n) indptr[i + 1] = incl;
        __syncthreads();
        if (tid == (int)blockDim.x - 1) s_carry = incl;
        __syncthreads();
    }
}

// ---- scratch rows (allocation order) -> final CSR (input order) ----
__global__ void fr_gather_kernel(const int32_t *__restrict__ scratch, const int64_t *__restrict__ read_off,
                                 const int64_t *__restrict__ indptr, int64_t n_reads, int32_t *indices) {
    const int lane = threadIdx.x & 31;
    int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; r < n_reads; r += nwarps) {
        const int64_t a = indptr[r], n = indptr[r + 1] - a, off = read_off[r];
        for (int64_t i = lane; i < n; i += 32) indices[a + i] = scratch[off + i];
    }
}

// kmer_searcher.cpp:327-328, 345-346, 56-86: per-feature document frequency.
__global__ void fr_docfreq_kernel(const int32_t *__restrict__ scratch, int64_t nnz, uint32_t *docfreq) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < nnz; i += stride) atomicAdd(&docfreq[scratch[i]], 1u);
}

int fr_extract_launch(fr_ctx *ctx, const uint8_t *d_bases, int64_t n_bases, const int64_t *d_offsets,
                      int64_t n_reads, uint32_t *d_docfreq, cudaStream_t st) {
    (void)d_docfreq;
    int64_t grid = n_reads < (int64_t)ctx->sm_count * 64 ? n_reads : (int64_t)ctx->sm_count * 64;
    if (grid < 1) grid = 1;
    static bool attr_set = false;
    if (!attr_set) {
        FR_CUDA(ctx, cudaFuncSetAttribute(fr_extract_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)sizeof(FrExtractShared)));
        attr_set = true;
    }
    fr_kt_begin(ctx, "fr_extract_kernel", st);
    fr_extract_kernel<<<(unsigned)grid, FR_THREADS, sizeof(FrExtractShared), st>>>(
        d_bases, n_bases, d_offsets, n_reads, ctx->table, ctx->d_scratch, ctx->scratch_cap, ctx->d_spill,
        ctx->spill_cap, ctx->d_read_off, ctx->d_read_cnt, ctx->d_counters);
    FR_LAUNCH_CHECK(ctx, "fr_extract_kernel");
    return FR_OK;
}

int fr_scan_counts(fr_ctx *ctx, const int32_t *d_cnt, int64_t n, int64_t *d_indptr, cudaStream_t st) {
    fr_kt_begin(ctx, "fr_scan_kernel", st);
    fr_scan_kernel<<<1, 1024, 0, st>>>(d_cnt, n, d_indptr);
    FR_LAUNCH_CHECK(ctx, "fr_scan_kernel");
    return FR_OK;
}

int fr_gather_rows(fr_ctx *ctx, int32_t *d_indices, cudaStream_t st) {
    int64_t n = ctx->last_n_reads;
    if (n <= 0) return FR_OK;
    int64_t blocks = (n * 32 + 255) / 256;
    if (blocks > (int64_t)ctx->sm_count * 32) blocks = (int64_t)ctx->sm_count * 32;
    fr_kt_begin(ctx, "fr_gather_kernel", st);
    fr_gather_kernel<<<(unsigned)blocks, 256, 0, st>>>(ctx->d_scratch, ctx->d_read_off, ctx->last_indptr, n,
                                                       d_indices);
    FR_LAUNCH_CHECK(ctx, "fr_gather_kernel");
    return FR_OK;
}

int fr_docfreq_launch(fr_ctx *ctx, int64_t nnz, uint32_t *d_docfreq, cudaStream_t st) {
    if (nnz <= 0) return FR_OK;
    int64_t blocks = (nnz + 255) / 256;
    if (blocks > (int64_t)ctx->sm_count * 16) blocks = (int64_t)ctx->sm_count * 16;
    fr_kt_begin(ctx, "fr_docfreq_kernel", st);
    fr_docfreq_kernel<<<(unsigned)blocks, 256, 0, st>>>(ctx->d_scratch, nnz, d_docfreq);
    FR_LAUNCH_CHECK(ctx, "fr_docfreq_kernel");
    return FR_OK;
}
