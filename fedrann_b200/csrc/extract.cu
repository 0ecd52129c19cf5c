// K1 + K3 -- reads -> CSR feature matrix.
// Replaces the consumer loop of kmer_searcher.cpp:295-354 (roll :314-323 and
// :333-341, probe :325-326 / :343-344, per-read dedupe :327 / :345) and the
// result collection of :90-96.
//
// One CTA (128 threads, 8 resident per SM: the phases of a read -- pack, filter, probe, sort --
// are separated by block barriers, and it is the number of reads in flight per SM that hides
// their latencies) per read.  The read is walked in tiles of FR_TILE = 4096 positions:
//   K1  the tile's ASCII is loaded with 16-byte vector loads and packed to
//       2-bit codes (first base in the most significant bits, so that a k-mer
//       is a funnel shift of two packed words) + a 1-bit "non-ACGT" mask;
//   K3a every thread owns 32 consecutive positions; their k-mers are cut out of
//       the packed words in registers (no loop-carried roll), hashed with one
//       32-bit multiply and tested against the one-bit filter (one 32-byte L2
//       sector per probe; the 32 probes of a thread are independent loads).
//       The result is a 32-bit pass mask per thread; the tile positions of the
//       survivors are queued in shared memory.  Positions whose window holds a
//       non-ACGT base skip the filter and are queued unconditionally;
//   K3b the queue is drained against the open-addressing table: the k-mer is
//       rebuilt from the packed tile with the reference's exact semantics
//       (UINT64_MAX quirk: bases at or before the last non-ACGT base of the
//       window read as T; reads shorter than k are left-padded with A);
//   K3c at the end of the read the hit list is sorted and de-duplicated: a
//       counting pass splits it into value-range buckets of <= 128 entries and
//       every warp sorts whole buckets in registers (shuffle bitonic network) --
//       no block barrier inside the sort.  The row goes to a bump-allocated
//       scratch CSR.
#include "common.cuh"

#include <cstdlib>
#include <cstring>

#define FR_THREADS 128
#define FR_PPT 32                       // positions per thread per tile
#define FR_TILE (FR_THREADS * FR_PPT)   // 4096 positions
#define FR_QCAP 1024                    // deferred k-mer queue; drained when a tile might not fit
#define FR_HCAP 2048                    // shared-memory hit list
#define FR_NWORDS (FR_TILE / 16 + 2)    // packed words incl. 32 positions of look-behind
#define FR_MAXB 32                      // most sort buckets (FR_HCAP / 64)
#define FR_BCAP 128                     // entries one warp sorts in registers

#define FR_FLAG_SCRATCH_OVERFLOW 1ull
#define FR_FLAG_SPILL_OVERFLOW 2ull

// ---- K1: 16 ASCII bytes -> 32 bits of codes + 16 invalid bits, first base most significant ----
// code = ((b>>1)^(b>>2))&3 maps A/a 0, C/c 1, G/g 2, T/t 3 (kmer_searcher.cpp:316-322)
__device__ __forceinline__ void fr_pack4(uint32_t w, uint32_t &codes8, uint32_t &inv4) {
    uint32_t t = ((w >> 1) ^ (w >> 2)) & 0x03030303u;
    codes8 = ((t << 6) | (t >> 4) | (t >> 14) | (t >> 24)) & 0xFFu;
    uint32_t x = w & 0xDFDFDFDFu;  // fold case
    uint32_t ok = __vcmpeq4(x, 0x41414141u) | __vcmpeq4(x, 0x43434343u) |
                  __vcmpeq4(x, 0x47474747u) | __vcmpeq4(x, 0x54545454u);
    uint32_t bad = ~ok;
    inv4 = ((bad << 3) & 8u) | ((bad >> 6) & 4u) | ((bad >> 15) & 2u) | ((bad >> 24) & 1u);
}

__device__ __forceinline__ void fr_pack16(uint4 v, uint32_t &codes, uint32_t &inv) {
    uint32_t c0, c1, c2, c3, i0, i1, i2, i3;
    fr_pack4(v.x, c0, i0);
    fr_pack4(v.y, c1, i1);
    fr_pack4(v.z, c2, i2);
    fr_pack4(v.w, c3, i3);
    codes = (c0 << 24) | (c1 << 16) | (c2 << 8) | c3;
    inv = (i0 << 12) | (i1 << 8) | (i2 << 4) | i3;
}

// ---- block-wide bitonic sort of n u32 (n padded with 0xFFFFFFFF to a power of two) ----
// (general path: hit lists that spilled to global memory, mid-read compactions, skewed buckets)
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
    uint32_t hits[FR_HCAP];
    uint32_t tmp[FR_HCAP];      // K3c: hits grouped by bucket
    uint32_t codes[FR_NWORDS];
    uint32_t inv[FR_NWORDS];
    uint32_t rcodes[FR_NWORDS]; // complemented codes, first base least significant (sampled libraries only)
    uint64_t kq[FR_QCAP];       // deferred k-mers awaiting their table probe
    int bcnt[FR_MAXB + 1];      // K3c: bucket sizes -> unique counts
    int bstart[FR_MAXB + 1];
    int bfill[FR_MAXB + 1];     // scatter cursors -> output offsets
    int scan[FR_THREADS / 32 + 2];
    int hcount;
    int qcount;
    int tcount;                 // survivors of the current tile
    int flag;
    long long misc;
};

// ---- K3c fast path: sort + unique of sh.hits[0..n), n <= FR_HCAP ----
// 128 values per warp, 4 per lane (element e = 4*lane + r), ascending.
__device__ __forceinline__ void fr_warp_sort128(uint32_t (&v)[4], int lane) {
#pragma unroll
    for (int k = 2; k <= 128; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= 4) {
                const int lj = j >> 2;
                const bool up = ((lane << 2) & k) == 0;
                const bool lower = (lane & lj) == 0;
                const bool keepmin = up == lower;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const uint32_t o = __shfl_xor_sync(0xFFFFFFFFu, v[r], lj);
                    v[r] = keepmin ? min(v[r], o) : max(v[r], o);
                }
            } else {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const int r2 = r ^ j;
                    if (r2 > r) {
                        const bool up = (((lane << 2) | r) & k) == 0;
                        const uint32_t a = v[r], b = v[r2];
                        const uint32_t mn = min(a, b), mx = max(a, b);
                        v[r] = up ? mn : mx;
                        v[r2] = up ? mx : mn;
                    }
                }
            }
        }
    }
}

__device__ __forceinline__ int fr_warp_excl_scan(int x, int lane, int &total) {
    int incl = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int y = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += y;
    }
    total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    return incl - x;
}

// Returns the unique count, or -1 when a bucket is too large (caller takes the general path;
// sh.hits is untouched in that case).
__device__ int fr_bucket_sort_unique(FrExtractShared &sh, int n, uint32_t n_features) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int lgB = 0;
    while ((64 << lgB) < n) ++lgB;                                  // <= 64 entries per bucket on average
    while (lgB > 0 && (n_features >> lgB) == 0) --lgB;
    const int B = 1 << lgB;
    // bucket = floor(v * B / F) (approximately; only monotonicity and b < B matter)
    const unsigned long long m64 = ((0x100000000ull / n_features) << lgB);
    const uint32_t M = m64 > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)m64;
    if (tid <= B) sh.bcnt[tid] = 0;
    if (tid == 0) sh.flag = 0;
    __syncthreads();
    for (int i = tid; i < n; i += FR_THREADS) atomicAdd(&sh.bcnt[__umulhi(sh.hits[i], M)], 1);
    __syncthreads();
    if (warp == 0) {
        int carry = 0;
        for (int b0 = 0; b0 < B; b0 += 32) {
            const int c = b0 + lane < B ? sh.bcnt[b0 + lane] : 0;
            int tot;
            const int ex = fr_warp_excl_scan(c, lane, tot);
            if (b0 + lane < B) { sh.bstart[b0 + lane] = carry + ex; sh.bfill[b0 + lane] = carry + ex; }
            if (c > FR_BCAP) sh.flag = 1;
            carry += tot;
        }
        if (lane == 0) sh.bstart[B] = carry;
    }
    __syncthreads();
    if (sh.flag) return -1;
    for (int i = tid; i < n; i += FR_THREADS) {
        const uint32_t v = sh.hits[i];
        sh.tmp[atomicAdd(&sh.bfill[__umulhi(v, M)], 1)] = v;
    }
    __syncthreads();
    for (int b = warp; b < B; b += FR_THREADS / 32) {
        const int a = sh.bstart[b], cnt = sh.bstart[b + 1] - a;
        uint32_t v[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) v[r] = 4 * lane + r < cnt ? sh.tmp[a + 4 * lane + r] : 0xFFFFFFFFu;
        fr_warp_sort128(v, lane);
        uint32_t prev = __shfl_up_sync(0xFFFFFFFFu, v[3], 1);
        int c = 0;
        bool keep[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int e = 4 * lane + r;
            keep[r] = e < cnt && (e == 0 || v[r] != prev);
            c += keep[r];
            prev = v[r];
        }
        int tot;
        int o = a + fr_warp_excl_scan(c, lane, tot);
        __syncwarp();
#pragma unroll
        for (int r = 0; r < 4; ++r)
            if (keep[r]) sh.tmp[o++] = v[r];
        if (lane == 0) sh.bcnt[b] = tot;
    }
    __syncthreads();
    if (warp == 0) {
        int carry = 0;
        for (int b0 = 0; b0 < B; b0 += 32) {
            const int c = b0 + lane < B ? sh.bcnt[b0 + lane] : 0;
            int tot;
            const int ex = fr_warp_excl_scan(c, lane, tot);
            if (b0 + lane < B) sh.bfill[b0 + lane] = carry + ex;
            carry += tot;
        }
        if (lane == 0) sh.bfill[B] = carry;
    }
    __syncthreads();
    for (int b = warp; b < B; b += FR_THREADS / 32) {
        const int a = sh.bstart[b], o = sh.bfill[b], c = sh.bcnt[b];
        for (int i = lane; i < c; i += 32) sh.hits[o + i] = sh.tmp[a + i];
    }
    const int u = sh.bfill[B];
    __syncthreads();
    return u;
}

struct FrHitList {
    uint32_t *ptr;  // shared memory first, a global spill slab for very long reads
    int cap;
    bool dead;      // spill pool exhausted: the host grows it and reruns
};

// The k-mer the reference looks up at tile position `pos` (kmer_searcher.cpp:314-349), rebuilt
// from the packed tile.  Returns false when the base at `pos` is not ACGT (no lookup, :325/:343).
__device__ __forceinline__ bool fr_kmer_at(const FrExtractShared &sh, int pos, int K, uint64_t kmask,
                                            uint64_t &kmer) {
    const int u = pos + 32, W = u >> 4, r = u & 15;
    const uint32_t w2 = sh.codes[W], w1 = sh.codes[W - 1], w0 = sh.codes[W - 2];
    const int shc = 30 - 2 * r;
    const uint32_t lo = __funnelshift_r(w2, w1, shc), hi = __funnelshift_r(w1, w0, shc);
    kmer = (((uint64_t)hi << 32) | lo) & kmask;
    const uint64_t iv = (((uint64_t)sh.inv[W - 2] << 32) | ((uint64_t)sh.inv[W - 1] << 16) | sh.inv[W]) >> (15 - r);
    const uint32_t ivw = (uint32_t)iv & (uint32_t)((1ull << K) - 1);   // bit d = base pos - d
    if (ivw & 1u) return false;
    if (ivw) {
        const int d = __ffs(ivw) - 1;                 // last non-ACGT base of the window is d bases back:
        kmer = (kmer | ~((1ull << (2 * d)) - 1)) & kmask;   // it and everything before it read as T
    }
    return true;
}

// append one hit to the read's list (room was made before the probes were issued)
__device__ __forceinline__ void fr_append_hit(FrExtractShared &sh, const FrHitList &hl, uint32_t idx) {
    if (hl.dead) return;
    const int pos = atomicAdd(&sh.hcount, 1);
    if (pos < hl.cap) hl.ptr[pos] = idx;
}

__device__ __forceinline__ void fr_probe_table(const FrTable &tb, uint64_t kmer, FrExtractShared &sh,
                                               const FrHitList &hl) {
    const uint32_t cap_mask = (uint32_t)((1ull << tb.lg_cap) - 1);
    uint32_t slot = fr_table_slot(fr_hash(kmer), tb.lg_cap);
    while (true) {
        uint64_t key;
        uint32_t val;
        if (tb.vbits) {
            const uint64_t w = __ldg(&tb.cslots[slot]);
            if (w == FR_EMPTY_KEY) return;
            key = w >> tb.vbits;
            val = (uint32_t)(w & ((1ull << tb.vbits) - 1));
        } else {
            const uint4 v = __ldg(&tb.slots[slot]);
            key = ((uint64_t)v.y << 32) | v.x;
            if (key == FR_EMPTY_KEY) return;
            val = v.z;
        }
        if (key == kmer) { fr_append_hit(sh, hl, val); return; }
        slot = (slot + 1) & cap_mask;
    }
}

// Makes room for `n` more hits (block-uniform): sort + unique of what is there, and if that is
// not enough the list moves to a global slab that can hold one entry per lookup of the read.
__device__ void fr_make_room(FrExtractShared &sh, FrHitList &hl, int n, int64_t n_lookups, uint32_t *spill,
                             int64_t spill_cap, unsigned long long *counters) {
    int hc = sh.hcount;
    __syncthreads();
    if (hc + n <= hl.cap || hl.dead) return;
    const int u = fr_block_sort_unique(hl.ptr, hc, sh.scan);
    if (threadIdx.x == 0) sh.hcount = u;
    __syncthreads();
    hc = u;
    if (hc + n <= hl.cap) return;
    long long need = 1;
    while (need < n_lookups) need <<= 1;
    if (threadIdx.x == 0) sh.misc = (long long)atomicAdd(&counters[1], (unsigned long long)need);
    __syncthreads();
    const long long off = sh.misc;
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

// K3b: the block drains the deferred k-mer queue against the table.  The first slot of FR_NB
// k-mers per thread is requested before any is examined: a probe is a random DRAM/L2 access and
// only the number of them in flight matters.
#define FR_NB 4
__device__ void fr_block_drain(FrExtractShared &sh, FrHitList &hl, const FrTable &tb, int64_t n_lookups,
                               uint32_t *spill, int64_t spill_cap, unsigned long long *counters) {
    const int n = sh.qcount;
    __syncthreads();
    if (n == 0) return;
    fr_make_room(sh, hl, n, n_lookups, spill, spill_cap, counters);
    const uint32_t cap_mask = (uint32_t)((1ull << tb.lg_cap) - 1);
    const int vb = tb.vbits;
    const uint64_t vmask = (1ull << vb) - 1;
    for (int i0 = 0; i0 < n; i0 += FR_THREADS * FR_NB) {
        uint64_t km[FR_NB];
        uint32_t slot[FR_NB];
        bool ok[FR_NB];
#pragma unroll
        for (int b = 0; b < FR_NB; ++b) {
            const int i = i0 + FR_THREADS * b + threadIdx.x;
            ok[b] = i < n;
            km[b] = ok[b] ? sh.kq[i] : 0ull;
            ok[b] = ok[b] && km[b] != FR_EMPTY_KEY;
            slot[b] = fr_table_slot(fr_hash(km[b]), tb.lg_cap);
        }
        if (vb) {
            uint64_t w[FR_NB];
#pragma unroll
            for (int b = 0; b < FR_NB; ++b) w[b] = ok[b] ? __ldg(&tb.cslots[slot[b]]) : FR_EMPTY_KEY;
#pragma unroll
            for (int b = 0; b < FR_NB; ++b) {
                uint64_t x = w[b];
                uint32_t sl = slot[b];
                while (x != FR_EMPTY_KEY) {
                    if ((x >> vb) == km[b]) { fr_append_hit(sh, hl, (uint32_t)(x & vmask)); break; }
                    sl = (sl + 1) & cap_mask;
                    x = __ldg(&tb.cslots[sl]);
                }
            }
        } else {
            uint4 v[FR_NB];
#pragma unroll
            for (int b = 0; b < FR_NB; ++b) v[b] = ok[b] ? __ldg(&tb.slots[slot[b]]) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int b = 0; b < FR_NB; ++b) {
                if (!ok[b]) continue;
                uint4 w = v[b];
                uint32_t sl = slot[b];
                while (true) {
                    const uint64_t key = ((uint64_t)w.y << 32) | w.x;
                    if (key == km[b]) { fr_append_hit(sh, hl, w.z); break; }
                    if (key == FR_EMPTY_KEY) break;
                    sl = (sl + 1) & cap_mask;
                    w = __ldg(&tb.slots[sl]);
                }
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) sh.qcount = 0;
    __syncthreads();
}

// K3a for one thread: 32 filter probes, all independent.  c0..c3 = packed codes of the 32
// look-behind positions (c0, c1) and of the thread's own 32 positions (c2, c3).
template <bool WIDE>
__device__ __forceinline__ uint32_t fr_filter32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t kmask_lo,
                                                uint32_t kmask_hi, const uint32_t *__restrict__ bloom, int wsh,
                                                int bsh) {
    uint32_t pass = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const int s = 62 - 2 * j;   // the k-mer ending at position j is (c0:c1:c2:c3) >> s
        uint32_t lo, hi = 0;
        if (s >= 32) {
            lo = __funnelshift_r(c2, c1, s - 32);
            if (WIDE) hi = __funnelshift_r(c1, c0, s - 32);
        } else {
            lo = __funnelshift_r(c3, c2, s);
            if (WIDE) hi = __funnelshift_r(c2, c1, s);
        }
        uint32_t h;
        if (WIDE) h = fr_bloom_hash(lo, hi & kmask_hi);
        else h = (lo & kmask_lo) * FR_BLOOM_MUL1;
        const uint32_t word = __ldg(bloom + (h >> wsh));
        const uint32_t t = __funnelshift_l(0u, word, h >> bsh);   // word << ((h >> bsh) & 31): tested bit -> sign
        pass = __funnelshift_l(t, pass, 1);                       // (pass << 1) | (t >> 31)
    }
    return pass;
}

// The same for a library that is a verified hash-threshold sample (FrTable::sampled): a position
// can only hit when its CANONICAL k-mer passes the sampler's test, which needs no memory access.
// The reverse complement comes from a second packed stream (complemented codes, first base in
// the LEAST significant bits): q0..q3 = that stream for the thread's 64 positions, already
// shifted right by 2*(33-K) bits, so that the reverse complement of the k-mer ending at
// position j is (q3:q2:q1:q0) >> 2j.
template <bool WIDE>
__device__ __forceinline__ uint32_t fr_filter32_sampled(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                        uint32_t q0, uint32_t q1, uint32_t q2, uint32_t q3,
                                                        uint32_t kmask_lo, uint32_t kmask_hi, uint64_t seed,
                                                        uint64_t threshold) {
    uint32_t pass = 0;
    const uint32_t seedmix = fr_sample_seedmix(seed), thr32 = (uint32_t)(threshold >> 32);
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const int s = 62 - 2 * j;
        uint32_t lo, hi = 0, rlo, rhi = 0;
        if (s >= 32) {
            lo = __funnelshift_r(c2, c1, s - 32);
            if (WIDE) hi = __funnelshift_r(c1, c0, s - 32);
        } else {
            lo = __funnelshift_r(c3, c2, s);
            if (WIDE) hi = __funnelshift_r(c2, c1, s);
        }
        if (2 * j >= 32) {
            rlo = __funnelshift_r(q1, q2, 2 * j - 32);
            if (WIDE) rhi = __funnelshift_r(q2, q3, 2 * j - 32);
        } else {
            rlo = __funnelshift_r(q0, q1, 2 * j);
            if (WIDE) rhi = __funnelshift_r(q1, q2, 2 * j);
        }
        uint32_t keep;
        if (WIDE) {
            const uint64_t fw = ((uint64_t)(hi & kmask_hi) << 32) | lo;
            const uint64_t rc = ((uint64_t)(rhi & kmask_hi) << 32) | rlo;
            const uint64_t can = fw < rc ? fw : rc;
            keep = fr_sample_hash((uint32_t)can, (uint32_t)(can >> 32), seedmix) < thr32 ? 1u : 0u;
        } else {
            const uint32_t can = min(lo & kmask_lo, rlo & kmask_lo);
            keep = fr_sample_hash(can, 0u, seedmix) < thr32 ? 1u : 0u;
        }
        pass = (pass << 1) | keep;
    }
    if (threshold == 0xFFFFFFFFFFFFFFFFull) pass = 0xFFFFFFFFu;     // fraction 1: everything is sampled
    return pass;
}

// K3a, core bitmap (FrTable::cores): the k-mers ending at W consecutive positions j .. j+W-1 all contain the
// c-mer (c = K - W + 1) that ends at position j, so ONE probe of the bitmap -- index = that core, or a hash of
// it when it is longer than 14 bases -- rejects W positions at once: 1/W random L2 sectors per position
// instead of one, and no hash arithmetic per position.  Returns the pass mask (bit 31-j = position j).
template <int W>
__device__ __forceinline__ uint32_t fr_core_groups(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint64_t cmask,
                                                   const uint32_t *__restrict__ cores, int lg, int hashed) {
    uint32_t pass = 0;
#pragma unroll
    for (int g = 0; g < 32 / W; ++g) {
        const int j = g * W, s = 62 - 2 * j;     // the core ending at position j is (c0:c1:c2:c3) >> s
        uint32_t lo, hi;
        if (s >= 32) { lo = __funnelshift_r(c2, c1, s - 32); hi = __funnelshift_r(c1, c0, s - 32); }
        else { lo = __funnelshift_r(c3, c2, s); hi = __funnelshift_r(c2, c1, s); }
        const uint64_t core = (((uint64_t)hi << 32) | lo) & cmask;
        const uint32_t idx = fr_core_index(core, lg, hashed);
        const uint32_t word = __ldg(cores + (idx >> 5));
        const uint32_t bit = (word >> (idx & 31)) & 1u;
        pass |= (0u - bit) & (((1u << W) - 1u) << (32 - W - j));
    }
    return pass;
}

__device__ __forceinline__ uint32_t fr_filter32_core(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                     const FrTable &tb) {
    const uint64_t cmask = tb.core_c >= 32 ? ~0ull : ((1ull << (2 * tb.core_c)) - 1);
    switch (tb.core_w) {
        case 1: return fr_core_groups<1>(c0, c1, c2, c3, cmask, tb.cores, tb.core_lg, tb.core_hashed);
        case 2: return fr_core_groups<2>(c0, c1, c2, c3, cmask, tb.cores, tb.core_lg, tb.core_hashed);
        case 4: return fr_core_groups<4>(c0, c1, c2, c3, cmask, tb.cores, tb.core_lg, tb.core_hashed);
        default: return fr_core_groups<8>(c0, c1, c2, c3, cmask, tb.cores, tb.core_lg, tb.core_hashed);
    }
}

// Second stage for a VERIFIED hash-threshold sample (FrTable::sampled): of the few positions the core bitmap
// let through, only those whose canonical k-mer passes the sampler's test can be in the library -- evaluated
// in registers, so that they never reach the table.
__device__ __forceinline__ uint32_t fr_refine_sampled(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t m,
                                                      int K, uint64_t kmask, uint64_t seed, uint64_t threshold) {
    if (threshold == 0xFFFFFFFFFFFFFFFFull) return m;
    const uint32_t seedmix = fr_sample_seedmix(seed), thr32 = (uint32_t)(threshold >> 32);
    uint32_t out = m;
    while (m) {
        const int j = __clz(m);
        m &= ~(0x80000000u >> j);
        const int s = 62 - 2 * j;
        uint32_t lo, hi;
        if (s >= 32) { lo = __funnelshift_r(c2, c1, s - 32); hi = __funnelshift_r(c1, c0, s - 32); }
        else { lo = __funnelshift_r(c3, c2, s); hi = __funnelshift_r(c2, c1, s); }
        const uint64_t fw = (((uint64_t)hi << 32) | lo) & kmask;
        uint64_t r = __brevll(~fw);                                     // reverse the bits ...
        r = ((r >> 1) & 0x5555555555555555ull) | ((r & 0x5555555555555555ull) << 1);   // ... each base's two bits back in order
        r >>= (64 - 2 * K);
        const uint64_t can = fw < r ? fw : r;
        if (!(fr_sample_hash((uint32_t)can, (uint32_t)(can >> 32), seedmix) < thr32)) out &= ~(0x80000000u >> j);
    }
    return out;
}

// CORE = true: first stage = core bitmap (any library), second stage = sampler test on the survivors when the
// library is a verified sample.  CORE = false: the per-position filters above (one-bit filter probe per
// position, or the sampler evaluated at every position) -- FR_EXTRACT_FILTER=old, for A/B measurements.
template <bool WIDE, bool SAMPLED, bool CORE>
__global__ void __launch_bounds__(FR_THREADS, 8)
fr_extract_kernel(const uint8_t *__restrict__ bases, int64_t n_bases, const int64_t *__restrict__ offsets,
                  int64_t n_reads, FrTable tb, int32_t *scratch, int64_t scratch_cap, uint32_t *spill,
                  int64_t spill_cap, int64_t *read_off, int32_t *read_cnt, unsigned long long *counters,
                  const int32_t *__restrict__ order, int dbg) {
    extern __shared__ __align__(16) unsigned char fr_extract_smem[];
    FrExtractShared &sh = *reinterpret_cast<FrExtractShared *>(fr_extract_smem);
    const int tid = threadIdx.x, lane = tid & 31;
    const int K = tb.k;
    const uint64_t kmask = (1ull << (2 * K)) - 1;
    const uint32_t kmask_lo = (uint32_t)kmask, kmask_hi = (uint32_t)(kmask >> 32);
    const int wsh = 32 - tb.lg_words, bsh = 27 - tb.lg_words;
    for (int64_t ri = blockIdx.x; ri < n_reads; ri += gridDim.x) {
        const int64_t r = order[ri];                       // longest reads first (fr_read_order_kernel)
        const int64_t s = offsets[r], e = offsets[r + 1];
        const int64_t len = e - s;
        const int64_t n_lookups = len >= K ? len - K + 1 : 1;
        const int64_t lo = len >= K ? s + K - 1 : e - 1;  // first lookup position
        FrHitList hl;
        hl.ptr = sh.hits; hl.cap = FR_HCAP; hl.dead = false;
        if (tid == 0) { sh.hcount = 0; sh.flag = 0; sh.qcount = 0; sh.tcount = 0; }
        __syncthreads();
        if (len == 0 && tid == 0) fr_probe_table(tb, 0ull, sh, hl);   // kmer_searcher.cpp:325 with an empty read
        const int64_t abase = s & ~15ll;
        const int64_t ntiles = len > 0 ? (e - abase + FR_TILE - 1) / FR_TILE : 0;
        for (int64_t tile = 0; tile < ntiles; ++tile) {
            const int64_t tile_start = abase + tile * FR_TILE;
            // ---- K1: pack this tile (plus 32 positions of look-behind); bases outside the read are A, valid ----
            for (int w = tid; w < FR_NWORDS; w += FR_THREADS) {
                const int64_t g = tile_start - 32 + 16ll * w;
                uint32_t c = 0, iv = 0;
                if (g + 16 > s && g < e) {
                    uint4 v;
                    if (g >= 0 && g + 16 <= n_bases) {
                        v = __ldcs(reinterpret_cast<const uint4 *>(bases + g));   // streamed once: evict first
                    } else {
                        uint32_t b[4] = {0, 0, 0, 0};
                        for (int j = 0; j < 16; ++j) {
                            int64_t p = g + j;
                            if (p >= 0 && p < n_bases) b[j >> 2] |= (uint32_t)bases[p] << (8 * (j & 3));
                        }
                        v = make_uint4(b[0], b[1], b[2], b[3]);
                    }
                    fr_pack16(v, c, iv);
                    if (g < s) {            // 1..15 leading positions belong to the previous read
                        const int nb = (int)(s - g);
                        c &= 0xFFFFFFFFu >> (2 * nb);
                        iv &= 0xFFFFu >> nb;
                    }
                }
                sh.codes[w] = c;
                sh.inv[w] = iv;
                if (SAMPLED && !CORE) {      // reverse the 2-bit groups and complement
                    uint32_t rv = __brev(c);
                    rv = ((rv >> 1) & 0x55555555u) | ((rv & 0x55555555u) << 1);
                    sh.rcodes[w] = ~rv;
                }
            }
            __syncthreads();
            // ---- K3a: filter (warps run on their own from here to the end of the tile) ----
            const int64_t cstart = tile_start + (int64_t)FR_PPT * tid;
            const int64_t jl = lo > cstart ? lo - cstart : 0, jh = e - cstart < FR_PPT ? e - cstart : FR_PPT;
            uint32_t pass = 0;
            if (jl < jh) {
                const int wi = 2 * tid;
                const uint32_t c0 = sh.codes[wi], c1 = sh.codes[wi + 1], c2 = sh.codes[wi + 2], c3 = sh.codes[wi + 3];
                const uint64_t iv64 = ((uint64_t)sh.inv[wi] << 48) | ((uint64_t)sh.inv[wi + 1] << 32) |
                                      ((uint64_t)sh.inv[wi + 2] << 16) | (uint64_t)sh.inv[wi + 3];
                const uint32_t pm = (0xFFFFFFFFu >> (int)jl) & (jh == 32 ? 0xFFFFFFFFu : ~(0xFFFFFFFFu >> (int)jh));
                uint32_t am = 0;                       // bit 31-j: the window ending at j holds a non-ACGT base
                if (iv64) {
                    for (int d = 0; d < K; ++d) am |= (uint32_t)(iv64 >> d);
                }
                if (CORE) {
                    pass = fr_filter32_core(c0, c1, c2, c3, tb) & pm & ~am;
                    if (SAMPLED && pass) pass = fr_refine_sampled(c0, c1, c2, c3, pass, K, kmask, tb.seed, tb.threshold);
                } else if (SAMPLED) {
                    // reverse-complement stream of the same 64 positions, shifted right by 2*(33-K) bits
                    const uint32_t r0 = sh.rcodes[wi], r1 = sh.rcodes[wi + 1], r2 = sh.rcodes[wi + 2],
                                   r3 = sh.rcodes[wi + 3];
                    const int bs = 2 * (33 - K), wo = bs >> 5, bb = bs & 31;
                    const uint32_t t0 = wo == 0 ? r0 : (wo == 1 ? r1 : r2), t1 = wo == 0 ? r1 : (wo == 1 ? r2 : r3),
                                   t2 = wo == 0 ? r2 : (wo == 1 ? r3 : 0u), t3 = wo == 0 ? r3 : 0u;
                    const uint32_t q0 = __funnelshift_r(t0, t1, bb), q1 = __funnelshift_r(t1, t2, bb),
                                   q2 = __funnelshift_r(t2, t3, bb), q3 = t3 >> bb;
                    pass = fr_filter32_sampled<WIDE>(c0, c1, c2, c3, q0, q1, q2, q3, kmask_lo, kmask_hi, tb.seed,
                                                     tb.threshold);
                } else if (!(dbg & 2))
                    pass = fr_filter32<WIDE>(c0, c1, c2, c3, kmask_lo, kmask_hi, tb.bloom, wsh, bsh);
                pass = (pass & pm & ~am) | (pm & am & ~(uint32_t)iv64);
                if (dbg & 1) pass = 0;
            }
            // ---- K3b: survivors are rebuilt as k-mers and deferred; the block probes the table in bulk ----
            {
                int total;
                int my = fr_warp_excl_scan(__popc(pass), lane, total);
                if (lane == 31 && total) atomicAdd(&sh.tcount, total);
                __syncthreads();
                const int tcount = sh.tcount, qc0 = sh.qcount;
                __syncthreads();                            // everyone holds the same counts before they move
                if (qc0 + tcount > FR_QCAP) {               // uniform: make the queue empty first
                    fr_block_drain(sh, hl, tb, n_lookups, spill, spill_cap, counters);
                }
                if (tcount <= FR_QCAP) {
                    int base = 0;
                    if (lane == 31 && total) base = atomicAdd(&sh.qcount, total);
                    my += __shfl_sync(0xFFFFFFFFu, base, 31);
                    while (pass) {
                        const int j = __clz(pass);
                        pass &= ~(0x80000000u >> j);
                        uint64_t kmer;
                        // a position whose own base is not ACGT was never selected, so this cannot fail;
                        // the slot is filled either way to keep the queue dense
                        if (!fr_kmer_at(sh, FR_PPT * tid + j, K, kmask, kmer)) kmer = FR_EMPTY_KEY;
                        sh.kq[my++] = kmer;
                    }
                } else {
                    // a tile with more survivors than the queue holds (dense libraries): probe directly
                    fr_make_room(sh, hl, tcount, n_lookups, spill, spill_cap, counters);
                    while (pass) {
                        const int j = __clz(pass);
                        pass &= ~(0x80000000u >> j);
                        uint64_t kmer;
                        if (fr_kmer_at(sh, FR_PPT * tid + j, K, kmask, kmer)) fr_probe_table(tb, kmer, sh, hl);
                    }
                }
            }
            __syncthreads();                                 // sh.codes may be overwritten now
            if (tid == 0) sh.tcount = 0;
        }
        fr_block_drain(sh, hl, tb, n_lookups, spill, spill_cap, counters);
        // ---- K3c: sort + unique, emit ----
        __syncthreads();
        int u = -1;
        const int hc = sh.hcount;
        if (dbg & 4) u = 0;
        else if (hl.ptr == sh.hits && hc > 1 && tb.n > 0) u = fr_bucket_sort_unique(sh, hc, (uint32_t)tb.n);
        if (u < 0) u = fr_block_sort_unique(hl.ptr, hc, sh.scan);
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

// ---- read order: one CTA per read means the longest read of a launch sets its tail.  Reads are
// issued longest first (CTAs start in blockIdx order): a counting sort by length / 1024, done by
// one CTA with shared-memory atomics (the order inside a bucket does not matter; rows still land
// in input order).
#define FR_LEN_BUCKETS 64
__global__ void __launch_bounds__(1024) fr_read_order_kernel(const int64_t *__restrict__ offsets, int64_t n_reads,
                                                            int32_t *__restrict__ order) {
    __shared__ int s_cnt[FR_LEN_BUCKETS];
    const int tid = threadIdx.x;
    if (tid < FR_LEN_BUCKETS) s_cnt[tid] = 0;
    __syncthreads();
    for (int64_t r = tid; r < n_reads; r += blockDim.x) {
        const int64_t b = (offsets[r + 1] - offsets[r]) >> 10;
        atomicAdd(&s_cnt[b < FR_LEN_BUCKETS - 1 ? (int)b : FR_LEN_BUCKETS - 1], 1);
    }
    __syncthreads();
    if (tid == 0) {          // start of each bucket, longest bucket first
        int run = 0;
        for (int b = FR_LEN_BUCKETS - 1; b >= 0; --b) { const int c = s_cnt[b]; s_cnt[b] = run; run += c; }
    }
    __syncthreads();
    for (int64_t r = tid; r < n_reads; r += blockDim.x) {
        const int64_t b = (offsets[r + 1] - offsets[r]) >> 10;
        order[atomicAdd(&s_cnt[b < FR_LEN_BUCKETS - 1 ? (int)b : FR_LEN_BUCKETS - 1], 1)] = (int32_t)r;
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

// fr_extract_async: the same gather without a host round trip.  status[0] collects the reasons the
// rows could not be delivered (1 scratch / 2 spill overflow of the extract pass, 4 indices too
// small), status[1] the number of entries; nothing is written when a bit is set.
__global__ void fr_gather_checked_kernel(const int32_t *__restrict__ scratch, const int64_t *__restrict__ read_off,
                                         int64_t *indptr, int64_t n_reads, int32_t *indices,
                                         int64_t cap, const unsigned long long *__restrict__ counters,
                                         long long *status) {
    const unsigned long long flags = counters[2];
    const long long nnz = (long long)counters[0];
    const long long bad = (long long)(flags & 3ull) | (nnz > cap ? 4ll : 0ll);
    if (blockIdx.x == 0 && threadIdx.x == 0) { status[0] |= bad; status[1] = nnz; }
    if (bad) {      // the rows are not delivered: make them empty so that a consumer queued behind stays in bounds
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n_reads; i += (int64_t)gridDim.x * blockDim.x)
            indptr[i] = 0;
        return;
    }
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
    int64_t grid = n_reads < (int64_t)ctx->sm_count * 128 ? n_reads : (int64_t)ctx->sm_count * 128;
    if (grid < 1) grid = 1;
    typedef void (*KernelFn)(const uint8_t *, int64_t, const int64_t *, int64_t, FrTable, int32_t *, int64_t, uint32_t *,
                             int64_t, int64_t *, int32_t *, unsigned long long *, const int32_t *, int);
    static const KernelFn fns[8] = {fr_extract_kernel<false, false, false>, fr_extract_kernel<true, false, false>,
                                    fr_extract_kernel<false, true, false>, fr_extract_kernel<true, true, false>,
                                    fr_extract_kernel<false, false, true>, fr_extract_kernel<true, false, true>,
                                    fr_extract_kernel<false, true, true>, fr_extract_kernel<true, true, true>};
    static bool attr_set = false;
    if (!attr_set) {
        for (int i = 0; i < 8; ++i)
            FR_CUDA(ctx, cudaFuncSetAttribute(fns[i], cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)sizeof(FrExtractShared)));
        attr_set = true;
    }
    static const int dbg = getenv("FR_EXTRACT_DEBUG") ? atoi(getenv("FR_EXTRACT_DEBUG")) : 0;   // timing experiments only
    // a k-mer of k <= 16 fits one 32-bit word; a verified sampled library skips the filter loads
    // first-stage filter: the core bitmap for a generic library; the register-only sampler test at every
    // position for a verified sample (no L2 traffic at all: measured faster than bitmap + refinement, 18.1 vs
    // 21.6 ms at the chr1 scale).  FR_EXTRACT_FILTER=old / core forces the per-position filters / the bitmap.
    static const char *ff = getenv("FR_EXTRACT_FILTER");
    const bool sampled = ctx->table.sampled && !(dbg & 8);
    const bool core = ff && !strcmp(ff, "old") ? false : (ff && !strcmp(ff, "core") ? true : !sampled);
    const KernelFn fn = fns[(ctx->table.k > 16 ? 1 : 0) + (sampled ? 2 : 0) + (core ? 4 : 0)];
    if (n_reads >= (1ll << 31)) return fr_fail(ctx, FR_ERR_ARG, "fr_extract: more than 2^31 reads in one call");
    if (n_reads > 0) {
        fr_kt_begin(ctx, "fr_read_order_kernel", st);
        fr_read_order_kernel<<<1, 1024, 0, st>>>(d_offsets, n_reads, ctx->d_read_order);
        FR_LAUNCH_CHECK(ctx, "fr_read_order_kernel");
    }
    fr_kt_begin(ctx, "fr_extract_kernel", st);
    fn<<<(unsigned)grid, FR_THREADS, sizeof(FrExtractShared), st>>>(
        d_bases, n_bases, d_offsets, n_reads, ctx->table, ctx->d_scratch, ctx->scratch_cap, ctx->d_spill,
        ctx->spill_cap, ctx->d_read_off, ctx->d_read_cnt, ctx->d_counters, ctx->d_read_order, dbg);
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

int fr_gather_rows_checked(fr_ctx *ctx, int64_t *d_indptr, int64_t n, int32_t *d_indices, int64_t cap,
                           long long *d_status, cudaStream_t st) {
    int64_t blocks = (n * 32 + 255) / 256;
    if (blocks > (int64_t)ctx->sm_count * 32) blocks = (int64_t)ctx->sm_count * 32;
    if (blocks < 1) blocks = 1;
    fr_kt_begin(ctx, "fr_gather_kernel", st);
    fr_gather_checked_kernel<<<(unsigned)blocks, 256, 0, st>>>(ctx->d_scratch, ctx->d_read_off, d_indptr, n, d_indices,
                                                               cap, ctx->d_counters, d_status);
    FR_LAUNCH_CHECK(ctx, "fr_gather_checked_kernel");
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
