// K5 (normalise) and the exact fp32 cosine-kNN scan.
//
// The scan is the arithmetic definition of the path's answer ("the reference's
// kNN backend in exact mode", nearest_neighbors.py:22-55 with metric="cosine";
// SURVEY.md 8(c)): dist = 1 - <x/|x|, y/|y|> in fp32, ascending, ties by lower
// index, zero-norm rows at distance 1 from everything.  It serves fr_knn
// mode 1 and the rows the tensor-core path (knn_tc.cu) cannot certify.
#include "common.cuh"

// ---- K5: one warp per row ----
__global__ void fr_normalize_kernel(const float *__restrict__ E, int64_t N, int D, int Dp,
                                    float *__restrict__ En, __half *__restrict__ Eh,
                                    float *__restrict__ norms) {
    const int lane = threadIdx.x & 31;
    int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; r < N; r += nwarps) {
        const float *x = E + r * D;
        float s = 0.0f;
        for (int j = lane; j < D; j += 32) { float v = x[j]; s = fmaf(v, v, s); }
        for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, d);
        const float nrm = sqrtf(s);
        const float inv = nrm > 0.0f ? 1.0f / nrm : 0.0f;
        if (lane == 0 && norms) norms[r] = nrm;
        for (int j = lane; j < Dp; j += 32) {
            float v = j < D ? x[j] * inv : 0.0f;
            if (j < D && En) En[r * D + j] = v;
            if (Eh) Eh[r * (int64_t)Dp + j] = __float2half_rn(v);
        }
    }
}

extern "C" int fr_normalize(fr_ctx *ctx, const float *d_E, int64_t N, int D, int Dp, float *d_En, void *d_Eh,
                            float *d_norms, void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (N < 0 || D <= 0 || Dp < D || (d_Eh && (Dp % 64) != 0))
        return fr_fail(ctx, FR_ERR_ARG, "fr_normalize: need D > 0, Dp >= D, Dp % 64 == 0");
    if (N == 0) return FR_OK;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t blocks = (N * 32 + 255) / 256;
    if (blocks > (int64_t)ctx->sm_count * 32) blocks = (int64_t)ctx->sm_count * 32;
    fr_kt_begin(ctx, "fr_normalize_kernel", (cudaStream_t)stream);
    fr_normalize_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_E, N, D, Dp, d_En, (__half *)d_Eh,
                                                                           d_norms);
    FR_LAUNCH_CHECK(ctx, "fr_normalize_kernel");
    return FR_OK;
}

// ---- exact scan: one warp = FR_QPW query rows against the whole database ----
#define FR_QPW 4
#define FR_EX_WARPS 4
#define FR_MAX_K 128

// sorted insertion into a warp-private list (dist asc, then idx asc); entries beyond k dropped
__device__ __forceinline__ void fr_list_insert(float *ld, int *li, int k, float d, int idx) {
    const int lane = threadIdx.x & 31;
    // position = number of entries ordered before (d, idx); db is scanned in ascending idx so
    // equal distances already in the list stay in front
    int pos = 0;
    for (int j = lane; j < k; j += 32) pos += (ld[j] <= d) ? 1 : 0;
    for (int s = 16; s > 0; s >>= 1) pos += __shfl_xor_sync(0xFFFFFFFFu, pos, s);
    if (pos >= k) return;
    // shift [pos, k-1) right by one, back to front in 32-wide waves
    for (int hi = k - 1; hi > pos; hi -= 32) {
        int j = hi - lane;
        float td = 0.f; int ti = 0;
        bool mv = j > pos;
        if (mv) { td = ld[j - 1]; ti = li[j - 1]; }
        __syncwarp();
        if (mv) { ld[j] = td; li[j] = ti; }
        __syncwarp();
    }
    if (lane == 0) { ld[pos] = d; li[pos] = idx; }
    __syncwarp();
}

// rows == nullptr: queries are q0 + [0, nq).  Otherwise rows[i] (local query index) for i < *n_rows.
__global__ void __launch_bounds__(FR_EX_WARPS * 32)
fr_knn_exact_kernel(const float *__restrict__ En, const float *__restrict__ norms, int64_t n_db, int D,
                    int64_t q0, int64_t nq, int k, const int32_t *__restrict__ rows,
                    const unsigned long long *__restrict__ n_rows, int64_t row_skip, int32_t *__restrict__ out_idx,
                    float *__restrict__ out_dist) {
    extern __shared__ __align__(16) unsigned char fr_ex_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *qbuf = reinterpret_cast<float *>(fr_ex_smem) + (size_t)warp * FR_QPW * D;
    float *ld = reinterpret_cast<float *>(fr_ex_smem) + (size_t)FR_EX_WARPS * FR_QPW * D +
                (size_t)warp * FR_QPW * FR_MAX_K;
    int *li = reinterpret_cast<int *>(reinterpret_cast<float *>(fr_ex_smem) + (size_t)FR_EX_WARPS * FR_QPW * D +
                                      (size_t)FR_EX_WARPS * FR_QPW * FR_MAX_K) +
              (size_t)warp * FR_QPW * FR_MAX_K;
    // with a row list the first row_skip entries were answered by the parallel fallback below
    const int64_t total = rows ? ((int64_t)(*n_rows) > row_skip ? (int64_t)(*n_rows) - row_skip : 0) : nq;
    if (rows) rows += row_skip;
    const int64_t ngroups = (total + FR_QPW - 1) / FR_QPW;
    for (int64_t g = (int64_t)blockIdx.x * FR_EX_WARPS + warp; g < ngroups; g += (int64_t)gridDim.x * FR_EX_WARPS) {
        int64_t qloc[FR_QPW];
        float qn[FR_QPW], worst[FR_QPW];
#pragma unroll
        for (int u = 0; u < FR_QPW; ++u) {
            int64_t i = g * FR_QPW + u;
            qloc[u] = i < total ? (rows ? (int64_t)rows[i] : i) : -1;
            qn[u] = qloc[u] >= 0 ? norms[q0 + qloc[u]] : 0.0f;
            for (int j = lane; j < D; j += 32)
                qbuf[u * D + j] = qloc[u] >= 0 ? En[(q0 + qloc[u]) * D + j] : 0.0f;
            for (int j = lane; j < k; j += 32) { ld[u * FR_MAX_K + j] = INFINITY; li[u * FR_MAX_K + j] = -1; }
            worst[u] = INFINITY;
        }
        __syncwarp();
        for (int64_t c = 0; c < n_db; ++c) {
            const float *y = En + c * D;
            float acc[FR_QPW];
#pragma unroll
            for (int u = 0; u < FR_QPW; ++u) acc[u] = 0.0f;
            for (int j = lane; j < D; j += 32) {
                const float yv = __ldg(&y[j]);
#pragma unroll
                for (int u = 0; u < FR_QPW; ++u) acc[u] = fmaf(qbuf[u * D + j], yv, acc[u]);
            }
#pragma unroll
            for (int u = 0; u < FR_QPW; ++u)
                for (int s = 16; s > 0; s >>= 1) acc[u] += __shfl_xor_sync(0xFFFFFFFFu, acc[u], s);
            const float cn = norms[c];
#pragma unroll
            for (int u = 0; u < FR_QPW; ++u) {
                if (qloc[u] < 0) continue;
                const float d = (qn[u] == 0.0f || cn == 0.0f) ? 1.0f : 1.0f - acc[u];
                if (d < worst[u]) {  // warp-uniform
                    fr_list_insert(ld + u * FR_MAX_K, li + u * FR_MAX_K, k, d, (int)c);
                    worst[u] = ld[u * FR_MAX_K + k - 1];
                }
            }
        }
#pragma unroll
        for (int u = 0; u < FR_QPW; ++u) {
            if (qloc[u] < 0) continue;
            for (int j = lane; j < k; j += 32) {
                out_idx[qloc[u] * k + j] = li[u * FR_MAX_K + j];
                out_dist[qloc[u] * k + j] = ld[u * FR_MAX_K + j];
            }
        }
        __syncwarp();
    }
}

int fr_knn_exact_launch(fr_ctx *ctx, const float *d_En, const float *d_norms, int64_t n_db, int D, int64_t q0,
                        int64_t nq, int k, const int32_t *d_rows, const unsigned long long *d_n_rows,
                        int64_t row_skip, int32_t *d_idx, float *d_dist, cudaStream_t st) {
    if (k < 1 || k > FR_MAX_K) return fr_fail(ctx, FR_ERR_ARG, "fr_knn: k must be in [1,128]");
    size_t smem = (size_t)FR_EX_WARPS * FR_QPW * D * sizeof(float) +
                  (size_t)FR_EX_WARPS * FR_QPW * FR_MAX_K * (sizeof(float) + sizeof(int));
    if (smem > 200 * 1024) return fr_fail(ctx, FR_ERR_ARG, "fr_knn: D too large for the exact scan");
    FR_CUDA(ctx, cudaFuncSetAttribute(fr_knn_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t groups = (nq + FR_QPW - 1) / FR_QPW;
    int64_t blocks = (groups + FR_EX_WARPS - 1) / FR_EX_WARPS;
    if (blocks > (int64_t)ctx->sm_count * 8) blocks = (int64_t)ctx->sm_count * 8;
    if (blocks < 1) blocks = 1;
    fr_kt_begin(ctx, "fr_knn_exact_kernel", st);
    fr_knn_exact_kernel<<<(unsigned)blocks, FR_EX_WARPS * 32, smem, st>>>(d_En, d_norms, n_db, D, q0, nq, k, d_rows,
                                                                          d_n_rows, row_skip, d_idx, d_dist);
    FR_LAUNCH_CHECK(ctx, "fr_knn_exact_kernel");
    return FR_OK;
}

// ---- parallel fallback for the few rows the tensor-core path could not certify ----
// Same arithmetic as the scan above (lane-strided FMA, xor-shuffle reduction), organised the
// other way round: one warp per DATABASE row computes the exact distance of every fallback
// query (their rows sit in shared memory) and stores orderable 64-bit keys (distance, index);
// then one CTA per query finds its k smallest keys exactly -- a bitwise binary search for the
// k-th smallest key over the whole key array, a gather of the k keys at or below it, a sort.
#define FR_FB_MAX 128       // queries answered per launch pair (fewer when n_db is large: <= 1 GB of keys).
                            // Large databases get several launch pairs (each returns at once when nothing is
                            // left for it): there one warp scanning the whole database per query is far too
                            // slow for a handful of rows.  The rest take the scan above.
#define FR_FB_QB 8          // queries per block of the distance kernel

__device__ __forceinline__ uint32_t fr_f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return u ^ ((uint32_t)((int32_t)u >> 31) | 0x80000000u);
}
__device__ __forceinline__ float fr_ord2f(uint32_t u) {
    return __uint_as_float(u ^ ((u >> 31) ? 0x80000000u : 0xFFFFFFFFu));
}

__global__ void __launch_bounds__(256)
fr_knn_fb_dist_kernel(const float *__restrict__ En, const float *__restrict__ norms, int64_t n_db, int D, int64_t q0,
                      const int32_t *__restrict__ rows, const unsigned long long *__restrict__ n_rows, int f_base,
                      int fb_max, unsigned long long *__restrict__ keys) {
    extern __shared__ __align__(16) unsigned char fr_fb_smem[];
    float *qbuf = reinterpret_cast<float *>(fr_fb_smem);          // [FR_FB_QB][D]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t total = min((int64_t)(*n_rows) - f_base, (int64_t)fb_max);   // rows of this batch
    const int f0 = blockIdx.y * FR_FB_QB;
    if (f0 >= total) return;
    rows += f_base;
    const int nf = (int)min((int64_t)FR_FB_QB, total - f0);
    float qn[FR_FB_QB];
    for (int u = 0; u < nf; ++u) {
        const int64_t q = q0 + rows[f0 + u];
        qn[u] = norms[q];
        for (int j = threadIdx.x; j < D; j += blockDim.x) qbuf[u * D + j] = En[q * D + j];
    }
    __syncthreads();
    for (int64_t c = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp; c < n_db; c += (int64_t)gridDim.x * (blockDim.x >> 5)) {
        const float *y = En + c * D;
        float acc[FR_FB_QB];
#pragma unroll
        for (int u = 0; u < FR_FB_QB; ++u) acc[u] = 0.0f;
        for (int j = lane; j < D; j += 32) {
            const float yv = __ldg(&y[j]);
#pragma unroll
            for (int u = 0; u < FR_FB_QB; ++u)
                if (u < nf) acc[u] = fmaf(qbuf[u * D + j], yv, acc[u]);
        }
#pragma unroll
        for (int u = 0; u < FR_FB_QB; ++u)
            for (int s = 16; s > 0; s >>= 1) acc[u] += __shfl_xor_sync(0xFFFFFFFFu, acc[u], s);
        const float cn = norms[c];
        if (lane == 0) {
#pragma unroll
            for (int u = 0; u < FR_FB_QB; ++u)
                if (u < nf) {
                    const float d = (qn[u] == 0.0f || cn == 0.0f) ? 1.0f : 1.0f - acc[u];
                    keys[(int64_t)(f0 + u) * n_db + c] = ((unsigned long long)fr_f2ord(d) << 32) | (unsigned long long)c;
                }
        }
    }
}

__global__ void __launch_bounds__(1024)
fr_knn_fb_select_kernel(const unsigned long long *__restrict__ keys, int64_t n_db, int k,
                        const int32_t *__restrict__ rows, const unsigned long long *__restrict__ n_rows, int f_base,
                        int fb_max, int32_t *__restrict__ out_idx, float *__restrict__ out_dist) {
    __shared__ int s_part[32];
    __shared__ int s_total;
    __shared__ int s_n;
    __shared__ unsigned long long s_keys[FR_MAX_K * 2];
    const int f = blockIdx.x;
    if (f >= min((int64_t)(*n_rows) - f_base, (int64_t)fb_max)) return;
    rows += f_base;
    const unsigned long long *kp = keys + (int64_t)f * n_db;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // X = k-th smallest key: the largest x with count(keys < x) <= k - 1
    unsigned long long X = 0;
    for (int b = 63; b >= 0; --b) {
        const unsigned long long cand = X | (1ull << b);
        int c = 0;
        for (int64_t i = threadIdx.x; i < n_db; i += blockDim.x) c += __ldcg(kp + i) < cand;
        c = __reduce_add_sync(0xFFFFFFFFu, c);
        if (lane == 0) s_part[warp] = c;
        __syncthreads();
        if (warp == 0) {
            int t = lane < (int)(blockDim.x >> 5) ? s_part[lane] : 0;
            t = __reduce_add_sync(0xFFFFFFFFu, t);
            if (lane == 0) s_total = t;
        }
        __syncthreads();
        if (s_total <= k - 1) X = cand;
        __syncthreads();
    }
    if (threadIdx.x == 0) s_n = 0;
    for (int i = threadIdx.x; i < FR_MAX_K * 2; i += blockDim.x) s_keys[i] = 0xFFFFFFFFFFFFFFFFull;
    __syncthreads();
    for (int64_t i = threadIdx.x; i < n_db; i += blockDim.x) {
        const unsigned long long key = __ldcg(kp + i);
        if (key <= X) { const int p = atomicAdd(&s_n, 1); if (p < FR_MAX_K * 2) s_keys[p] = key; }
    }
    __syncthreads();
    // bitonic sort of 256 keys (keys are distinct: exactly k of them are real)
    for (int kk = 2; kk <= FR_MAX_K * 2; kk <<= 1)
        for (int j = kk >> 1; j > 0; j >>= 1) {
            const int i = threadIdx.x;
            if (i < FR_MAX_K * 2) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const unsigned long long x = s_keys[i], y = s_keys[ixj];
                    const bool up = (i & kk) == 0;
                    if ((x > y) == up) { s_keys[i] = y; s_keys[ixj] = x; }
                }
            }
            __syncthreads();
        }
    const int64_t q = rows[f];
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const unsigned long long key = s_keys[j];
        out_idx[q * k + j] = (int32_t)(uint32_t)key;
        out_dist[q * k + j] = fr_ord2f((uint32_t)(key >> 32));
    }
}

// First level of the selection: the key array of one fallback query is cut into parts of FR_FB_PART keys;
// one CTA per part holds its keys in registers and finds the part's k smallest by the same bitwise
// search (64 block-wide counts, no memory traffic), then writes them (padded with all-ones) for the second
// level, which is fr_knn_fb_select_kernel over parts x FR_MAX_K keys.  A single CTA searching a million
// keys 64 times took 11 ms per uncertified row at the chr1 scale; this is ~0.1 ms.
#define FR_FB_KPT 16
#define FR_FB_PART (1024 * FR_FB_KPT)
__global__ void __launch_bounds__(1024)
fr_knn_fb_part_kernel(const unsigned long long *__restrict__ keys, int64_t n_db, int k,
                      const unsigned long long *__restrict__ n_rows, int f_base, int fb_max, int parts,
                      unsigned long long *__restrict__ part_keys) {
    __shared__ int s_cnt[64];
    __shared__ int s_n;
    const int f = blockIdx.y, part = blockIdx.x;
    if (f >= min((int64_t)(*n_rows) - f_base, (int64_t)fb_max)) return;
    const unsigned long long *kp = keys + (int64_t)f * n_db;
    const int64_t base = (int64_t)part * FR_FB_PART;
    unsigned long long v[FR_FB_KPT];
#pragma unroll
    for (int i = 0; i < FR_FB_KPT; ++i) {
        const int64_t j = base + (int64_t)i * 1024 + threadIdx.x;
        v[i] = j < n_db ? __ldcg(kp + j) : 0xFFFFFFFFFFFFFFFFull;
    }
    if (threadIdx.x < 64) s_cnt[threadIdx.x] = 0;
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    unsigned long long X = 0;       // the part's k-th smallest key: the largest x with count(keys < x) <= k - 1
    for (int b = 63; b >= 0; --b) {
        const unsigned long long cand = X | (1ull << b);
        int c = 0;
#pragma unroll
        for (int i = 0; i < FR_FB_KPT; ++i) c += v[i] < cand;
        c = __reduce_add_sync(0xFFFFFFFFu, c);
        if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_cnt[b], c);
        __syncthreads();
        if (s_cnt[b] <= k - 1) X = cand;
    }
    unsigned long long *out = part_keys + ((int64_t)f * parts + part) * FR_MAX_K;
#pragma unroll
    for (int i = 0; i < FR_FB_KPT; ++i)
        if (v[i] <= X && v[i] != 0xFFFFFFFFFFFFFFFFull) {
            const int q = atomicAdd(&s_n, 1);
            if (q < FR_MAX_K) out[q] = v[i];
        }
    __syncthreads();
    for (int j = min(s_n, FR_MAX_K) + threadIdx.x; j < FR_MAX_K; j += blockDim.x) out[j] = 0xFFFFFFFFFFFFFFFFull;
}

int fr_knn_fallback_launch(fr_ctx *ctx, const float *d_En, const float *d_norms, int64_t n_db, int D, int64_t q0,
                           int k, const int32_t *d_rows, const unsigned long long *d_n_rows, int32_t *d_idx,
                           float *d_dist, int *rows_covered, cudaStream_t st) {
    *rows_covered = 0;
    if (k < 1 || k > FR_MAX_K || n_db < k) return FR_OK;
    // scratch: fb_max x n_db keys, at most 1 GB, then the first-level selections (fb_max x parts x FR_MAX_K keys)
    int fb_max = FR_FB_MAX;
    while (fb_max > 8 && (int64_t)fb_max * n_db * 8 > (1ll << 30)) fb_max >>= 1;
    if ((int64_t)fb_max * n_db * 8 > (1ll << 31)) return FR_OK;
    const int parts = (int)((n_db + FR_FB_PART - 1) / FR_FB_PART);
    const int64_t keys_b = (int64_t)fb_max * n_db * 8;
    const int64_t part_b = (int64_t)fb_max * parts * FR_MAX_K * 8;
    int rc = fr_ensure(ctx, &ctx->d_knn_fb, &ctx->knn_fb_cap, keys_b + part_b, "kNN fallback keys");
    if (rc) return rc;
    unsigned long long *d_keys = (unsigned long long *)ctx->d_knn_fb;
    unsigned long long *d_part = (unsigned long long *)((unsigned char *)ctx->d_knn_fb + keys_b);
    const size_t smem = (size_t)FR_FB_QB * D * sizeof(float);
    FR_CUDA(ctx, cudaFuncSetAttribute(fr_knn_fb_dist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)(ctx->sm_count * 4), (unsigned)((fb_max + FR_FB_QB - 1) / FR_FB_QB));
    int batches = (int)(n_db / 250000);
    batches = batches < 1 ? 1 : (batches > 64 ? 64 : batches);
    for (int b = 0; b < batches; ++b) {
        fr_kt_begin(ctx, "fr_knn_fb_dist_kernel", st);
        fr_knn_fb_dist_kernel<<<grid, 256, smem, st>>>(d_En, d_norms, n_db, D, q0, d_rows, d_n_rows, b * fb_max, fb_max,
                                                       d_keys);
        FR_LAUNCH_CHECK(ctx, "fr_knn_fb_dist_kernel");
        if (parts > 1) {
            fr_kt_begin(ctx, "fr_knn_fb_part_kernel", st);
            fr_knn_fb_part_kernel<<<dim3((unsigned)parts, (unsigned)fb_max), 1024, 0, st>>>(
                d_keys, n_db, k, d_n_rows, b * fb_max, fb_max, parts, d_part);
            FR_LAUNCH_CHECK(ctx, "fr_knn_fb_part_kernel");
        }
        fr_kt_begin(ctx, "fr_knn_fb_select_kernel", st);
        fr_knn_fb_select_kernel<<<fb_max, 1024, 0, st>>>(parts > 1 ? d_part : d_keys,
                                                         parts > 1 ? (int64_t)parts * FR_MAX_K : n_db, k, d_rows,
                                                         d_n_rows, b * fb_max, fb_max, d_idx, d_dist);
        FR_LAUNCH_CHECK(ctx, "fr_knn_fb_select_kernel");
    }
    *rows_covered = batches * fb_max;
    return FR_OK;
}
