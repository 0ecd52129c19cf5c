// fr_knn_exact_noalloc: the exact fp32 cosine-kNN scan for callers that must not allocate.
//
// No counterpart in the reference; added for code under CUDA graph
// capture, tight serving loops and harnesses that forbid cudaMalloc.  Same answer as fr_knn mode 1
// (knn_exact.cu: dist = 1 - <x/|x|, y/|y|> in fp32 with the same lane-strided FMA order and xor-shuffle
// reduction, ascending, ties by lower index, zero-norm rows at distance 1), organised like the parallel
// fallback there: a distance kernel stores one orderable 64-bit key (distance, index) per (query, database
// row), then one CTA per query picks its k smallest keys by a bitwise search and sorts them.
//
// The key matrix is the scratch.  It lives in a file-scope __device__ array of FR_NOALLOC_MAX_KEYS keys
// (16 MiB) that every GPU which loads the library holds until exit; the entry point calls no cudaMalloc /
// cudaFree / allocating library routine.  Every call uses the same array:
//   * nothing in it has to start at zero -- the selection of a call reads exactly the nq x n_db keys the
//     distance kernel of the same call wrote before it on the same stream, so stale contents are never seen;
//   * two calls that overlap on one GPU (different streams, or different threads) RACE on it.  Serialise them.
#include "common.cuh"

#include <mutex>

#define FR_NOALLOC_MAX_KEYS (1ll << 21)   // nq * n_db <= 2 097 152 keys = 16 MiB
#define FR_NOALLOC_MAX_NQ 4096
#define FR_NOALLOC_MAX_D 1024             // 8 query rows in 32 KB of shared memory (no opt-in needed)
#define FR_NOALLOC_MAX_K 128
#define FR_NOALLOC_QB 8                   // queries per block of the distance kernel
#define FR_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned long long fr_noalloc_keys[FR_NOALLOC_MAX_KEYS];

namespace {

// In host code `fr_noalloc_keys` is only a stand-in for the symbol: the device address is asked for once per
// device, with that device current, and handed to the kernels as an argument.
std::mutex g_mu;
unsigned long long *g_keys_ptr[FR_NOALLOC_MAX_DEVICES] = {nullptr};

__device__ __forceinline__ uint32_t na_f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return u ^ ((uint32_t)((int32_t)u >> 31) | 0x80000000u);
}
__device__ __forceinline__ float na_ord2f(uint32_t u) {
    return __uint_as_float(u ^ ((u >> 31) ? 0x80000000u : 0xFFFFFFFFu));
}

// one warp per database row, FR_NOALLOC_QB query rows of the block in shared memory
__global__ void __launch_bounds__(256)
fr_knn_noalloc_dist_kernel(const float *__restrict__ En, const float *__restrict__ norms, int64_t n_db, int D,
                           int64_t q0, int64_t nq, unsigned long long *__restrict__ keys) {
    extern __shared__ __align__(16) unsigned char fr_na_smem[];
    float *qbuf = reinterpret_cast<float *>(fr_na_smem);          // [FR_NOALLOC_QB][D]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t f0 = (int64_t)blockIdx.y * FR_NOALLOC_QB;
    if (f0 >= nq) return;
    const int nf = (int)min((int64_t)FR_NOALLOC_QB, nq - f0);
    float qn[FR_NOALLOC_QB];
#pragma unroll
    for (int u = 0; u < FR_NOALLOC_QB; ++u) qn[u] = 0.0f;
    for (int u = 0; u < nf; ++u) {
        const int64_t q = q0 + f0 + u;
        qn[u] = norms[q];
        for (int j = threadIdx.x; j < D; j += blockDim.x) qbuf[u * D + j] = En[q * D + j];
    }
    __syncthreads();
    const int wpb = blockDim.x >> 5;
    for (int64_t c = (int64_t)blockIdx.x * wpb + warp; c < n_db; c += (int64_t)gridDim.x * wpb) {
        const float *y = En + c * D;
        float acc[FR_NOALLOC_QB];
#pragma unroll
        for (int u = 0; u < FR_NOALLOC_QB; ++u) acc[u] = 0.0f;
        for (int j = lane; j < D; j += 32) {
            const float yv = __ldg(&y[j]);
#pragma unroll
            for (int u = 0; u < FR_NOALLOC_QB; ++u)
                if (u < nf) acc[u] = fmaf(qbuf[u * D + j], yv, acc[u]);
        }
#pragma unroll
        for (int u = 0; u < FR_NOALLOC_QB; ++u)
            for (int s = 16; s > 0; s >>= 1) acc[u] += __shfl_xor_sync(0xFFFFFFFFu, acc[u], s);
        const float cn = norms[c];
        if (lane == 0) {
#pragma unroll
            for (int u = 0; u < FR_NOALLOC_QB; ++u)
                if (u < nf) {
                    const float d = (qn[u] == 0.0f || cn == 0.0f) ? 1.0f : 1.0f - acc[u];
                    keys[(f0 + u) * n_db + c] = ((unsigned long long)na_f2ord(d) << 32) | (unsigned long long)c;
                }
        }
    }
}

// one CTA per query: X = its k-th smallest key (the largest x with count(keys < x) <= k - 1), then the keys
// at or below X, sorted.  Keys are distinct (the index is part of the key).
__global__ void __launch_bounds__(1024)
fr_knn_noalloc_select_kernel(const unsigned long long *__restrict__ keys, int64_t n_db, int k,
                             int32_t *__restrict__ out_idx, float *__restrict__ out_dist) {
    __shared__ int s_part[32];
    __shared__ int s_total;
    __shared__ int s_n;
    __shared__ unsigned long long s_keys[FR_NOALLOC_MAX_K * 2];
    const int64_t f = blockIdx.x;
    const unsigned long long *kp = keys + f * n_db;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
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
    for (int i = threadIdx.x; i < FR_NOALLOC_MAX_K * 2; i += blockDim.x) s_keys[i] = 0xFFFFFFFFFFFFFFFFull;
    __syncthreads();
    for (int64_t i = threadIdx.x; i < n_db; i += blockDim.x) {
        const unsigned long long key = __ldcg(kp + i);
        if (key <= X) { const int p = atomicAdd(&s_n, 1); if (p < FR_NOALLOC_MAX_K * 2) s_keys[p] = key; }
    }
    __syncthreads();
    for (int kk = 2; kk <= FR_NOALLOC_MAX_K * 2; kk <<= 1)
        for (int j = kk >> 1; j > 0; j >>= 1) {
            const int i = threadIdx.x;
            if (i < FR_NOALLOC_MAX_K * 2) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const unsigned long long x = s_keys[i], y = s_keys[ixj];
                    const bool up = (i & kk) == 0;
                    if ((x > y) == up) { s_keys[i] = y; s_keys[ixj] = x; }
                }
            }
            __syncthreads();
        }
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const unsigned long long key = s_keys[j];
        const bool none = key == 0xFFFFFFFFFFFFFFFFull;       // fewer than k database rows: as the scan reports it
        out_idx[f * k + j] = none ? -1 : (int32_t)(uint32_t)key;
        out_dist[f * k + j] = none ? INFINITY : na_ord2f((uint32_t)(key >> 32));
    }
}

}  // namespace

extern "C" int fr_knn_exact_noalloc(fr_ctx *ctx, const float *d_En, const float *d_norms, int64_t n_db, int D,
                                    int64_t q0, int64_t nq, int k, int32_t *d_idx, float *d_dist, void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (!d_En || !d_norms || !d_idx || !d_dist) return fr_fail(ctx, FR_ERR_ARG, "fr_knn_exact_noalloc: null pointer");
    if (n_db < 1 || D < 1 || q0 < 0 || nq < 0 || q0 + nq > n_db)
        return fr_fail(ctx, FR_ERR_ARG, "fr_knn_exact_noalloc: need n_db >= 1, D >= 1, 0 <= q0, q0 + nq <= n_db");
    if (k < 1 || k > FR_NOALLOC_MAX_K)
        return fr_fail(ctx, FR_ERR_ARG, "fr_knn_exact_noalloc: k must be in [1,128]");
    if (D > FR_NOALLOC_MAX_D) return fr_fail(ctx, FR_ERR_ARG, "fr_knn_exact_noalloc: D exceeds the maximum of 1024");
    if (nq > FR_NOALLOC_MAX_NQ)
        return fr_fail(ctx, FR_ERR_ARG, "fr_knn_exact_noalloc: nq exceeds the maximum of 4096 queries per call");
    if (nq * n_db > FR_NOALLOC_MAX_KEYS)
        return fr_fail(ctx, FR_ERR_ARG,
                       "fr_knn_exact_noalloc: nq * n_db exceeds the static scratch of 2097152 keys (16 MiB); "
                       "use fr_knn, or fewer queries per call");
    if (nq == 0) return FR_OK;
    if (ctx->device < 0 || ctx->device >= FR_NOALLOC_MAX_DEVICES)
        return fr_fail(ctx, FR_ERR_ARG, "fr_knn_exact_noalloc: device ordinal out of range");
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned long long *keys = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_mu);
        if (!g_keys_ptr[ctx->device]) {
            void *p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, fr_noalloc_keys);
            if (e != cudaSuccess || !p) return fr_cuda_fail(ctx, e, "fr_knn_exact_noalloc: cudaGetSymbolAddress");
            g_keys_ptr[ctx->device] = static_cast<unsigned long long *>(p);
        }
        keys = g_keys_ptr[ctx->device];
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = (size_t)FR_NOALLOC_QB * D * sizeof(float);
    int64_t gx = (n_db + 7) / 8;                                  // 8 warps = 8 database rows per block and turn
    if (gx > (int64_t)ctx->sm_count * 4) gx = (int64_t)ctx->sm_count * 4;
    dim3 grid((unsigned)gx, (unsigned)((nq + FR_NOALLOC_QB - 1) / FR_NOALLOC_QB));
    fr_knn_noalloc_dist_kernel<<<grid, 256, smem, st>>>(d_En, d_norms, n_db, D, q0, nq, keys);
    FR_LAUNCH_CHECK(ctx, "fr_knn_noalloc_dist_kernel");
    fr_knn_noalloc_select_kernel<<<(unsigned)nq, 1024, 0, st>>>(keys, n_db, k, d_idx, d_dist);
    FR_LAUNCH_CHECK(ctx, "fr_knn_noalloc_select_kernel");
    return FR_OK;
}
