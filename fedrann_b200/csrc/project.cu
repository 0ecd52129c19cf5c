// K4 -- E = A.P, the projection of feature_extraction.py:204-213 with the row
// pairing of :136-140 (row 2r = forward hits, row 2r+1 = index-swapped hits).
//
// scipy's csr_matmat walks A's columns in ascending order and adds P's
// entries one at a time into a fp32 accumulator, so for every output column j
// the additions happen in ascending feature order.  Additions to DIFFERENT
// columns commute, so a warp may apply 32 staged entries at once as long as
// entries that share a column are applied in their original order; that is
// what the match_any ranking below does.  The result is bit-identical to the
// reference, not merely close.
//
// One CTA (2 warps) per read: warp 0 accumulates the forward row, warp 1 the
// reverse row, each in its own D-float shared-memory accumulator.
#include "common.cuh"

#define FR_PJ_STAGE 512
#define FR_PJ_U 4

__device__ __forceinline__ void fr_apply_group(float *acc, int j, float v, bool valid) {
    // lanes with equal j must add in lane order
    unsigned peers = __match_any_sync(0xFFFFFFFFu, valid ? j : -1 - (int)(threadIdx.x & 31));
    const unsigned lt = (1u << (threadIdx.x & 31)) - 1;
    int rank = __popc(peers & lt);
    int maxr = rank;
    for (int d = 16; d > 0; d >>= 1) maxr = max(maxr, __shfl_xor_sync(0xFFFFFFFFu, maxr, d));
    for (int round = 0; round <= maxr; ++round) {
        if (valid && rank == round) acc[j] = __fadd_rn(acc[j], v);
        __syncwarp();
    }
}

__global__ void __launch_bounds__(64)
fr_project_kernel(const int64_t *__restrict__ indptr, const int32_t *__restrict__ indices, int64_t n_reads,
                  const int64_t *__restrict__ p_indptr, const int32_t *__restrict__ p_indices,
                  const float *__restrict__ p_data, int64_t nk, int D, float *__restrict__ E) {
    extern __shared__ __align__(16) unsigned char fr_pj_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *acc = reinterpret_cast<float *>(fr_pj_smem) + (size_t)warp * D;
    int *st_col = reinterpret_cast<int *>(fr_pj_smem + (size_t)2 * D * sizeof(float)) + warp * FR_PJ_STAGE;
    float *st_val = reinterpret_cast<float *>(fr_pj_smem + (size_t)2 * D * sizeof(float) +
                                              2 * FR_PJ_STAGE * sizeof(int)) + warp * FR_PJ_STAGE;
    for (int64_t r = blockIdx.x; r < n_reads; r += gridDim.x) {
        for (int j = lane; j < D; j += 32) acc[j] = 0.0f;
        __syncwarp();
        const int64_t a = indptr[r];
        const int n = (int)(indptr[r + 1] - a);
        const int32_t *row = indices + a;
        // split = number of hits < nk (the row is ascending): counted with coalesced loads (one
        // round trip instead of a dependent binary search; the row is then in L1 for the passes below)
        int split = 0;
        for (int t = lane; t < n; t += 32) split += (int64_t)row[t] < nk;
        split = __reduce_add_sync(0xFFFFFFFFu, split);
        const int nhi = n - split;
        // FR_PJ_U groups of 32 hits per pass: their P.indptr lookups and P entries are independent
        // loads, so the three dependent round trips (hit -> indptr -> entries) are paid once per 128 hits
        for (int base = 0; base < n; base += 32 * FR_PJ_U) {
            int64_t pa[FR_PJ_U];
            int cnt[FR_PJ_U], excl[FR_PJ_U];
#pragma unroll
            for (int u = 0; u < FR_PJ_U; ++u) {
                const int t = base + 32 * u + lane;
                int64_t feat = -1;
                if (t < n) {
                    if (warp == 0) feat = row[t];
                    else  // swapped set in ascending order: [hits >= nk] - nk, then [hits < nk] + nk
                        feat = t < nhi ? (int64_t)row[split + t] - nk : (int64_t)row[t - nhi] + nk;
                }
                pa[u] = 0;
                cnt[u] = 0;
                if (feat >= 0) {
                    pa[u] = __ldg(&p_indptr[feat]);
                    cnt[u] = (int)(__ldg(&p_indptr[feat + 1]) - pa[u]);
                }
            }
            int total = 0;
#pragma unroll
            for (int u = 0; u < FR_PJ_U; ++u) {     // staging order = hit order: group u, then lane
                int incl = cnt[u];
                for (int d = 1; d < 32; d <<= 1) {
                    int y = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                    if (lane >= d) incl += y;
                }
                excl[u] = total + incl - cnt[u];
                total += __shfl_sync(0xFFFFFFFFu, incl, 31);
            }
            if (total == 0) continue;
            if (total <= FR_PJ_STAGE) {
#pragma unroll
                for (int u = 0; u < FR_PJ_U; ++u)
                    for (int m = 0; m < cnt[u]; ++m) {
                        st_col[excl[u] + m] = __ldg(&p_indices[pa[u] + m]);
                        st_val[excl[u] + m] = __ldg(&p_data[pa[u] + m]);
                    }
                __syncwarp();
                for (int g = 0; g < total; g += 32) {
                    const int e = g + lane;
                    const bool ev = e < total;
                    int j = ev ? st_col[e] : 0;
                    float v = ev ? st_val[e] : 0.0f;
                    fr_apply_group(acc, j, v, ev);
                }
                __syncwarp();
            } else {
                // pathological density: strictly sequential, group by group, lane by lane
#pragma unroll
                for (int u = 0; u < FR_PJ_U; ++u)
                    for (int l = 0; l < 32; ++l) {
                        if (lane == l)
                            for (int m = 0; m < cnt[u]; ++m) {
                                int j = __ldg(&p_indices[pa[u] + m]);
                                acc[j] = __fadd_rn(acc[j], __ldg(&p_data[pa[u] + m]));
                            }
                        __syncwarp();
                    }
            }
        }
        __syncwarp();
        float *out = E + ((size_t)2 * r + warp) * D;
        for (int j = lane; j < D; j += 32) out[j] = acc[j];
        __syncwarp();
    }
}

extern "C" int fr_project(fr_ctx *ctx, const int64_t *d_indptr, const int32_t *d_indices, int64_t n_reads,
                          const int64_t *d_p_indptr, const int32_t *d_p_indices, const float *d_p_data,
                          int64_t nk, int D, float *d_E, void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (n_reads < 0 || D <= 0 || D > 8192 || nk < 0)
        return fr_fail(ctx, FR_ERR_ARG, "fr_project: bad argument (D must be in [1,8192])");
    if (n_reads == 0) return FR_OK;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t smem = (size_t)2 * D * sizeof(float) + 2 * FR_PJ_STAGE * (sizeof(int) + sizeof(float));
    FR_CUDA(ctx, cudaFuncSetAttribute(fr_project_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t grid = n_reads < (int64_t)ctx->sm_count * 32 ? n_reads : (int64_t)ctx->sm_count * 32;
    fr_kt_begin(ctx, "fr_project_kernel", (cudaStream_t)stream);
    fr_project_kernel<<<(unsigned)grid, 64, smem, (cudaStream_t)stream>>>(
        d_indptr, d_indices, n_reads, d_p_indptr, d_p_indices, d_p_data, nk, D, d_E);
    FR_LAUNCH_CHECK(ctx, "fr_project_kernel");
    return FR_OK;
}
