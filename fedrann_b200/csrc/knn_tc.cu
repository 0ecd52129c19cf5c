// placeholder replaced by the tcgen05 kernel (see git history): routes to the exact scan
#include "common.cuh"
int fr_knn_exact_launch(fr_ctx *ctx, const float *d_En, const float *d_norms, int64_t n_db, int D, int64_t q0,
                        int64_t nq, int k, const int32_t *d_rows, const unsigned long long *d_n_rows,
                        int32_t *d_idx, float *d_dist, cudaStream_t st);
int fr_knn_tc_launch(fr_ctx *ctx, const float *d_En, const void *d_Eh, const float *d_norms, int64_t n_db, int D,
                     int Dp, int64_t q0, int64_t nq, int k, int32_t *d_idx, float *d_dist, cudaStream_t st) {
    (void)d_Eh; (void)Dp;
    return fr_knn_exact_launch(ctx, d_En, d_norms, n_db, D, q0, nq, k, nullptr, nullptr, d_idx, d_dist, st);
}
