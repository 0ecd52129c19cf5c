// fr_knn dispatcher: tensor-core candidate pass + exact re-rank (mode 0) or the
// exact fp32 scan (mode 1).  Replaces NNDescent_ava.get_neighbors
// (nearest_neighbors.py:22-55) with an exact search.
#include "common.cuh"

int fr_knn_exact_launch(fr_ctx *ctx, const float *d_En, const float *d_norms, int64_t n_db, int D, int64_t q0,
                        int64_t nq, int k, const int32_t *d_rows, const unsigned long long *d_n_rows,
                        int64_t row_skip, int32_t *d_idx, float *d_dist, cudaStream_t st);
int fr_knn_wait_rows(fr_ctx *ctx, cudaStream_t st);
int fr_knn_tc_launch(fr_ctx *ctx, const float *d_En, const void *d_Eh, const float *d_norms, int64_t n_db, int D,
                     int Dp, int64_t q0, int64_t nq, int k, int32_t *d_idx, float *d_dist, cudaStream_t st);

extern "C" int fr_knn(fr_ctx *ctx, const float *d_En, const void *d_Eh, const float *d_norms, int64_t n_db, int D,
                      int Dp, int64_t q0, int64_t nq, int k, int mode, int32_t *d_idx, float *d_dist,
                      void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (!d_En || !d_norms || !d_idx || !d_dist || n_db < 0 || nq < 0 || q0 < 0 || q0 + nq > n_db || D <= 0)
        return fr_fail(ctx, FR_ERR_ARG, "fr_knn: bad argument");
    if (nq == 0) return FR_OK;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    ctx->knn_fallback_rows = 0;
    ctx->knn_fb_pending = false;
    cudaStream_t st = (cudaStream_t)stream;
    if (mode == 1) {
        int rc = fr_knn_wait_rows(ctx, st);
        if (rc) return rc;
    }
    ctx->knn_last_path = 1;
    if (mode == 1) ctx->knn_shard_flags = nullptr;      // (the scan waited for every row above)
    if (mode == 1)
        return fr_knn_exact_launch(ctx, d_En, d_norms, n_db, D, q0, nq, k, nullptr, nullptr, 0, d_idx, d_dist, st);
    if (mode != 0) return fr_fail(ctx, FR_ERR_ARG, "fr_knn: mode must be 0 or 1");
    if (!d_Eh || Dp < D || (Dp % 64) != 0)
        return fr_fail(ctx, FR_ERR_ARG, "fr_knn: mode 0 needs the fp16 copy with Dp % 64 == 0");
    return fr_knn_tc_launch(ctx, d_En, d_Eh, d_norms, n_db, D, Dp, q0, nq, k, d_idx, d_dist, st);
}

extern "C" int fr_knn_rows_ready_after(fr_ctx *ctx, void *event) {
    if (!ctx) return FR_ERR_ARG;
    ctx->knn_en_event = (cudaEvent_t)event;
    return FR_OK;
}

// the next consumer of d_En on `st` waits for the event announced by fr_knn_rows_ready_after
int fr_knn_wait_rows(fr_ctx *ctx, cudaStream_t st) {
    if (ctx->knn_en_event) {
        FR_CUDA(ctx, cudaStreamWaitEvent(st, ctx->knn_en_event, 0));
        ctx->knn_en_event = nullptr;
    }
    return FR_OK;
}

extern "C" int fr_knn_last_path(const fr_ctx *ctx) { return ctx ? ctx->knn_last_path : -1; }

extern "C" int64_t fr_knn_last_fallback_rows(fr_ctx *ctx) {
    if (!ctx) return 0;
    if (ctx->knn_fb_pending) {
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->knn_fb_stream);
        ctx->knn_fallback_rows = (int64_t)ctx->h_counters[4];
        ctx->knn_fb_pending = false;
    }
    return ctx->knn_fallback_rows;
}
