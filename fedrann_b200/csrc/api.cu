// C ABI: context management and the extract (K1+K3) entry points.
#include "common.cuh"

#include <cstdio>
#include <cstring>
#include <new>

int fr_fail(fr_ctx *ctx, int code, const char *what) {
    if (ctx) ctx->err = what;
    return code;
}

int fr_cuda_fail(fr_ctx *ctx, cudaError_t e, const char *what) {
    if (ctx) {
        ctx->err = std::string(what) + ": " + cudaGetErrorString(e);
    }
    return e == cudaErrorMemoryAllocation ? FR_ERR_NOMEM : FR_ERR_CUDA;
}

// grow-only device buffer
int fr_ensure(fr_ctx *ctx, void **ptr, int64_t *cap, int64_t need_bytes, const char *what) {
    if (*ptr && *cap >= need_bytes) return FR_OK;
    if (*ptr) { cudaFree(*ptr); *ptr = nullptr; *cap = 0; }
    if (need_bytes <= 0) need_bytes = 16;
    cudaError_t e = cudaMalloc(ptr, (size_t)need_bytes);
    if (e != cudaSuccess) { *ptr = nullptr; return fr_cuda_fail(ctx, e, what); }
    *cap = need_bytes;
    return FR_OK;
}

// ---- per-kernel timing: one event pair around every launch while enabled ----
void fr_kt_begin(fr_ctx *ctx, const char *name, cudaStream_t st) {
    if (!ctx->kt_on) return;
    fr_ctx::KtRec r;
    r.name = name;
    if (cudaEventCreate(&r.e0) != cudaSuccess || cudaEventCreate(&r.e1) != cudaSuccess) return;
    cudaEventRecord(r.e0, st);
    ctx->kt.push_back(r);
    ctx->kt_stream = st;
    ctx->kt_open = true;
}

void fr_kt_end(fr_ctx *ctx) {
    cudaEventRecord(ctx->kt.back().e1, ctx->kt_stream);
    ctx->kt_open = false;
}

static void fr_kt_clear(fr_ctx *ctx) {
    for (auto &r : ctx->kt) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
    ctx->kt.clear();
    ctx->kt_open = false;
}

extern "C" int fr_profile_enable(fr_ctx *ctx, int on) {
    if (!ctx) return FR_ERR_ARG;
    fr_kt_clear(ctx);
    ctx->kt_on = on != 0;
    return FR_OK;
}

extern "C" int fr_profile_read(fr_ctx *ctx, char *buf, int64_t cap, int64_t *len_out) {
    if (!ctx || !len_out) return FR_ERR_ARG;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    struct Agg { const char *name; int64_t n; double ms; };
    std::vector<Agg> agg;
    for (auto &r : ctx->kt) {
        FR_CUDA(ctx, cudaEventSynchronize(r.e1));
        float ms = 0.f;
        FR_CUDA(ctx, cudaEventElapsedTime(&ms, r.e0, r.e1));
        size_t i = 0;
        for (; i < agg.size(); ++i) if (std::string(agg[i].name) == r.name) break;
        if (i == agg.size()) agg.push_back({r.name, 0, 0.0});
        agg[i].n++;
        agg[i].ms += ms;
    }
    std::string out;
    char line[256];
    for (auto &a : agg) {
        snprintf(line, sizeof line, "%s %lld %.6f\n", a.name, (long long)a.n, a.ms);
        out += line;
    }
    *len_out = (int64_t)out.size();
    if (buf && cap > 0) {
        size_t n = out.size() < (size_t)cap - 1 ? out.size() : (size_t)cap - 1;
        memcpy(buf, out.data(), n);
        buf[n] = 0;
    }
    fr_kt_clear(ctx);
    return FR_OK;
}

extern "C" int fr_version(void) { return 100; }

extern "C" int fr_create(int device, fr_ctx **out) {
    if (!out) return FR_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev)
        return FR_ERR_NO_DEVICE;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return FR_ERR_NO_DEVICE;
    if (prop.major != 10) return FR_ERR_NO_DEVICE;  // sm_100a code only
    if (cudaSetDevice(device) != cudaSuccess) return FR_ERR_CUDA;
    fr_ctx *ctx = new (std::nothrow) fr_ctx();
    if (!ctx) return FR_ERR_NOMEM;
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    if (cudaMalloc(&ctx->d_counters, 8 * sizeof(unsigned long long)) != cudaSuccess ||
        cudaMallocHost(&ctx->h_counters, 8 * sizeof(unsigned long long)) != cudaSuccess) {
        delete ctx;
        return FR_ERR_NOMEM;
    }
    *out = ctx;
    return FR_OK;
}

extern "C" void fr_destroy(fr_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaFree(ctx->d_bloom); cudaFree(ctx->d_cores); cudaFree(ctx->d_slots); cudaFree(ctx->d_cslots);
    cudaFree(ctx->d_scratch); cudaFree(ctx->d_spill); cudaFree(ctx->d_read_off);
    cudaFree(ctx->d_read_cnt); cudaFree(ctx->d_read_order); cudaFree(ctx->d_counters); cudaFree(ctx->d_scan_tmp);
    cudaFree(ctx->d_knn_ws); cudaFree(ctx->d_knn_fb); cudaFree(ctx->d_count_slots);
    if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
    fr_kt_clear(ctx);
    if (ctx->knn_inc && ctx->knn_inc_free) ctx->knn_inc_free(ctx->knn_inc);
    delete ctx;
}

extern "C" const char *fr_last_error(const fr_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }
extern "C" int64_t fr_launch_count(const fr_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" int fr_extract_count(fr_ctx *ctx, const uint8_t *d_bases, int64_t n_bases, const int64_t *d_offsets,
                                int64_t n_reads, int64_t *d_indptr, uint32_t *d_docfreq, int64_t *nnz_out,
                                void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (!ctx->has_table) return fr_fail(ctx, FR_ERR_STATE, "fr_extract_count: no table (call fr_table_build)");
    if (n_reads < 0 || n_bases < 0 || !d_offsets || !d_indptr || !nnz_out)
        return fr_fail(ctx, FR_ERR_ARG, "fr_extract_count: bad argument");
    if (((uintptr_t)d_bases & 15) != 0)
        return fr_fail(ctx, FR_ERR_ARG, "fr_extract_count: d_bases must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    ctx->last_n_reads = -1;
    // per-read bookkeeping
    if (ctx->reads_cap < n_reads || !ctx->d_read_off) {
        if (ctx->d_read_off) cudaFree(ctx->d_read_off);
        if (ctx->d_read_cnt) cudaFree(ctx->d_read_cnt);
        if (ctx->d_read_order) cudaFree(ctx->d_read_order);
        ctx->d_read_off = nullptr; ctx->d_read_cnt = nullptr; ctx->d_read_order = nullptr;
        int64_t cap = n_reads > 0 ? n_reads : 1;
        FR_CUDA(ctx, cudaMalloc(&ctx->d_read_off, (size_t)cap * 8));
        FR_CUDA(ctx, cudaMalloc(&ctx->d_read_cnt, (size_t)cap * 4));
        FR_CUDA(ctx, cudaMalloc(&ctx->d_read_order, (size_t)cap * 4));
        ctx->reads_cap = cap;
    }
    // first guess: one hit per 8 bases (sample fractions <= 0.05 give <= 1/20)
    int64_t want = n_bases / 8 + n_reads + 1024;
    if (ctx->scratch_cap < want) {
        int64_t capb = ctx->scratch_cap * 4;
        int rc = fr_ensure(ctx, (void **)&ctx->d_scratch, &capb, want * 4, "scratch");
        if (rc) return rc;
        ctx->scratch_cap = capb / 4;
    }
    for (int attempt = 0; attempt < 3; ++attempt) {
        FR_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, 8 * sizeof(unsigned long long), st));
        int rc = fr_extract_launch(ctx, d_bases, n_bases, d_offsets, n_reads, d_docfreq, st);
        if (rc) return rc;
        FR_CUDA(ctx, cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 8 * sizeof(unsigned long long),
                                     cudaMemcpyDeviceToHost, st));
        FR_CUDA(ctx, cudaStreamSynchronize(st));
        const unsigned long long used = ctx->h_counters[0], spill = ctx->h_counters[1],
                                 flags = ctx->h_counters[2];
        if (flags == 0) {
            rc = fr_scan_counts(ctx, ctx->d_read_cnt, n_reads, d_indptr, st);
            if (rc) return rc;
            if (d_docfreq) {
                rc = fr_docfreq_launch(ctx, (int64_t)used, d_docfreq, st);
                if (rc) return rc;
            }
            ctx->last_n_reads = n_reads;
            ctx->last_indptr = d_indptr;
            *nnz_out = (int64_t)used;
            return FR_OK;
        }
        // grow what overflowed and run again (results so far are discarded)
        if (flags & 2ull) {
            int64_t capb = ctx->spill_cap * 4;
            rc = fr_ensure(ctx, (void **)&ctx->d_spill, &capb, (int64_t)spill * 4 + 1024, "spill");
            if (rc) return rc;
            ctx->spill_cap = capb / 4;
        }
        if ((flags & 1ull) || (flags & 2ull)) {
            // with dead rows the count is a lower bound: size for the worst case on the next try
            int64_t need = (flags & 2ull) ? (int64_t)used + (int64_t)spill : (int64_t)used;
            if (need > ctx->scratch_cap) {
                int64_t capb = ctx->scratch_cap * 4;
                rc = fr_ensure(ctx, (void **)&ctx->d_scratch, &capb, need * 4 + 1024, "scratch");
                if (rc) return rc;
                ctx->scratch_cap = capb / 4;
            }
        }
    }
    return fr_fail(ctx, FR_ERR_STATE, "fr_extract_count: scratch sizing did not converge");
}

extern "C" int fr_extract_async(fr_ctx *ctx, const uint8_t *d_bases, int64_t n_bases, const int64_t *d_offsets,
                                int64_t n_reads, int64_t *d_indptr, int32_t *d_indices, int64_t cap,
                                int64_t *d_status, void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (!ctx->has_table) return fr_fail(ctx, FR_ERR_STATE, "fr_extract_async: no table (call fr_table_build)");
    if (n_reads < 0 || n_bases < 0 || cap < 0 || !d_offsets || !d_indptr || !d_indices || !d_status)
        return fr_fail(ctx, FR_ERR_ARG, "fr_extract_async: bad argument");
    if (((uintptr_t)d_bases & 15) != 0)
        return fr_fail(ctx, FR_ERR_ARG, "fr_extract_async: d_bases must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    ctx->last_n_reads = -1;
    if (ctx->reads_cap < n_reads || !ctx->d_read_off) {       // (allocation = a device-wide sync, first call only)
        if (ctx->d_read_off) cudaFree(ctx->d_read_off);
        if (ctx->d_read_cnt) cudaFree(ctx->d_read_cnt);
        if (ctx->d_read_order) cudaFree(ctx->d_read_order);
        ctx->d_read_off = nullptr; ctx->d_read_cnt = nullptr; ctx->d_read_order = nullptr;
        int64_t rcap = n_reads > 0 ? n_reads : 1;
        FR_CUDA(ctx, cudaMalloc(&ctx->d_read_off, (size_t)rcap * 8));
        FR_CUDA(ctx, cudaMalloc(&ctx->d_read_cnt, (size_t)rcap * 4));
        FR_CUDA(ctx, cudaMalloc(&ctx->d_read_order, (size_t)rcap * 4));
        ctx->reads_cap = rcap;
    }
    // scratch for as many hits as the caller has room for (at least the first guess of fr_extract_count)
    int64_t want = n_bases / 8 + n_reads + 1024;
    if (want < cap + n_reads) want = cap + n_reads;
    if (ctx->scratch_cap < want) {
        int64_t capb = ctx->scratch_cap * 4;
        int rc = fr_ensure(ctx, (void **)&ctx->d_scratch, &capb, want * 4, "scratch");
        if (rc) return rc;
        ctx->scratch_cap = capb / 4;
    }
    FR_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, 8 * sizeof(unsigned long long), st));
    int rc = fr_extract_launch(ctx, d_bases, n_bases, d_offsets, n_reads, nullptr, st);
    if (rc) return rc;
    rc = fr_scan_counts(ctx, ctx->d_read_cnt, n_reads, d_indptr, st);
    if (rc) return rc;
    return fr_gather_rows_checked(ctx, d_indptr, n_reads, d_indices, cap, (long long *)d_status, st);
}

extern "C" int fr_extract_fill(fr_ctx *ctx, int32_t *d_indices, void *stream) {
    if (!ctx) return FR_ERR_ARG;
    if (ctx->last_n_reads < 0) return fr_fail(ctx, FR_ERR_STATE, "fr_extract_fill: no preceding fr_extract_count");
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    return fr_gather_rows(ctx, d_indices, (cudaStream_t)stream);
}
