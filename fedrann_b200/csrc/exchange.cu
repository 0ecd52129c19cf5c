// Multi-GPU exchange of the embedding database through PEER MEMORY (NVLink / NVSwitch), one rank per GPU.
//
// reads shard over the GPUs, every GPU extracts / projects / normalises its rows, then
// every GPU searches ITS query rows against ALL rows.  The reference has no counterpart (it is one
// process: fedrann/feature_extraction.py:235-290 is its only parallel site), so this replaces the "one
// all-gather of embeddings" of the build plan -- and it is not a collective call: every rank owns one
// device slab (cudaMalloc, exported with CUDA IPC) that holds the WHOLE database
//     En f32 [N, D] | Eh f16 [N, Dp] | norms f32 [N] | shard flags u32 | barrier words u32
// and normalises its own rows straight into its row range of that slab.  fr_exchange_pull then
//   1. crosses a flag barrier (a 1-CTA kernel: release-store of the epoch into every peer's barrier word,
//      acquire-spin on the local words): every rank's rows are written;
//   2. queues, on a copy stream, one peer-to-peer copy per remote shard in RING order (rank+1, rank+2, ...)
//      -- copy engines over NVLink, no SM is taken from the search -- first the fp16 rows the tensor-core
//      pass needs, each followed by a stream memory operation that publishes "shard s has landed" in the
//      local flag word; then the norms and the fp32 rows the re-rank needs, and an event;
//   3. hands the flags to the NEXT fr_knn call of the context: K6 starts on the local shard at once, its
//      TMA producer spins on a shard's flag only when it reaches the first database tile of that shard (the
//      tile order starts at the local shard and follows the ring), and K7 waits for the event.  The
//      transfer is hidden behind the search tile by tile instead of preceding it.
// fr_exchange_release (after the search) is the second barrier: nobody overwrites rows a peer may still
// be pulling.
#include "common.cuh"

#include <cuda.h>
#include <cstdio>
#include <cstring>

struct fr_exchange {
    fr_ctx *ctx = nullptr;
    int rank = 0, world = 1;
    int D = 0, Dp = 0;
    int64_t bounds[FR_MAX_RANKS + 1] = {0};
    int64_t N = 0;
    unsigned char *slab = nullptr;                 // local
    size_t slab_bytes = 0;
    size_t off_En = 0, off_Eh = 0, off_norms = 0, off_flags = 0, off_sync = 0;
    unsigned char *peer[FR_MAX_RANKS] = {nullptr};  // peer[r] = rank r's slab in this process (peer[rank] = slab)
    uint32_t **d_peer_sync = nullptr;              // device array [world] of pointers to every rank's barrier words
    bool connected = false;
    uint32_t epoch = 0;                            // barrier epoch (two barriers per step)
    uint32_t data_epoch = 0;                       // value the shard flags take this step
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_start = nullptr, ev_rows = nullptr;
    bool pulled = false;
};

namespace {

typedef CUresult (*WriteValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

WriteValue32Fn get_write_value_fn() {
    static WriteValue32Fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<WriteValue32Fn>(p);
    }
    return fn;
}

// publishes a flag value in stream order when the driver offers no stream memory operation
__global__ void fr_set_flag_kernel(uint32_t *flag, uint32_t value) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}

// Flag barrier over peer memory: thread t announces this rank's arrival (epoch) in rank t's barrier word
// [rank], then waits until rank t's arrival shows up in the local word [t].  Everything this GPU wrote before
// (earlier kernels of the stream) is visible to a peer that has seen the word.
__global__ void fr_xbarrier_kernel(uint32_t *const *peer_sync, int rank, int world, uint32_t epoch,
                                   unsigned long long *fail) {
    const int t = threadIdx.x;
    if (t >= world) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peer_sync[t] + rank), "r"(epoch) : "memory");
    const uint32_t *mine = peer_sync[rank] + t;
    const long long t0 = clock64();
    while (true) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
        if ((int32_t)(v - epoch) >= 0) break;
        if (clock64() - t0 > (120ll << 30)) {      // ~60 s: a peer never arrived -- fail the launch, do not hang
            atomicAdd(fail, 1ull);
            __trap();
        }
        __nanosleep(200);
    }
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

extern "C" int fr_exchange_create(fr_ctx *ctx, int rank, int world, const int64_t *row_bounds, int D, int Dp,
                                  fr_exchange **out, unsigned char *handle_out) {
    if (!ctx || !out || !row_bounds || !handle_out) return FR_ERR_ARG;
    *out = nullptr;
    if (world < 1 || world > FR_MAX_RANKS || rank < 0 || rank >= world || D <= 0 || Dp < D || (Dp % 64) != 0)
        return fr_fail(ctx, FR_ERR_ARG, "fr_exchange_create: bad argument (world <= 16, Dp % 64 == 0)");
    if (row_bounds[0] != 0) return fr_fail(ctx, FR_ERR_ARG, "fr_exchange_create: row_bounds must start at 0");
    for (int r = 0; r < world; ++r)
        if (row_bounds[r + 1] < row_bounds[r])
            return fr_fail(ctx, FR_ERR_ARG, "fr_exchange_create: row_bounds must not decrease");
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    fr_exchange *x = new fr_exchange();
    x->ctx = ctx; x->rank = rank; x->world = world; x->D = D; x->Dp = Dp;
    memcpy(x->bounds, row_bounds, sizeof(int64_t) * (world + 1));
    x->N = row_bounds[world];
    const size_t n = (size_t)(x->N > 0 ? x->N : 1);
    x->off_En = 0;
    x->off_Eh = align_up(x->off_En + n * D * 4, 1024);
    x->off_norms = align_up(x->off_Eh + n * Dp * 2, 1024);
    x->off_flags = align_up(x->off_norms + n * 4, 1024);
    x->off_sync = x->off_flags + 256;
    x->slab_bytes = x->off_sync + 256;
    cudaError_t e = cudaMalloc(&x->slab, x->slab_bytes);
    if (e != cudaSuccess) { delete x; return fr_cuda_fail(ctx, e, "fr_exchange_create: slab"); }
    // flags and barrier words start at 0 (epochs start at 1); the rows need no initialisation
    e = cudaMemset(x->slab + x->off_flags, 0, 512);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, x->slab);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&x->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&x->ev_start, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&x->ev_rows, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaMalloc(&x->d_peer_sync, sizeof(uint32_t *) * FR_MAX_RANKS);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        cudaFree(x->slab);
        delete x;
        return fr_cuda_fail(ctx, e, "fr_exchange_create");
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == FR_IPC_HANDLE_BYTES, "CUDA IPC handle size");
    memcpy(handle_out, &h, FR_IPC_HANDLE_BYTES);
    x->peer[rank] = x->slab;
    *out = x;
    return FR_OK;
}

extern "C" int fr_exchange_connect(fr_exchange *x, const unsigned char *handles) {
    if (!x || !handles) return FR_ERR_ARG;
    fr_ctx *ctx = x->ctx;
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)r * FR_IPC_HANDLE_BYTES, FR_IPC_HANDLE_BYTES);
        void *p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return fr_cuda_fail(ctx, e, "fr_exchange_connect: cudaIpcOpenMemHandle (peer access over NVLink)");
        x->peer[r] = (unsigned char *)p;
    }
    uint32_t *ptrs[FR_MAX_RANKS] = {nullptr};
    for (int r = 0; r < x->world; ++r) ptrs[r] = (uint32_t *)(x->peer[r] + x->off_sync);
    FR_CUDA(ctx, cudaMemcpy(x->d_peer_sync, ptrs, sizeof(ptrs), cudaMemcpyHostToDevice));
    x->connected = true;
    return FR_OK;
}

extern "C" void *fr_exchange_rows_f32(fr_exchange *x) { return x ? x->slab + x->off_En : nullptr; }
extern "C" void *fr_exchange_rows_f16(fr_exchange *x) { return x ? x->slab + x->off_Eh : nullptr; }
extern "C" void *fr_exchange_norms(fr_exchange *x) { return x ? x->slab + x->off_norms : nullptr; }

static int xbarrier(fr_exchange *x, cudaStream_t st) {
    fr_ctx *ctx = x->ctx;
    if (x->world == 1) return FR_OK;
    ++x->epoch;
    fr_kt_begin(ctx, "fr_xbarrier_kernel", st);
    fr_xbarrier_kernel<<<1, 32, 0, st>>>(x->d_peer_sync, x->rank, x->world, x->epoch, ctx->d_counters + 5);
    FR_LAUNCH_CHECK(ctx, "fr_xbarrier_kernel");
    return FR_OK;
}

extern "C" int fr_exchange_pull(fr_exchange *x, void *stream) {
    if (!x) return FR_ERR_ARG;
    fr_ctx *ctx = x->ctx;
    if (x->world > 1 && !x->connected) return fr_fail(ctx, FR_ERR_STATE, "fr_exchange_pull: fr_exchange_connect first");
    if (x->pulled) return fr_fail(ctx, FR_ERR_STATE, "fr_exchange_pull: fr_exchange_release the previous step first");
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    int rc = xbarrier(x, st);                       // every rank's rows are in its slab
    if (rc) return rc;
    ++x->data_epoch;
    uint32_t *flags = (uint32_t *)(x->slab + x->off_flags);
    if (x->world > 1) {
        FR_CUDA(ctx, cudaEventRecord(x->ev_start, st));
        FR_CUDA(ctx, cudaStreamWaitEvent(x->copy_stream, x->ev_start, 0));
        WriteValue32Fn wv = get_write_value_fn();
        // fp16 rows first, ring order: what the tensor-core pass consumes, in the order it consumes it
        for (int i = 1; i < x->world; ++i) {
            const int s = (x->rank + i) % x->world;
            const int64_t a = x->bounds[s], b = x->bounds[s + 1];
            if (b > a)
                FR_CUDA(ctx, cudaMemcpyAsync(x->slab + x->off_Eh + (size_t)a * x->Dp * 2,
                                             x->peer[s] + x->off_Eh + (size_t)a * x->Dp * 2, (size_t)(b - a) * x->Dp * 2,
                                             cudaMemcpyDeviceToDevice, x->copy_stream));
            bool done = false;
            if (wv) done = wv((CUstream)x->copy_stream, (CUdeviceptr)(flags + s), x->data_epoch, 0) == CUDA_SUCCESS;
            if (!done) {
                fr_set_flag_kernel<<<1, 1, 0, x->copy_stream>>>(flags + s, x->data_epoch);
                FR_CUDA(ctx, cudaGetLastError());
            }
        }
        // then what the re-rank needs: norms and fp32 rows of every remote shard
        for (int i = 1; i < x->world; ++i) {
            const int s = (x->rank + i) % x->world;
            const int64_t a = x->bounds[s], b = x->bounds[s + 1];
            if (b <= a) continue;
            FR_CUDA(ctx, cudaMemcpyAsync(x->slab + x->off_norms + (size_t)a * 4, x->peer[s] + x->off_norms + (size_t)a * 4,
                                         (size_t)(b - a) * 4, cudaMemcpyDeviceToDevice, x->copy_stream));
            FR_CUDA(ctx, cudaMemcpyAsync(x->slab + x->off_En + (size_t)a * x->D * 4,
                                         x->peer[s] + x->off_En + (size_t)a * x->D * 4, (size_t)(b - a) * x->D * 4,
                                         cudaMemcpyDeviceToDevice, x->copy_stream));
        }
        FR_CUDA(ctx, cudaEventRecord(x->ev_rows, x->copy_stream));
        // one-shot hand-over to the next fr_knn of this context
        ctx->knn_en_event = x->ev_rows;
        ctx->knn_shard_flags = flags;
        ctx->knn_n_shards = x->world;
        ctx->knn_shard_epoch = x->data_epoch;
        ctx->knn_local_shard = x->rank;
        memcpy(ctx->knn_shard_bounds, x->bounds, sizeof(int64_t) * (x->world + 1));
    }
    x->pulled = true;
    return FR_OK;
}

extern "C" int fr_exchange_release(fr_exchange *x, void *stream) {
    if (!x) return FR_ERR_ARG;
    fr_ctx *ctx = x->ctx;
    if (!x->pulled) return fr_fail(ctx, FR_ERR_STATE, "fr_exchange_release: nothing was pulled");
    FR_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (x->world > 1) {
        // this rank's pulls are complete, then every rank's: the rows may be overwritten by the next step
        FR_CUDA(ctx, cudaStreamWaitEvent(st, x->ev_rows, 0));
        if (ctx->knn_en_event == x->ev_rows) {      // no search consumed the hand-over
            ctx->knn_en_event = nullptr;
            ctx->knn_shard_flags = nullptr;
        }
    }
    int rc = xbarrier(x, st);
    x->pulled = false;
    return rc;
}

extern "C" void fr_exchange_destroy(fr_exchange *x) {
    if (!x) return;
    cudaSetDevice(x->ctx->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < x->world; ++r)
        if (r != x->rank && x->peer[r]) cudaIpcCloseMemHandle(x->peer[r]);
    if (x->ctx->knn_en_event == x->ev_rows) { x->ctx->knn_en_event = nullptr; x->ctx->knn_shard_flags = nullptr; }
    cudaFree(x->d_peer_sync);
    cudaFree(x->slab);
    if (x->copy_stream) cudaStreamDestroy(x->copy_stream);
    if (x->ev_start) cudaEventDestroy(x->ev_start);
    if (x->ev_rows) cudaEventDestroy(x->ev_rows);
    delete x;
}
