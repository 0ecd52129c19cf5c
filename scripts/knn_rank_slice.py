"""What ONE rank of a G-GPU all-vs-all search does (BASELINE.json configs[3] and the 8 M end of configs[4]),
measured on one B200: N unit-normalised standard-normal rows x D are the whole (all-gathered) database, the
rank searches its own N/G query rows against it (pipeline.run_hot_path: q0 = rank * N/G, nq = N/G).
    python scripts/knn_rank_slice.py N D k G [rank]     -> one JSON line
    python scripts/knn_rank_slice.py c4                 -> 6 M x 256, k = 20, 8 ranks (whole-human HiFi 30x)
    python scripts/knn_rank_slice.py sweep8m            -> 8 M x {64,128,256,512} x {10,50}, 8 ranks
The job's rows/s at G GPUs is N / (slice time) when the ranks finish together (equal-count shards, same work)."""
import json, os, sys
import torch
sys.path.insert(0, ".")
from fedrann_b200.engine import Engine


def run(eng, N, D, k, G, rank=None):
    rank = G // 2 if rank is None else rank
    nq = N // G
    q0 = rank * nq
    g = torch.Generator(device="cuda")
    g.manual_seed(1000 + D // 64)
    X = torch.randn((N, D), device="cuda", generator=g)     # 8 M x 512: 16 GB + 16 GB (En) + 8 GB (Eh) of 180
    zf = float(os.environ.get("KNN_SLICE_ZERO_FRAC", "0"))   # projected embeddings are 49-67 % exact zeros
    if zf > 0:
        X.mul_((torch.rand((N, D), device="cuda", generator=g) >= zf).to(X.dtype))
    En, Eh, norms = eng.normalize(X)
    del X
    torch.cuda.empty_cache()
    eng.knn(En, Eh, norms, k, q0=q0, nq=min(nq, 4096))          # warm-up: module load, first launches
    torch.cuda.synchronize()
    eng.profile(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    idx, dist = eng.knn(En, Eh, norms, k, q0=q0, nq=nq)
    e1.record()
    torch.cuda.synchronize()
    kt = eng.kernel_times()
    eng.profile(False)
    fb = eng.knn_fallback_rows
    fb_ms = sum(kt[n][1] for n in ("fr_knn_fb_dist_kernel", "fr_knn_fb_select_kernel", "fr_knn_exact_kernel") if n in kt)
    ms = e0.elapsed_time(e1)
    # spot check against the exact fp32 scan on 64 of the slice's rows
    s0 = nq // 3
    i1, d1 = eng.knn(En, Eh, norms, k, q0=q0 + s0, nq=64, mode=1)
    torch.cuda.synchronize()
    ok = bool(torch.equal(i1, idx[s0:s0 + 64]) and torch.equal(d1, dist[s0:s0 + 64]))
    k6 = kt["fr_knn_tc_kernel"][1]
    print(json.dumps({"N_db": N, "D": D, "k": k, "ranks": G, "rank": rank, "q0": q0, "nq": nq, "slice_ms": round(ms, 1),
                      "job_rows_per_s_at_G": round(N / ms * 1e3), "k6_ms": round(k6, 1),
                      "k6_tflops": round(2.0 * nq * N * D / k6 / 1e9, 1),
                      "k7_ms": round(kt["fr_knn_rerank_kernel"][1], 1), "fallback_rows": fb,
                      "fallback_ms": round(fb_ms, 1), "bit_equal_exact_scan_64_rows": ok,
                      "db_bytes_fp32_fp16": N * D * 4 + N * Eh.shape[1] * 2, "zero_frac": zf}), flush=True)
    del En, Eh, norms, idx, dist
    torch.cuda.empty_cache()


if __name__ == "__main__":
    eng = Engine(0)
    a = sys.argv[1:]
    if a and a[0] == "c4":
        run(eng, 6_000_000, 256, 20, 8)
    elif a and a[0] == "sweep8m":
        for D in (64, 128, 256, 512):
            for k in (10, 50):
                run(eng, 8_000_000, D, k, 8)
    elif a and a[0] == "diag":                  # scripts/gpu_knn_diag.sh
        for z in ("0", "0.6"):
            os.environ["KNN_SLICE_ZERO_FRAC"] = z
            for D in (64, 128, 256, 512):
                run(eng, 2_000_000, D, 10, 8)
    else:
        N, D, k, G = (int(x) for x in a[:4])
        run(eng, N, D, k, G, int(a[4]) if len(a) > 4 else None)
