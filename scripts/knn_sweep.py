"""kNN-only sweep (BASELINE.json configs[4]): N unit-normalised standard-normal rows x D, exact cosine kNN.
    python scripts/knn_sweep.py [N ...]      -> one JSON line per (N, D, k)
    (KNN_SWEEP_D / KNN_SWEEP_K: comma-separated subsets)"""
import json, os, sys
import torch
sys.path.insert(0, ".")
from fedrann_b200.engine import Engine


def run(eng, N, D, k):
    g = torch.Generator(device="cuda")
    g.manual_seed(1000 + D // 64)
    X = torch.randn((N, D), device="cuda", generator=g)
    En, Eh, norms = eng.normalize(X)
    del X
    eng.knn(En, Eh, norms, k)                      # warm-up (workspace allocation)
    torch.cuda.synchronize()
    eng.profile(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    idx, dist = eng.knn(En, Eh, norms, k)
    e1.record()
    torch.cuda.synchronize()
    kt = eng.kernel_times()
    eng.profile(False)
    fb = eng.knn_fallback_rows                     # (the spot check below resets the counter)
    fb_ms = sum(kt[n][1] for n in ("fr_knn_fb_dist_kernel", "fr_knn_fb_select_kernel", "fr_knn_exact_kernel") if n in kt)
    ms = e0.elapsed_time(e1)
    # spot check against the exact scan on 64 query rows
    q0 = N // 3
    i1, d1 = eng.knn(En, Eh, norms, k, q0=q0, nq=64, mode=1)
    torch.cuda.synchronize()
    ok = bool(torch.equal(i1, idx[q0:q0 + 64]) and torch.equal(d1, dist[q0:q0 + 64]))
    k6 = kt["fr_knn_tc_kernel"][1]
    print(json.dumps({"N": N, "D": D, "k": k, "ms": round(ms, 2), "rows_per_s": round(N / ms * 1e3),
                      "k6_ms": round(k6, 2), "k6_tflops": round(2.0 * N * N * D / k6 / 1e9, 1),
                      "k7_ms": round(kt["fr_knn_rerank_kernel"][1], 2), "fallback_rows": fb, "fallback_ms": round(fb_ms, 2),
                      "bit_equal_exact_scan_64_rows": ok}), flush=True)
    del En, Eh, norms, idx, dist
    torch.cuda.empty_cache()


if __name__ == "__main__":
    Ns = [int(a) for a in sys.argv[1:]] or [1_000_000]
    eng = Engine(0)
    for N in Ns:
        for D in [int(x) for x in os.environ.get("KNN_SWEEP_D", "64,128,256,512").split(",")]:
            for k in [int(x) for x in os.environ.get("KNN_SWEEP_K", "10,50").split(",")]:
                run(eng, N, D, k)
