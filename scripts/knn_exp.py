"""K6 timing experiments (FR_KNN_EXP, results invalid): what bounds the epilogue on long streams.
   FR_KNN_EXP=1 no threshold filter (TMEM loads only), =2 no TMEM loads either (barriers + MMA only)."""
import json, os, sys
import torch
sys.path.insert(0, ".")
from fedrann_b200.engine import Engine

eng = Engine(0)
N, nq = 2_000_000, 250_000
for D in (64, 128, 256, 512):
    g = torch.Generator(device="cuda"); g.manual_seed(1000 + D // 64)
    X = torch.randn((N, D), device="cuda", generator=g)
    En, Eh, norms = eng.normalize(X); del X
    out = {"N_db": N, "nq": nq, "D": D}
    for exp in ("0", "1", "2"):
        if exp == "0": os.environ.pop("FR_KNN_EXP", None)
        else: os.environ["FR_KNN_EXP"] = exp
        eng.knn(En, Eh, norms, 10, q0=N // 2, nq=4096); torch.cuda.synchronize()
        eng.profile(True)
        eng.knn(En, Eh, norms, 10, q0=N // 2, nq=nq); torch.cuda.synchronize()
        k6 = eng.kernel_times()["fr_knn_tc_kernel"][1]; eng.profile(False)
        out[f"exp{exp}_k6_ms"] = round(k6, 1); out[f"exp{exp}_tflops"] = round(2.0 * nq * N * D / k6 / 1e9, 1)
    print(json.dumps(out), flush=True)
    del En, Eh, norms; torch.cuda.empty_cache()
