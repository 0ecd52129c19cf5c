"""Dev check of fr_knn mode 0 (tcgen05) against mode 1 (exact scan): agreement + timing."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from fedrann_b200.engine import Engine

def run(N, D, k, seed=0, sparse=True, check=True):
    eng = Engine(0)
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    X = torch.randn((N, D), device="cuda", generator=g)
    if sparse:
        X[torch.rand((N, D), device="cuda", generator=g) < 0.5] = 0
    X[5] = 0; X[11] = X[10]
    En, Eh, norms = eng.normalize(X)
    torch.cuda.synchronize()
    eng.profile(True)
    t0 = time.time()
    i0, d0 = eng.knn(En, Eh, norms, k, mode=0)
    torch.cuda.synchronize()
    t1 = time.time()
    kt = eng.kernel_times()
    fb = eng.knn_fallback_rows
    msg = f"N={N} D={D} k={k}: wall {1e3*(t1-t0):.2f} ms fallback_rows={fb} " + " ".join(f"{a}={b[1]:.3f}ms" for a, b in kt.items())
    if "fr_knn_tc_kernel" in kt:
        msg += f" | TC {2.0*N*N*D/kt['fr_knn_tc_kernel'][1]/1e9:.1f} TFLOP/s"
    if check:
        i1, d1 = eng.knn(En, Eh, norms, k, mode=1)
        torch.cuda.synchronize()
        msg += f" | idx agree {(i0==i1).float().mean().item():.6f} dist maxdiff {(d0-d1).abs().max().item():.3g} bit-equal {torch.equal(d0,d1)}"
    print(msg, flush=True)
    eng.close()

if __name__ == "__main__":
    run(300, 64, 20)
    run(1500, 500, 50)
    run(700, 48, 10)
    run(5000, 256, 50)
    run(20000, 500, 50, check=True)
    run(50000, 500, 50, check=False)
    run(50000, 500, 50, check=False)
    run(100000, 256, 20, check=False)
    run(100000, 64, 10, check=False)
