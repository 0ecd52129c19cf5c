"""Incremental vs one-shot kNN on random unit rows: python scripts/knn_inc_check.py [N D k phases]
(FR_KNN_DEBUG=1 prints the per-launch cycle counters)"""
import sys, torch
sys.path.insert(0, ".")
from fedrann_b200.engine import Engine

N, D, k, phases = (int(a) for a in (sys.argv[1:5] if len(sys.argv) >= 5 else (50000, 500, 50, 8)))
eng = Engine(0)
g = torch.Generator(device="cuda"); g.manual_seed(5)
En, Eh, norms = eng.normalize(torch.randn((N, D), device="cuda", generator=g))
i0, d0 = eng.knn(En, Eh, norms, k)
bnd = [(N * j // phases) // 2 * 2 for j in range(phases + 1)]
for rep in range(2):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    eng.knn_inc_begin(En, Eh, norms, k, bnd)
    for j in range(phases):
        eng.knn_inc_phase(j)
    i1, d1 = eng.knn_inc_finish()
    e1.record(); torch.cuda.synchronize()
    print("incremental %d phases: %.3f ms, equal %s, fallback rows %d" % (
        phases, e0.elapsed_time(e1), bool(torch.equal(i0, i1) and torch.equal(d0, d1)), eng.knn_fallback_rows))
