"""End-to-end (host buffers in, host results out) timing of run_hot_path_host on the c2 workload for a
few chunk / kNN-phase schedules:   python scripts/e2e_tune.py   -> one JSON line per schedule"""
import json, sys
import torch
sys.path.insert(0, ".")
import bench
from fedrann_b200.engine import Engine
from fedrann_b200.pipeline import HostBuffers, run_hot_path, run_hot_path_host


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    w = bench.WORKLOADS[name]
    wl = bench.Workload(name, 0, 1, dev)
    eng = Engine(0)
    eng.build_table(wl.lib.all_codes(), w["k"], sampler=(w["seed"], w["fraction"]))
    P_dev = eng.upload_projection(wl.P)
    offsets_d = torch.from_numpy(wl.offsets).to(dev)
    nk, k_nn = wl.lib.nk, w["knn_k"]
    ref = run_hot_path(eng, wl.bases, offsets_d, P_dev, nk, k_nn)
    hb = HostBuffers.from_device(wl.bases, wl.offsets, 2 * wl.n_reads_local, k_nn)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    # device-resident: the incremental search alone against the one-shot search
    En, Eh, norms = eng.normalize(ref.embedding)
    N = En.shape[0]

    def timed(fn, reps=5):
        fn()
        out = []
        for _ in range(reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); r = fn(); e1.record()
            torch.cuda.synchronize()
            out.append(e0.elapsed_time(e1))
        return sorted(out)[len(out) // 2], r

    ms0, (i0, d0) = timed(lambda: eng.knn(En, Eh, norms, k_nn))
    print(json.dumps({"knn": "one shot", "ms": round(ms0, 3)}), flush=True)
    for phases in ():
        bnd = [(N * j // phases) // 2 * 2 for j in range(phases + 1)]

        def inc():
            eng.knn_inc_begin(En, Eh, norms, k_nn, bnd)
            for j in range(phases):
                eng.knn_inc_phase(j)
            return eng.knn_inc_finish()
        ms, (i1, d1) = timed(inc)
        print(json.dumps({"knn": "incremental", "phases": phases, "ms": round(ms, 3), "fallback_rows": eng.knn_fallback_rows,
                          "equal": bool(torch.equal(i0, i1) and torch.equal(d0, d1))}), flush=True)
    del En, Eh, norms
    best = (0.125, 0.25, 0.375, 0.5, 0.625, 0.8)
    schedules = [
        dict(chunks=8, knn_overlap=False, direct_out=False),
        dict(chunks=8, knn_overlap=False),
        dict(chunks=best, knn_phase_ends=(5, 6), direct_out=False),
        dict(chunks=best, knn_phase_ends=(5, 6)),
        dict(chunks=(0.125, 0.25, 0.375, 0.5, 0.625, 0.8, 0.92), knn_phase_ends=(5, 6, 7)),
        dict(chunks=(0.15, 0.3, 0.45, 0.6, 0.8), knn_phase_ends=(4, 5)),
    ]
    for kw in schedules:
        for _ in range(2):
            run_hot_path_host(eng, hb, P_dev, nk, k_nn, **kw)
        flush.zero_()
        tr = []
        run_hot_path_host(eng, hb, P_dev, nk, k_nn, trace=tr, **kw)
        torch.cuda.synchronize()
        t0 = tr[0][1]
        print(json.dumps({"schedule": {k: (list(v) if isinstance(v, tuple) else v) for k, v in kw.items()},
                          "timeline_ms": {lab: round(t0.elapsed_time(ev), 2) for lab, ev in tr[1:]}}), flush=True)
    for kw in schedules:
        for _ in range(2):
            res = run_hot_path_host(eng, hb, P_dev, nk, k_nn, **kw)
        same = bool(torch.equal(hb.nbr_idx, ref.nbr_idx.cpu()) and torch.equal(hb.nbr_dist, ref.nbr_dist.cpu()))
        evs = []
        for _ in range(8):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run_hot_path_host(eng, hb, P_dev, nk, k_nn, **kw)
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        eng.profile(True)
        run_hot_path_host(eng, hb, P_dev, nk, k_nn, **kw)
        torch.cuda.synchronize()
        kt = {k: (v[0], round(v[1], 3)) for k, v in eng.kernel_times().items()}
        eng.profile(False)
        ms = sorted(a.elapsed_time(b) for a, b in evs)
        print(json.dumps({"schedule": {k: (list(v) if isinstance(v, tuple) else v) for k, v in kw.items()},
                          "ms_median": round(ms[len(ms) // 2], 3), "ms_min": round(ms[0], 3),
                          "reads_per_s": round(wl.n_reads_total / ms[len(ms) // 2] * 1e3),
                          "kernels(launches, ms)": kt, "fallback_rows": eng.knn_fallback_rows, "equal_to_device_path": same}), flush=True)


if __name__ == "__main__":
    main()
