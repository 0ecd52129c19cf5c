#!/bin/bash
# round 2, call H (last GPU minutes): K6 without run-time debug counters + per-chunk sparse path: parity, c3, slices,
# one rank's share of c4 (750 k x 6 M x 256, k = 20), long-stream ncu capture of K6
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02h_pytest_parity.log 2>&1; rc=$?
echo "pytest rc=$rc" >> gpurun_out/r02h_pytest_parity.log; tail -3 gpurun_out/r02h_pytest_parity.log
[ $rc -ne 0 ] && exit 1
timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02h_bench_c3.json 2> gpurun_out/r02h_bench_c3.err; echo "c3 rc=$?"
for D in 64 256; do
  timeout 60 python scripts/knn_rank_slice.py 2000000 $D 10 8 >> gpurun_out/r02h_knn_rank_slice_2M.jsonl 2>> gpurun_out/r02h_slice.err
done
timeout 90 python scripts/knn_rank_slice.py c4 > gpurun_out/r02h_c4_rank_slice.jsonl 2>> gpurun_out/r02h_slice.err; echo "c4 slice rc=$?"
CMD2="python scripts/knn_rank_slice.py 1000000 256 50 8"
timeout 60 $CMD2 > gpurun_out/r02h_plain2.log 2>&1 &&
timeout 150 ncu --section SpeedOfLight --section SourceCounters --section WarpStateStats --section MemoryWorkloadAnalysis --clock-control none --import-source on -k regex:fr_knn_tc_kernel -s 1 -c 1 -f -o gpurun_out/r02h_prof_k6_long $CMD2 > gpurun_out/r02h_ncu_k6.log 2>&1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02h_bench_*.json")):
    try:
        d = json.load(open(f))
        r = d["roofline"]
        print(f, d["n_gpus"], round(d["value"]), round(d["ms_per_step"], 2), (d.get("e2e") or {}).get("value"), d.get("parity_ok"), d.get("digest", "")[:12], round(r["achieved"]), round(r["frac"], 3), d["knn_fallback_rows"], d["clocks"])
        print("   ", {k: round(v, 3) for k, v in d["kernels_ms_per_step"].items()})
    except Exception as e:
        print(f, "unreadable", e)
for f in ("gpurun_out/r02h_knn_rank_slice_2M.jsonl", "gpurun_out/r02h_c4_rank_slice.jsonl", "gpurun_out/r02h_plain2.log"):
    for l in open(f):
        try:
            d = json.loads(l); print(d["N_db"], d["D"], d["k"], d["k6_ms"], d["k6_tflops"], d["k7_ms"], d["fallback_rows"], d["bit_equal_exact_scan_64_rows"])
        except Exception:
            pass
PY
