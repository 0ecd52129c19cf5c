#!/bin/bash
# round 2, call C: pair-mode K6 with default-semantics remote arrives -- parity, then A/B on c3 / c2 / 2 M slices
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "knn" > gpurun_out/r02c_pytest_knn.log 2>&1; rc=$?
echo "pytest knn rc=$rc" >> gpurun_out/r02c_pytest_knn.log; tail -3 gpurun_out/r02c_pytest_knn.log
[ $rc -ne 0 ] && exit 1
for pair in 1 0; do
  FR_KNN_PAIR=$pair timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02c_bench_c3_pair$pair.json 2> gpurun_out/r02c_bench_c3_pair$pair.err; echo "c3 pair=$pair rc=$?"
  FR_KNN_PAIR=$pair timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02c_bench_c2_pair$pair.json 2> gpurun_out/r02c_bench_c2_pair$pair.err; echo "c2 pair=$pair rc=$?"
done
for D in 64 128 256 512; do
  FR_KNN_PAIR=1 timeout 120 python scripts/knn_rank_slice.py 2000000 $D 10 8 >> gpurun_out/r02c_knn_rank_slice_2M_pair1.jsonl 2>> gpurun_out/r02c_slice.err
done
FR_KNN_DEBUG=1 timeout 120 python scripts/knn_rank_slice.py 2000000 256 10 8 > gpurun_out/r02c_knn_dbg_256.log 2>&1
FR_KNN_DEBUG=1 timeout 120 python scripts/knn_rank_slice.py 2000000 64 10 8 > gpurun_out/r02c_knn_dbg_64.log 2>&1
FR_KNN_DEBUG=1 FR_KNN_PAIR=0 timeout 120 python scripts/knn_rank_slice.py 2000000 256 10 8 > gpurun_out/r02c_knn_dbg_256_pair0.log 2>&1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02c_bench_*.json")):
    try:
        d = json.load(open(f))
        r = d["roofline"]
        print(f, round(d["value"]), round(d["ms_per_step"], 2), d.get("parity_ok"), d.get("digest", "")[:12], r["kernel"], round(r["achieved"]), round(r["frac"], 3),
              {k: round(v, 2) for k, v in d["kernels_ms_per_step"].items() if "knn" in k}, d["knn_fallback_rows"], d["clocks"])
    except Exception as e:
        print(f, "unreadable", e)
for f in sorted(glob.glob("gpurun_out/r02c_knn_rank_slice*.jsonl")):
    for l in open(f):
        d = json.loads(l); print(f[-12:], d["D"], d["k6_ms"], d["k6_tflops"], d["k7_ms"], d["fallback_rows"], d["fallback_ms"], d["bit_equal_exact_scan_64_rows"])
PY
grep "fr_knn dbg" gpurun_out/r02c_knn_dbg_256.log gpurun_out/r02c_knn_dbg_64.log gpurun_out/r02c_knn_dbg_256_pair0.log | tail -9
