#!/bin/bash
# round 2, call D: core-bitmap extraction filter + hierarchical sparse epilogue path + adaptive host-path schedule
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02d_pytest_gpu.log 2>&1; rc=$?
echo "pytest rc=$rc" >> gpurun_out/r02d_pytest_gpu.log; tail -4 gpurun_out/r02d_pytest_gpu.log
[ $rc -ne 0 ] && exit 1
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02d_bench_c3.json 2> gpurun_out/r02d_bench_c3.err; echo "c3 rc=$?"
timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02d_bench_c2.json 2> gpurun_out/r02d_bench_c2.err; echo "c2 rc=$?"
FR_EXTRACT_FILTER=old timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02d_bench_c3_oldfilter.json 2> gpurun_out/r02d_bench_c3_oldfilter.err; echo "c3 old rc=$?"
FR_EXTRACT_FILTER=old timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02d_bench_c2_oldfilter.json 2> gpurun_out/r02d_bench_c2_oldfilter.err; echo "c2 old rc=$?"
for D in 64 128 256 512; do
  timeout 120 python scripts/knn_rank_slice.py 2000000 $D 10 8 >> gpurun_out/r02d_knn_rank_slice_2M.jsonl 2>> gpurun_out/r02d_slice.err
done
FR_KNN_DEBUG=1 timeout 120 python scripts/knn_rank_slice.py 2000000 256 10 8 > gpurun_out/r02d_knn_dbg_256.log 2>&1
FR_KNN_DEBUG=1 timeout 120 python scripts/knn_rank_slice.py 2000000 64 10 8 > gpurun_out/r02d_knn_dbg_64.log 2>&1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02d_bench_*.json")):
    try:
        d = json.load(open(f))
        r = d["roofline"]
        print(f, round(d["value"]), round(d["ms_per_step"], 2), (d.get("e2e") or {}).get("value"), d.get("parity_ok"), d.get("digest", "")[:12], r["kernel"], round(r["achieved"]), round(r["frac"], 3))
        print("   ", {k: round(v, 2) for k, v in d["kernels_ms_per_step"].items()}, d["knn_fallback_rows"], d["clocks"])
        print("   ", {k: (round(v["frac"], 3), round(v["kernel_ms"], 3)) for k, v in d["stage_rooflines"].items() if v and "frac" in v})
    except Exception as e:
        print(f, "unreadable", e)
for l in open("gpurun_out/r02d_knn_rank_slice_2M.jsonl"):
    d = json.loads(l); print(d["D"], d["k6_ms"], d["k6_tflops"], d["k7_ms"], d["fallback_rows"], d["bit_equal_exact_scan_64_rows"])
PY
grep "fr_knn dbg" gpurun_out/r02d_knn_dbg_256.log gpurun_out/r02d_knn_dbg_64.log | tail -6
