#!/bin/bash
# round 2, call A: GPU parity tests (incl. the at-scale kNN cases), bench on the new default (c3) with the
# digest / parity block, the host-only reference arm, c2 for comparison with round 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r02a_smi.csv
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02a_pytest_gpu.log
tail -5 gpurun_out/r02a_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02a_bench_c3.json 2> gpurun_out/r02a_bench_c3.err; echo "c3 rc=$?"
tail -3 gpurun_out/r02a_bench_c3.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02a_bench_c3_ref.json 2> gpurun_out/r02a_bench_c3_ref.err; echo "ref rc=$?"
timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02a_bench_c2.json 2> gpurun_out/r02a_bench_c2.err; echo "c2 rc=$?"
python - <<'PY'
import json
for f in ("r02a_bench_c3", "r02a_bench_c3_ref", "r02a_bench_c2"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f))
        print(f, round(d["value"]), d["ms_per_step"], (d.get("e2e") or {}).get("value"), d.get("parity_ok"), d.get("digest"),
              d.get("parity_rows_checked"), (d.get("roofline") or {}).get("frac"), d.get("stage_ms"), d.get("knn_fallback_rows"))
        print("   kernels", d.get("kernels_ms_per_step"))
        print("   stage_roofs", {k: (v or {}).get("frac") for k, v in (d.get("stage_rooflines") or {}).items()})
    except Exception as e:
        print(f, "unreadable", e)
PY
