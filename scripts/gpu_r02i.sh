#!/bin/bash
# round 2, call I (final state): full GPU test suite, smoke, c2 bench, launch list of the default bench command
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/r02i_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02i_pytest_gpu.log; tail -3 gpurun_out/r02i_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02i_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02i_smoke.log
timeout 120 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02i_bench_c2.json 2> gpurun_out/r02i_bench_c2.err; echo "c2 rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
timeout 120 $CMD > gpurun_out/r02i_plain.log 2>&1 &&
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02i_launches_c3.csv $CMD > gpurun_out/r02i_ncu_launches.log 2>&1
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02i_bench_c2.json")); r = d["roofline"]
print("c2", round(d["value"]), round(d["ms_per_step"], 2), d["e2e"]["value"], d.get("parity_ok"), d.get("digest", "")[:12], round(r["achieved"]), round(r["frac"], 3))
print({k: round(v, 3) for k, v in d["kernels_ms_per_step"].items()})
print({k: (round(v["frac"], 3), round(v["kernel_ms"], 3)) for k, v in d["stage_rooflines"].items() if v and "frac" in v})
PY
tail -5 gpurun_out/r02i_launches_c3.csv
