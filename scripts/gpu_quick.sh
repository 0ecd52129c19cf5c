#!/bin/bash
# quick GPU check: parity tests + one bench line
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_quick.err
python -c "
import json;d=json.load(open('gpurun_out/bench_quick.json'));print(d['value'],d['e2e']['value'],d['roofline']['frac']);print(d['kernels_ms_per_step']);print(d['stage_ms'])"
