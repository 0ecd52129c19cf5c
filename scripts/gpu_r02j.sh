#!/bin/bash
# round 2, call J: launch list (this library's kernels) and one `ncu --set full` of K6 for the DEFAULT bench command
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
timeout 100 $CMD > gpurun_out/r02j_plain.json 2> gpurun_out/r02j_plain.err &&
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:fr_ -c 120 --csv --log-file gpurun_out/r02j_launches_c3.csv $CMD > gpurun_out/r02j_ncu_launches.log 2>&1 &&
timeout 150 ncu --set full --clock-control none --import-source on -k regex:fr_knn_tc_kernel -s 3 -c 1 -f -o gpurun_out/r02j_prof_k6_c3 $CMD > gpurun_out/r02j_ncu_k6.log 2>&1
ls -la gpurun_out/r02j*; tail -3 gpurun_out/r02j_launches_c3.csv | cut -c1-300
