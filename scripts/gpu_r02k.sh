#!/bin/bash
# round 2, call K: `ncu --set full` of the final K6 on c2 (a 2.7 ms launch; the c3 launch is 380 ms x 37 passes)
mkdir -p gpurun_out
CMD="python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
timeout 60 $CMD > gpurun_out/r02k_plain.json 2> gpurun_out/r02k_plain.err &&
timeout 100 ncu --set full --clock-control none --import-source on -k regex:fr_knn_tc_kernel -s 3 -c 1 -f -o gpurun_out/r02k_prof_k6_c2 $CMD > gpurun_out/r02k_ncu_k6.log 2>&1
ls -la gpurun_out/r02k*
