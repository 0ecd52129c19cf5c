#!/bin/bash
# where K6's time goes on long streams (2 M database rows, one of 8 ranks' query slice), per embedding width:
# per-role wait-cycle counters (FR_KNN_DEBUG), clocks and power sampled alongside, dense Gaussian vs 60 % zero rows
mkdir -p gpurun_out
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_event_reasons.active --format=csv,noheader -lms 200 > gpurun_out/diag_clocks.csv &
SMI=$!
timeout 200 python scripts/knn_rank_slice.py diag > gpurun_out/knn_diag_plain.log 2>&1
FR_KNN_DEBUG=1 timeout 200 python scripts/knn_rank_slice.py diag > gpurun_out/knn_diag_debug.log 2>&1
kill $SMI
sort gpurun_out/diag_clocks.csv | uniq -c | sort -rn | head -12
cut -c1-260 gpurun_out/knn_diag_plain.log; cat gpurun_out/knn_diag_debug.log | cut -c1-400
