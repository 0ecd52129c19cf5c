#!/bin/bash
# round 2, call L (last GPU seconds): parity of the final build, then the sparse-path criterion on the chr1 stream length
mkdir -p gpurun_out
timeout 60 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "knn" > gpurun_out/r02l_pytest_knn.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02l_pytest_knn.log; tail -2 gpurun_out/r02l_pytest_knn.log
for sm in 8 32 16 64; do
  FR_KNN_SPARSE_MAX=$sm timeout 40 python scripts/knn_rank_slice.py 1000000 256 50 8 2>/dev/null | sed "s/^/sparse_max=$sm /" >> gpurun_out/r02l_sparse_max.log
done
for sm in 8 32; do
  FR_KNN_SPARSE_MAX=$sm timeout 30 python scripts/knn_rank_slice.py 50000 500 50 1 2>/dev/null | sed "s/^/sparse_max=$sm /" >> gpurun_out/r02l_sparse_max.log
done
cut -c1-260 gpurun_out/r02l_sparse_max.log
