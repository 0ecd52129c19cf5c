#!/bin/bash
# round 2, call F (2 GPUs): (1) the peer-memory exchange: 2 ranks = 1 rank bit for bit, c3 / c2 at 2 ranks with both
# transports; (2) K6 with a candidate buffer per epilogue warp; (3) core-bitmap sizes for generic libraries;
# (4) ncu --set full of the extract kernel and of K6 (source-level), each after the plain run of the same command
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02f_topo.txt 2>&1
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r02f_pytest_multi.log 2>&1; rc=$?
echo "pytest multi rc=$rc" >> gpurun_out/r02f_pytest_multi.log; tail -12 gpurun_out/r02f_pytest_multi.log
run2() {  # name, exchange, workload, steps
  FEDRANN_EXCHANGE=$2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) \
    bench.py --gpus 2 --workload $3 --steps $4 --warmup 3 --no-cpu-baseline > gpurun_out/r02f_bench_$1.json 2> gpurun_out/r02f_bench_$1.err; echo "$1 rc=$?"
}
run2 c3_n2_peer peer c3 3
run2 c3_n2_nccl nccl c3 3
run2 c2_n2_peer peer c2 10
run2 c2_n2_nccl nccl c2 10
export CUDA_VISIBLE_DEVICES=0
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02f_pytest_parity.log 2>&1; echo "pytest parity rc=$?" >> gpurun_out/r02f_pytest_parity.log; tail -3 gpurun_out/r02f_pytest_parity.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02f_bench_c3_n1.json 2> gpurun_out/r02f_bench_c3_n1.err; echo "c3 n1 rc=$?"
for lg in 26 27 28; do
  FR_CORE_LG=$lg timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02f_bench_c2_corelg$lg.json 2> gpurun_out/r02f_bench_c2_corelg$lg.err; echo "c2 lg=$lg rc=$?"
done
for D in 64 128 256 512; do
  timeout 120 python scripts/knn_rank_slice.py 2000000 $D 10 8 >> gpurun_out/r02f_knn_rank_slice_2M.jsonl 2>> gpurun_out/r02f_slice.err
done
FR_KNN_DEBUG=1 timeout 120 python scripts/knn_rank_slice.py 2000000 256 10 8 > gpurun_out/r02f_knn_dbg_256.log 2>&1
CMD1="python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
$CMD1 > gpurun_out/r02f_plain1.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fr_extract_kernel -s 3 -c 1 -f -o gpurun_out/r02f_prof_extract $CMD1 > gpurun_out/r02f_ncu_extract.log 2>&1
CMD2="python scripts/knn_rank_slice.py 400000 256 10 8"
$CMD2 > gpurun_out/r02f_plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fr_knn_tc_kernel -s 1 -c 1 -f -o gpurun_out/r02f_prof_k6 $CMD2 > gpurun_out/r02f_ncu_k6.log 2>&1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02f_bench_*.json")):
    try:
        d = json.load(open(f))
        r = d["roofline"]
        print(f, d["n_gpus"], round(d["value"]), round(d["ms_per_step"], 2), (d.get("e2e") or {}).get("value"), d.get("parity_ok"), d.get("digest", "")[:12], round(r["achieved"]), round(r["frac"], 3))
        print("   ", {k: round(v, 3) for k, v in d["stage_ms"].items()})
        print("   ", {k: round(v, 3) for k, v in d["kernels_ms_per_step"].items()})
        print("   ", {k: (round(v["frac"], 3), round(v["kernel_ms"], 3)) for k, v in d["stage_rooflines"].items() if v and "frac" in v})
    except Exception as e:
        print(f, "unreadable", e)
for l in open("gpurun_out/r02f_knn_rank_slice_2M.jsonl"):
    d = json.loads(l); print(d["D"], d["k6_ms"], d["k6_tflops"], d["k7_ms"], d["fallback_rows"], d["bit_equal_exact_scan_64_rows"])
PY
grep "fr_knn dbg" gpurun_out/r02f_knn_dbg_256.log | tail -3
ls -la gpurun_out/*.ncu-rep
