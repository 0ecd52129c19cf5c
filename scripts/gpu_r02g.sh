#!/bin/bash
# round 2, call G (8 GPUs): determinism at 8 / 3 ranks, c3 at 8 and 4 ranks, the north-star target c4 (3 M reads,
# 6 M x 256 rows, k = 20) end to end on 8 ranks, c3 at 8 ranks with the NCCL all-gather for comparison
mkdir -p gpurun_out
port() { echo $((29700 + RANDOM % 200)); }
for n in 8 3; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $(port) \
    scripts/multi_check.py --out gpurun_out/r02g_multi_check_n$n.json > gpurun_out/r02g_multi_check_n$n.log 2>&1; echo "multi_check n=$n rc=$?"
done
cat gpurun_out/r02g_multi_check_n8.json gpurun_out/r02g_multi_check_n3.json; echo
run() {  # name, n, exchange, workload, steps
  FEDRANN_EXCHANGE=$3 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $(port) \
    bench.py --gpus $2 --workload $4 --steps $5 --warmup 3 --no-cpu-baseline > gpurun_out/r02g_bench_$1.json 2> gpurun_out/r02g_bench_$1.err; echo "$1 rc=$?"
}
run c3_n8_peer 8 peer c3 5
run c4_n8_peer 8 peer c4 3
[ -s gpurun_out/r02g_bench_c4_n8_peer.json ] || tail -30 gpurun_out/r02g_bench_c4_n8_peer.err
run c3_n4_peer 4 peer c3 5
run c3_n8_nccl 8 nccl c3 5
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02g_bench_*.json")):
    try:
        d = json.load(open(f))
        r = d["roofline"]
        print(f, d["n_gpus"], round(d["value"]), round(d["ms_per_step"], 2), (d.get("e2e") or {}).get("value"), d.get("parity_ok"), d.get("digest", "")[:12], round(r["achieved"]), round(r["frac"], 3), d["knn_fallback_rows"], d["clocks"])
        print("   ", {k: round(v, 3) for k, v in d["stage_ms"].items()})
        print("   ", {k: round(v, 3) for k, v in d["kernels_ms_per_step"].items()})
    except Exception as e:
        print(f, "unreadable", e)
PY
