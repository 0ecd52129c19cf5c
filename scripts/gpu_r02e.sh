#!/bin/bash
# round 2, call E (2 GPUs): the peer-memory exchange -- determinism first (2 ranks = 1 rank, peer and nccl
# transports), then c3 / c2 at 2 ranks with both transports
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02e_topo.txt 2>&1
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r02e_pytest_multi.log 2>&1; rc=$?
echo "pytest rc=$rc" >> gpurun_out/r02e_pytest_multi.log; tail -15 gpurun_out/r02e_pytest_multi.log
run() {  # name, exchange, workload, steps
  FEDRANN_EXCHANGE=$2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) \
    bench.py --gpus 2 --workload $3 --steps $4 --warmup 3 --no-cpu-baseline > gpurun_out/r02e_bench_$1.json 2> gpurun_out/r02e_bench_$1.err; echo "$1 rc=$?"
}
run c3_n2_peer peer c3 3
[ $rc -ne 0 ] && tail -20 gpurun_out/r02e_bench_c3_n2_peer.err
run c3_n2_nccl nccl c3 3
run c2_n2_peer peer c2 10
run c2_n2_nccl nccl c2 10
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02e_bench_*.json")):
    try:
        d = json.load(open(f))
        r = d["roofline"]
        print(f, d["n_gpus"], round(d["value"]), round(d["ms_per_step"], 2), (d.get("e2e") or {}).get("value"), d.get("parity_ok"), d.get("digest", "")[:12], round(r["achieved"]), round(r["frac"], 3))
        print("   ", {k: round(v, 3) for k, v in d["stage_ms"].items()})
        print("   ", {k: round(v, 3) for k, v in d["kernels_ms_per_step"].items()})
    except Exception as e:
        print(f, "unreadable", e)
PY
