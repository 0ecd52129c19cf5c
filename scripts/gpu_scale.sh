#!/bin/bash
# strong-scaling bench lines at N ranks (the driver's launch line); WORKLOAD=c3 STEPS=3 for the chr1 scale
N=${1:-8}
W=${WORKLOAD:-c2}
S=${STEPS:-10}
mkdir -p gpurun_out
for n in 1 2 4 8; do
  [ $n -gt $N ] && break
  if [ $n -eq 1 ]; then
    python bench.py --gpus 1 --workload $W --steps $S --warmup 3 --no-cpu-baseline > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n)) bench.py --gpus $n --workload $W --steps $S --warmup 3 --no-cpu-baseline > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  echo "n=$n rc=$?"
  python -c "
import json;d=json.load(open('gpurun_out/scale_n$n.json'));print(d['n_gpus'],round(d['value']),round(d['e2e']['value']),d['ms_per_step'],d['stage_ms'])"
done
