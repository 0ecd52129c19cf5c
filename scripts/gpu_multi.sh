#!/bin/bash
# multi-GPU check: bench at N ranks + CLI determinism (1 rank vs N ranks)
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_n$N.err
python - <<'PY'
import sys; sys.path.insert(0, '.')
from fedrann_b200 import synthetic as S
g = S.random_genome(300000, 3)
rs = S.simulate_reads(g, 15, 3, mean_len=6000, len_model="lognormal", min_len=100, sub=0.02, ins=0.01, dele=0.01, n_frac=0.001)
open('/tmp/reads.fasta', 'wb').write(rs.fasta())
print('reads', rs.n_reads)
PY
ARGS="-i /tmp/reads.fasta -k 15 --kmer-sample-fraction 0.05 --seed 602 -n 128 --nndescent-n-neighbors 20"
python -m fedrann_b200 $ARGS -o /tmp/out1 > gpurun_out/cli1.log 2>&1; echo "cli1 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 -m fedrann_b200 $ARGS -o /tmp/outN > gpurun_out/cliN.log 2>&1; echo "cliN rc=$?"
tail -3 gpurun_out/cliN.log
md5sum /tmp/out1/overlaps.tsv /tmp/outN/overlaps.tsv; wc -l /tmp/out1/overlaps.tsv
cmp /tmp/out1/overlaps.tsv /tmp/outN/overlaps.tsv && echo "CLI output identical at 1 and $N ranks"
python -c "
import json;d=json.load(open('gpurun_out/bench_n$N.json'));print(d['n_gpus'],d['value'],d['e2e']['value'],d['ms_per_step']);print(d['stage_ms'])"
