#!/bin/bash
# K6 selection-window experiments on the c2 embeddings: time + fallback rows per (klo, khi)
for CFG in "80 128" "80 96" "80 88" "72 84" "66 76"; do
set -- $CFG
FR_KNN_KLO=$1 FR_KNN_KHI=$2 python - <<'PY'
import os, sys, torch
sys.path.insert(0, '.')
import bench
from fedrann_b200.engine import Engine
dev = torch.device('cuda', 0)
if not os.path.exists('/tmp/c2_E.pt'):
    wl = bench.Workload('c2', 0, 1, dev); w = bench.WORKLOADS['c2']
    eng = Engine(0); eng.build_table(wl.lib.all_codes(), 15, sampler=(w['seed'], w['fraction']))
    ip, ix = eng.extract(wl.bases, torch.from_numpy(wl.offsets).to(dev))
    E = eng.project(ip, ix, eng.upload_projection(wl.P), wl.lib.nk)
    torch.save(E.cpu(), '/tmp/c2_E.pt')
else:
    eng = Engine(0)
E = torch.load('/tmp/c2_E.pt').to(dev)
En, Eh, norms = eng.normalize(E)
for i in range(2): eng.knn(En, Eh, norms, 50)
torch.cuda.synchronize()
eng.profile(True)
for i in range(5): idx, d = eng.knn(En, Eh, norms, 50)
torch.cuda.synchronize()
kt = eng.kernel_times()
print('klo/khi', os.environ['FR_KNN_KLO'], os.environ['FR_KNN_KHI'], {k: round(v[1] / v[0], 3) for k, v in kt.items()}, 'fallback', eng.knn_fallback_rows)
PY
done
