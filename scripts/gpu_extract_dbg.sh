#!/bin/bash
# timing experiments on the extract kernel: FR_EXTRACT_DEBUG bits (1 no table probes, 2 no filter loads, 4 no sort, 8 generic filter)
mkdir -p gpurun_out
for CFG in 0 1 4; do for FL in 1; do export FR_FLUSH=$FL
FR_EXTRACT_DEBUG=$CFG python - <<'PY'
import os, torch, numpy as np, sys
sys.path.insert(0, '.')
import bench
from fedrann_b200.engine import Engine
dev = torch.device('cuda', 0)
wl = bench.Workload('c2', 0, 1, dev)
w = bench.WORKLOADS['c2']
eng = Engine(0)
eng.build_table(wl.lib.all_codes(), 15, sampler=(w['seed'], w['fraction']))
off = torch.from_numpy(wl.offsets).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for i in range(3): eng.extract(wl.bases, off)
eng.profile(True)
for i in range(5):
    if os.environ['FR_FLUSH'] == '1': flush.zero_()
    ip, ix = eng.extract(wl.bases, off)
torch.cuda.synchronize()
kt = eng.kernel_times()
print('dbg', os.environ['FR_EXTRACT_DEBUG'], 'flush', os.environ['FR_FLUSH'], 'sampled', eng.sampled, round(kt['fr_extract_kernel'][1] / kt['fr_extract_kernel'][0], 4), 'nnz', ix.numel())
PY
done; done
