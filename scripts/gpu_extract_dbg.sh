#!/bin/bash
mkdir -p gpurun_out
for D in 0 1 2 3 4 7; do
FR_EXTRACT_DEBUG=$D python - <<'PY'
import os, torch, numpy as np, sys
sys.path.insert(0, '.')
import bench
from fedrann_b200.engine import Engine
dev = torch.device('cuda', 0)
wl = bench.Workload('c2', 0, 1, dev)
eng = Engine(0)
eng.build_table(wl.lib.all_codes(), 15)
off = torch.from_numpy(wl.offsets).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for i in range(3): eng.extract(wl.bases, off)
eng.profile(True)
for i in range(5):
    flush.zero_(); eng.extract(wl.bases, off)
torch.cuda.synchronize()
kt = eng.kernel_times()
print('dbg', os.environ['FR_EXTRACT_DEBUG'], {k: round(v[1] / v[0], 4) for k, v in kt.items()}, 'maxlen', int(np.diff(wl.offsets).max()))
PY
done
