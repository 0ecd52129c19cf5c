"""Sharded runs give the single-GPU answer bit for bit; needs >= 2 GPUs on the box
(skipped on the one-GPU box of the round-end run; `gpurun --gpus N -- python -m pytest tests/test_multi_gpu.py`).
The world-size-2 gloo test of the host-side logic is tests/test_host_cpu.py."""
import json
import os
import subprocess
import sys
from os.path import abspath, dirname, join

import pytest
import torch

ROOT = dirname(dirname(abspath(__file__)))
pytestmark = pytest.mark.gpu

N_GPUS = torch.cuda.device_count() if torch.cuda.is_available() else 0


def _run(n, exchange, tmp_path, port):
    out = str(tmp_path / ("multi_%d_%s.json" % (n, exchange)))
    env = dict(os.environ, FEDRANN_EXCHANGE=exchange)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", str(port), join(ROOT, "scripts", "multi_check.py"), "--out", out]
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return json.load(open(out))


@pytest.mark.skipif(N_GPUS < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("exchange", ["peer", "nccl"])
def test_two_ranks_equal_one(exchange, tmp_path):
    line = _run(2, exchange, tmp_path, 29611)
    assert line["ok"] and line["world"] == 2 and line["exchange"] == exchange


@pytest.mark.skipif(N_GPUS < 3, reason="needs at least 3 GPUs")
def test_three_ranks_uneven_shards(tmp_path):
    line = _run(3, "peer", tmp_path, 29612)
    assert line["ok"] and len(set(line["rows_per_rank"])) > 1        # shards of different heights


@pytest.mark.skipif(N_GPUS < 8, reason="needs 8 GPUs")
def test_eight_ranks_equal_one(tmp_path):
    line = _run(8, "peer", tmp_path, 29613)
    assert line["ok"] and line["world"] == 8
