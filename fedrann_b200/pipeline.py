"""The hot path end to end on one rank: extract -> project -> normalise ->
(all-gather) -> kNN.  Mirrors stages 1b, 3 and 4 of run_fedrann_pipeline
(fedrann/__main__.py:321-365) with device-resident intermediates.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch

from .engine import Engine


@dataclass
class StageTimes:
    """CUDA-event timers per stage (ms), accumulated over calls."""
    enabled: bool = False
    ms: dict = field(default_factory=dict)
    _open: list = field(default_factory=list)

    def start(self, name):
        if not self.enabled:
            return
        e0 = torch.cuda.Event(enable_timing=True)
        e0.record()
        self._open.append((name, e0))

    def stop(self):
        if not self.enabled:
            return
        name, e0 = self._open.pop()
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record()
        self.ms.setdefault(name, []).append((e0, e1))

    def totals(self):
        torch.cuda.synchronize()
        return {k: sum(a.elapsed_time(b) for a, b in v) for k, v in self.ms.items()}


@dataclass
class HotPathResult:
    indptr: torch.Tensor
    indices: torch.Tensor
    embedding: torch.Tensor          # (2R, D) fp32, this rank's rows
    nbr_idx: torch.Tensor            # (2R, k) int32, global row ids
    nbr_dist: torch.Tensor           # (2R, k) fp32
    row_offset: int                  # global row id of this rank's first embedding row


def run_hot_path(eng: Engine, bases: torch.Tensor, offsets: torch.Tensor, P_dev, nk: int, k_nn: int,
                 *, mode: int = 0, timers: StageTimes | None = None, docfreq: torch.Tensor | None = None,
                 group=None) -> HotPathResult:
    """bases/offsets: this rank's reads on the device.  With a process group the
    normalised embeddings of all ranks are all-gathered (NCCL) and this rank
    searches its own rows against the whole database; neighbour ids are global
    (rank-major) row numbers."""
    t = timers or StageTimes(False)
    t.start("extract")
    indptr, indices = eng.extract(bases, offsets, docfreq)
    t.stop()
    t.start("project")
    E = eng.project(indptr, indices, P_dev, nk)
    t.stop()
    t.start("normalize")
    En, Eh, norms = eng.normalize(E)
    t.stop()
    q0, nq = 0, En.shape[0]
    if group is not None and torch.distributed.get_world_size(group) > 1:
        t.start("allgather")
        from .parallel import all_gather_rows
        En, Eh, norms, q0 = all_gather_rows(En, Eh, norms, group)
        t.stop()
    t.start("knn")
    k_eff = min(k_nn, En.shape[0])
    idx, dist = eng.knn(En, Eh, norms, k_eff, q0=q0, nq=nq, mode=mode)
    t.stop()
    return HotPathResult(indptr, indices, E, idx, dist, q0)
