"""The hot path end to end on one rank: extract -> project -> normalise ->
(all-gather) -> kNN.  Mirrors stages 1b, 3 and 4 of run_fedrann_pipeline
(fedrann/__main__.py:321-365) with device-resident intermediates.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch

from .engine import Engine


# Gather the fp32 rows on a side stream while K6 runs (parallel.all_gather_rows(overlap=True)).
# Off by default: measured on 4 x B200 the NCCL kernel and the persistent K6 CTAs compete for SMs
# (K6 +8 %) and the step gets slower; at 2 ranks it gains 7 %.
OVERLAP_ROW_GATHER = False


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
                 group=None, rows_per_rank=None) -> HotPathResult:
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
    q0, nq, ready = 0, En.shape[0], []
    if group is not None and torch.distributed.get_world_size(group) > 1:
        t.start("allgather")
        from .parallel import all_gather_rows
        En, Eh, norms, q0, *ready = all_gather_rows(En, Eh, norms, group, rows_per_rank, overlap=OVERLAP_ROW_GATHER)
        t.stop()
    t.start("knn")
    k_eff = min(k_nn, En.shape[0])
    idx, dist = eng.knn(En, Eh, norms, k_eff, q0=q0, nq=nq, mode=mode, rows_ready=ready[0] if ready else None)
    t.stop()
    return HotPathResult(indptr, indices, E, idx, dist, q0)


@dataclass
class HostBuffers:
    """Pinned host memory either side of the path: the reads going in and the
    neighbour lists coming out (what the CLI hands to / gets from the engine)."""
    bases: torch.Tensor        # uint8 pinned
    offsets: torch.Tensor      # int64 pinned [R+1]
    nbr_idx: torch.Tensor      # int32 pinned (2R, k)
    nbr_dist: torch.Tensor     # fp32 pinned (2R, k)

    @classmethod
    def from_device(cls, bases_d: torch.Tensor, offsets_np, n_rows: int, k: int) -> "HostBuffers":
        b = torch.empty(bases_d.numel(), dtype=torch.uint8, pin_memory=True)
        b.copy_(bases_d)
        o = torch.empty(len(offsets_np), dtype=torch.int64, pin_memory=True)
        o.copy_(torch.from_numpy(offsets_np))
        return cls(b, o, torch.empty((n_rows, k), dtype=torch.int32, pin_memory=True),
                   torch.empty((n_rows, k), dtype=torch.float32, pin_memory=True))

    @property
    def h2d_bytes(self) -> int:
        return self.bases.numel() + 8 * self.offsets.numel()

    @property
    def d2h_bytes(self) -> int:
        return 4 * self.nbr_idx.numel() + 4 * self.nbr_dist.numel()


def run_hot_path_host(eng: Engine, hb: HostBuffers, P_dev, nk: int, k_nn: int, *, mode: int = 0, group=None,
                      timers: StageTimes | None = None, chunks: int = 8, rows_per_rank=None) -> HotPathResult:
    """The call a host program makes: reads in pinned host memory -> device ->
    extract/project/kNN -> neighbour lists back in pinned host memory.  Returns
    after the results have landed.

    The reads cross PCIe in `chunks` pieces on a copy stream while the compute
    stream extracts and projects the pieces that have arrived (reads are
    independent); the kNN starts when the last piece is projected."""
    dev = torch.device("cuda", eng.device)
    n = hb.bases.numel()
    R = hb.offsets.numel() - 1
    cur = torch.cuda.current_stream()
    copy_stream = _copy_stream(dev)
    bases_d = torch.empty(n + 16, dtype=torch.uint8, device=dev)[:n]
    offsets_d = hb.offsets.to(dev, non_blocking=True)
    chunks = max(1, min(chunks, R))
    # read ranges balanced by bases
    off = hb.offsets
    bounds = [0]
    for c in range(1, chunks):
        target = int(off[0]) + (n * c) // chunks
        bounds.append(max(bounds[-1], int(torch.searchsorted(off, torch.tensor(target)).item())))
    bounds.append(R)
    copy_stream.wait_stream(cur)
    events = []
    with torch.cuda.stream(copy_stream):
        for c in range(chunks):
            b0, b1 = int(off[bounds[c]]), int(off[bounds[c + 1]])
            if b1 > b0:
                bases_d[b0:b1].copy_(hb.bases[b0:b1], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
            events.append(ev)
    D = P_dev[3]
    E = torch.empty((2 * R, D), dtype=torch.float32, device=dev)
    ips, ixs, base = [], [], 0
    for c in range(chunks):
        r0, r1 = bounds[c], bounds[c + 1]
        cur.wait_event(events[c])
        if r1 == r0:
            continue
        ip, ix = eng.extract(bases_d, offsets_d[r0:r1 + 1])
        eng.project(ip, ix, P_dev, nk, out=E[2 * r0:2 * r1])
        ips.append(ip[1:] + base if ips else ip)
        ixs.append(ix)
        base += int(ix.numel())
    bases_d.record_stream(copy_stream)
    indptr = torch.cat(ips) if ips else torch.zeros(1, dtype=torch.int64, device=dev)
    indices = torch.cat(ixs) if ixs else torch.zeros(0, dtype=torch.int32, device=dev)
    En, Eh, norms = eng.normalize(E)
    q0, nq, ready = 0, En.shape[0], []
    if group is not None and torch.distributed.get_world_size(group) > 1:
        from .parallel import all_gather_rows
        En, Eh, norms, q0, *ready = all_gather_rows(En, Eh, norms, group, rows_per_rank, overlap=OVERLAP_ROW_GATHER)
    k_eff = min(k_nn, En.shape[0])
    idx, dist = eng.knn(En, Eh, norms, k_eff, q0=q0, nq=nq, mode=mode, rows_ready=ready[0] if ready else None)
    res = HotPathResult(indptr, indices, E, idx, dist, q0)
    hb.nbr_idx[:, : res.nbr_idx.shape[1]].copy_(res.nbr_idx, non_blocking=True)
    hb.nbr_dist[:, : res.nbr_dist.shape[1]].copy_(res.nbr_dist, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return res


_copy_streams: dict = {}


def _copy_stream(dev):
    s = _copy_streams.get(dev.index)
    if s is None:
        s = _copy_streams[dev.index] = torch.cuda.Stream(device=dev)
    return s
