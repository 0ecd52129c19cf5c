"""The hot path end to end on one rank: extract -> project -> normalise ->
(all-gather) -> kNN.  Mirrors stages 1b, 3 and 4 of run_fedrann_pipeline
(fedrann/__main__.py:321-365) with device-resident intermediates.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field

import torch

from .engine import Engine


# How the ranks' embedding rows reach each other:
#   "peer" (default)  parallel.RowExchange: peer-memory pulls over NVLink by the copy engines, consumed by K6 shard
#                     by shard while it runs (csrc/exchange.cu) -- no collective, no SM taken from the search;
#   "nccl"            three all_gather_into_tensor calls before the search (round 1; 9.9 ms of a 74 ms step at the
#                     chr1 scale on 8 ranks), kept for A/B measurements and for boxes without peer access.
EXCHANGE = os.environ.get("FEDRANN_EXCHANGE", "peer")
OVERLAP_ROW_GATHER = os.environ.get("FEDRANN_OVERLAP_GATHER", "0") == "1"      # nccl mode: fp32 rows on a side stream


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
    db: tuple | None = None          # (En, Eh, norms) of the WHOLE database the search ran against


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
    multi = group is not None and torch.distributed.get_world_size(group) > 1
    xch = None
    if multi and EXCHANGE == "peer":
        from .parallel import row_exchange
        xch = row_exchange(eng, _rows_per_rank(E.shape[0], group, rows_per_rank), E.shape[1], group)
    t.start("normalize")
    En, Eh, norms = eng.normalize(E, out=xch.local() if xch is not None else None)
    t.stop()
    q0, nq, ready = 0, En.shape[0], []
    if xch is not None:
        t.start("exchange")          # the barrier + queueing; the transfers themselves run under the search
        xch.pull()
        En, Eh, norms, q0 = xch.En, xch.Eh, xch.norms, xch.q0
        t.stop()
    elif multi:
        t.start("allgather")
        from .parallel import all_gather_rows
        En, Eh, norms, q0, *ready = all_gather_rows(En, Eh, norms, group, rows_per_rank, overlap=OVERLAP_ROW_GATHER)
        t.stop()
    t.start("knn")
    k_eff = min(k_nn, En.shape[0])
    idx, dist = eng.knn(En, Eh, norms, k_eff, q0=q0, nq=nq, mode=mode, rows_ready=ready[0] if ready else None)
    if xch is not None:
        xch.release()
    t.stop()
    return HotPathResult(indptr, indices, E, idx, dist, q0, (En, Eh, norms))


def _rows_per_rank(n_local: int, group, rows_per_rank):
    """Rows of every rank (rank-major); exchanged once when the caller does not know them."""
    if rows_per_rank is not None:
        return [int(c) for c in rows_per_rank]
    world = torch.distributed.get_world_size(group)
    out = [None] * world
    torch.distributed.all_gather_object(out, int(n_local), group=group)
    return [int(c) for c in out]


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


# planning constants of the host path's schedule (orders of magnitude are what matters): pinned host -> device copy
# rate and the distance kernel's rate on this machine
_H2D_BYTES_PER_S = 50e9
_K6_FLOPS_PER_S = 1.0e15


def phase_schedule(n_bases: int, n_rows: int, Dp: int):
    """Where the host path cuts the reads: (chunk cuts, phase ends) as cumulative fractions of the bases.

    The copy of the reads is LINEAR in the fraction f that has arrived (C * f), the distance work that can
    be done with it QUADRATIC (W * f^2), so early on the GPU can only wait for PCIe and at the end the copy
    is long over: whatever the schedule, the step cannot end before max_f (C f + W (1 - f^2)), reached at
    f* = C / 2W.  The first search phase is therefore placed at f* (a small read set like the E. coli
    configuration is copy-bound, f* ~ 0.8: few late phases; the chr1 scale is compute-bound, f* ~ 0.15:
    start early), and every later phase where the copy will have got to by the time the previous phase's
    work is done, at least 15 % further (every phase costs a pair of launches and a candidate buffer per
    row).  Reads between phase ends are still copied / extracted / projected in pieces of <= 15 %."""
    C = n_bases / _H2D_BYTES_PER_S
    W = 2.0 * float(n_rows) * float(n_rows) * Dp / _K6_FLOPS_PER_S
    if W <= 0 or C <= 0:
        return [], [1.0]
    p = min(0.6, max(0.1, C / (2.0 * W)))
    ends, t_free, prev = [], 0.0, 0.0
    while p < 0.999 and len(ends) < 6:
        ends.append(p)
        t_free = max(t_free, C * p) + W * (p * p - prev * prev)
        prev = p
        p = max(p + 0.15, min(1.0, t_free / C))
        if p > 0.9:
            break
    ends.append(1.0)
    cuts, last = [], 0.0
    for e in ends:
        n_extra = int((e - last - 1e-9) / 0.15)
        for i in range(1, n_extra + 1):
            cuts.append(last + (e - last) * i / (n_extra + 1))
        if e < 1.0:
            cuts.append(e)
        last = e
    return cuts, ends


def run_hot_path_host(eng: Engine, hb: HostBuffers, P_dev, nk: int, k_nn: int, *, mode: int = 0, group=None,
                      timers: StageTimes | None = None, chunks=None, rows_per_rank=None,
                      knn_overlap: bool = True, knn_phase_ends=None, async_extract: bool = True,
                      direct_out: bool = True, trace: list | None = None) -> HotPathResult:
    """The call a host program makes: reads in pinned host memory -> device ->
    extract/project/kNN -> neighbour lists back in pinned host memory.  Returns
    after the results have landed.

    The reads cross PCIe in `chunks` pieces (an int: equal pieces; a sequence: the
    cumulative fractions of the bases at which to cut) on a copy stream while the compute
    stream extracts, projects and normalises the pieces that have arrived (reads
    are independent).  On one GPU the kNN is incremental as well (knn_overlap,
    fr_knn_inc_*): after chunk c the new embedding rows are searched against all
    rows present and the earlier rows against the new ones, so only the last
    chunk's share of the distance work is left when the copy ends.
    chunks / knn_phase_ends: where to cut the reads and after which chunk counts a search phase runs.
    Default (both None): phase_schedule() above, from the size of the read set and of the distance work
    (E. coli configuration: cuts at 15/30/45/60/.../89 %, phases after 60 / 78 / 100 % -- every launch of
    the distance kernel costs ~0.1 ms of ramp and tail and the rows of a new chunk start their thresholds
    from -inf on all CTAs at once, so early, small phases cost more than they hide there
    (scripts/e2e_tune.py: 9.3 ms without overlap, 8.8 ms with phases after 60 / 80 / 100 %, 10.8 ms with a
    phase after each of 8 equal chunks); chr1 scale: phases after ~15 / 30 / 52 / 100 %, which keeps the
    GPU busy from the moment a sixth of the reads has arrived).
    async_extract: chunks go through fr_extract_async (room for one hit per 8 bases, no host
    synchronisation per chunk: the host queues the whole step and the GPU is only ever waiting for
    PCIe); a chunk that does not fit is reported at the end and the step is redone synchronously.
    direct_out: the re-rank kernel writes the neighbour lists straight into the pinned host buffers
    (posted PCIe writes while it runs) instead of a device-to-host copy after it; the result's
    nbr_idx / nbr_dist are then those host tensors."""
    dev = torch.device("cuda", eng.device)
    n = hb.bases.numel()
    R = hb.offsets.numel() - 1
    cur = torch.cuda.current_stream()

    def mark(label, stream=None):           # trace: (label, event) pairs for a timeline of the step
        if trace is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(stream or cur)
            trace.append((label, ev))
    mark("start")
    copy_stream = _copy_stream(dev)
    bases_d = torch.empty(n + 16, dtype=torch.uint8, device=dev)[:n]
    offsets_d = hb.offsets.to(dev, non_blocking=True)
    # read ranges by bases: `chunks` equal pieces, or the cumulative fractions given
    chunks_arg = chunks
    auto_ends = None
    if chunks is None:
        chunks, auto_ends = phase_schedule(n, 2 * R, (P_dev[3] + 63) // 64 * 64)
    fracs = [c / chunks for c in range(1, chunks)] if isinstance(chunks, int) else [float(f) for f in chunks if 0 < f < 1]
    fracs = fracs[: max(0, R - 1)]
    off = hb.offsets
    bounds = [0]
    for f in fracs:
        target = int(off[0]) + int(n * f)
        bounds.append(max(bounds[-1], int(torch.searchsorted(off, torch.tensor(target)).item())))
    bounds.append(R)
    chunks = len(bounds) - 1
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
            mark("h2d %d" % c, copy_stream)
    D = P_dev[3]
    E = torch.empty((2 * R, D), dtype=torch.float32, device=dev)
    multi = group is not None and torch.distributed.get_world_size(group) > 1
    k_eff = min(k_nn, 2 * R)
    inc = knn_overlap and mode == 0 and not multi and chunks > 1 and Engine.knn_inc_supported(2 * R, D, k_eff)
    if inc:
        if auto_ends is not None and knn_phase_ends is None and len(fracs) == chunks - 1:
            # phase ends as chunk counts: the chunk whose cut is the phase's fraction
            default_ends = [1 + min(range(len(fracs)), key=lambda i: abs(fracs[i] - e)) if e < 1.0 and fracs else chunks
                            for e in auto_ends]
        else:
            default_ends = range(max(1, chunks - 2), chunks + 1)
        ends = sorted(set(int(e) for e in (knn_phase_ends or default_ends) if 1 <= int(e) < chunks) | {chunks})
        ends = ends[-64:]                       # fr_knn_inc_begin takes at most 64 phases
        Dp = (D + 63) // 64 * 64
        En = torch.empty((2 * R, D), dtype=torch.float32, device=dev)
        Eh = torch.empty((2 * R, Dp), dtype=torch.float16, device=dev)
        norms = torch.empty(2 * R, dtype=torch.float32, device=dev)
        eng.knn_inc_begin(En, Eh, norms, k_eff, [0] + [2 * bounds[e] for e in ends])
        phase = 0
    ips, ixs, base = [], [], 0
    status = torch.zeros(2 * chunks, dtype=torch.int64, device=dev) if async_extract else None
    for c in range(chunks):
        r0, r1 = bounds[c], bounds[c + 1]
        cur.wait_event(events[c])
        if r1 > r0:
            if async_extract:
                room = (int(off[r1]) - int(off[r0])) // 8 + (r1 - r0) + 1024
                ix = torch.empty(room, dtype=torch.int32, device=dev)
                ip = eng.extract_async(bases_d, offsets_d[r0:r1 + 1], ix, status[2 * c:2 * c + 2])
                ips.append(ip)
                ixs.append(ix)
            else:
                ip, ix = eng.extract(bases_d, offsets_d[r0:r1 + 1])
                ips.append(ip[1:] + base if ips else ip)
                ixs.append(ix)
                base += int(ix.numel())
            mark("extract %d" % c)
            eng.project(ip, ix, P_dev, nk, out=E[2 * r0:2 * r1])
            mark("project %d" % c)
            if inc:
                eng.normalize(E[2 * r0:2 * r1], out=(En[2 * r0:2 * r1], Eh[2 * r0:2 * r1], norms[2 * r0:2 * r1]))
        if inc and c + 1 == ends[phase]:
            eng.knn_inc_phase(phase)
            mark("knn phase %d" % phase)
            phase += 1
    bases_d.record_stream(copy_stream)
    if not async_extract:
        indptr = torch.cat(ips) if ips else torch.zeros(1, dtype=torch.int64, device=dev)
        indices = torch.cat(ixs) if ixs else torch.zeros(0, dtype=torch.int32, device=dev)
    direct = direct_out and hb.nbr_idx.shape[1] == k_eff and hb.nbr_idx.is_pinned() and hb.nbr_dist.is_pinned()
    out = (hb.nbr_idx, hb.nbr_dist) if direct else None
    if inc:
        idx, dist = eng.knn_inc_finish(out=out)
        q0 = 0
    else:
        xch = None
        if multi and EXCHANGE == "peer":
            from .parallel import row_exchange
            xch = row_exchange(eng, _rows_per_rank(E.shape[0], group, rows_per_rank), E.shape[1], group)
        En, Eh, norms = eng.normalize(E, out=xch.local() if xch is not None else None)
        q0, nq, ready = 0, En.shape[0], []
        if xch is not None:
            xch.pull()
            En, Eh, norms, q0 = xch.En, xch.Eh, xch.norms, xch.q0
        elif multi:
            from .parallel import all_gather_rows
            En, Eh, norms, q0, *ready = all_gather_rows(En, Eh, norms, group, rows_per_rank, overlap=OVERLAP_ROW_GATHER)
        if direct and (En.shape[0] < k_nn or nq != hb.nbr_idx.shape[0]):
            direct, out = False, None
        k_eff = min(k_nn, En.shape[0])
        idx, dist = eng.knn(En, Eh, norms, k_eff, q0=q0, nq=nq, mode=mode, rows_ready=ready[0] if ready else None,
                            out=out)
        if xch is not None:
            xch.release()
    mark("knn done")
    if not direct:
        hb.nbr_idx[:, : idx.shape[1]].copy_(idx, non_blocking=True)
        hb.nbr_dist[:, : dist.shape[1]].copy_(dist, non_blocking=True)
        mark("d2h")
    if async_extract:
        st_h = _pinned_i64(dev, 2 * chunks)
        st_h.copy_(status, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        st = st_h.tolist()
        if any(st[0::2]):           # a chunk had more hits than room: redo the step the synchronous way
            if inc:
                eng.knn_fallback_rows   # (drains the engine's pending counter read)
            return run_hot_path_host(eng, hb, P_dev, nk, k_nn, mode=mode, group=group, timers=timers, chunks=chunks_arg,
                                     rows_per_rank=rows_per_rank, knn_overlap=knn_overlap,
                                     knn_phase_ends=knn_phase_ends, async_extract=False, direct_out=direct_out)
        # CSR of the whole read set from the chunks' rows (queued; the neighbour lists have landed)
        nz = [st[2 * c + 1] for c in range(chunks) if bounds[c + 1] > bounds[c]]
        parts, base = [], 0
        for ip, n_c in zip(ips, nz):
            parts.append(ip[1:] + base if parts else ip)
            base += n_c
        indptr = torch.cat(parts) if parts else torch.zeros(1, dtype=torch.int64, device=dev)
        indices = torch.cat([ix[:n_c] for ix, n_c in zip(ixs, nz)]) if ixs else torch.zeros(0, dtype=torch.int32, device=dev)
    else:
        torch.cuda.current_stream().synchronize()
    return HotPathResult(indptr, indices, E, idx, dist, q0, (En, Eh, norms))


_pinned: dict = {}


def _pinned_i64(dev, n: int) -> torch.Tensor:
    t = _pinned.get((dev.index, n))
    if t is None:
        t = _pinned[(dev.index, n)] = torch.empty(n, dtype=torch.int64, pin_memory=True)
    return t


_copy_streams: dict = {}


def _copy_stream(dev):
    s = _copy_streams.get(dev.index)
    if s is None:
        s = _copy_streams[dev.index] = torch.cuda.Stream(device=dev)
    return s
