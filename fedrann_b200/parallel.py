"""Multi-GPU plumbing: one process per GPU, torch.distributed (NCCL on GPUs,
gloo in the CPU tests).  The path shards by reads: each rank
extracts and projects a contiguous range of reads, the normalised embedding
shards are all-gathered once, and every rank searches its own query rows
against the whole database.  There is no other data-path collective.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist


def init_from_env(backend: str | None = None):
    """Initialise torch.distributed from torchrun's environment; returns
    (rank, world, local_rank).  Single-process runs return (0, 1, 0)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend=backend)
    return rank, world, local


def shard_reads(offsets: np.ndarray, world: int) -> np.ndarray:
    """Contiguous read ranges with equal read COUNTS (+-1): returns int64[world+1] read
    boundaries (rank r owns reads [b[r], b[r+1])).  Equal counts balance the dominant stage
    exactly (every rank searches the same number of query rows) and let the embedding shards
    be all-gathered in one collective without padding; bases per rank then differ by
    O(1/sqrt(reads per rank)), which is what the extraction stage sees."""
    R = len(np.asarray(offsets)) - 1
    return np.array([(R * r) // world for r in range(world + 1)], dtype=np.int64)


def _all_to_all_rows(parts, group):
    """parts[r] = int64 [n_r, 2] rows this rank sends to rank r -> list of the blocks received from every
    rank.  NCCL: one size exchange + one all_to_all_single of the data.  gloo (CPU tests) has no
    all-to-all: the blocks travel through all_gather_object there."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if parts[0].is_cuda:
        dev = parts[0].device
        send_n = torch.tensor([p.shape[0] for p in parts], dtype=torch.int64, device=dev)
        recv_n = torch.empty(world, dtype=torch.int64, device=dev)
        dist.all_to_all_single(recv_n, send_n, group=group)
        recv_counts = [int(x) for x in recv_n.cpu()]
        send = torch.cat(parts).contiguous()
        recv = torch.empty((sum(recv_counts), 2), dtype=torch.int64, device=dev)
        dist.all_to_all_single(recv, send, output_split_sizes=recv_counts,
                               input_split_sizes=[p.shape[0] for p in parts], group=group)
        return list(recv.split(recv_counts, dim=0))
    everything = [None] * world
    dist.all_gather_object(everything, [p.numpy() for p in parts], group=group)
    return [torch.from_numpy(everything[r][rank]) for r in range(world)]


def partitioned_kmer_counts(codes: torch.Tensor, counts: torch.Tensor, min_multiplicity: int, group=None):
    """The multi-GPU half of N1 ("hash-partitioned across GPUs with one all-to-all"; replaces the
    single jellyfish table of fedrann/count_kmers.py:76-123).  Every rank has counted the sampled canonical
    k-mers of ITS reads (codes int64, counts int64, any order).  Rank hash(code) % world owns a k-mer: one
    all-to-all sends every (code, count) pair to its owner, the owner sums the partial counts and applies the
    multiplicity cut, and the survivors -- the library, a small fraction of what was counted -- are all-gathered.
    -> (codes int64 ascending, counts int64), identical on every rank."""
    world = dist.get_world_size(group) if group is not None else 1
    if world == 1:
        keep = counts >= min_multiplicity
        c, n = codes[keep], counts[keep]
        order = torch.argsort(c)
        return c[order], n[order]
    from .synthetic_device import splitmix64_t
    owner = (splitmix64_t(codes) & 0x7FFFFFFFFFFFFFFF) % world
    order = torch.argsort(owner, stable=True)
    pairs = torch.stack([codes[order], counts[order]], dim=1)
    sizes = torch.bincount(owner, minlength=world).cpu().tolist()
    got = _all_to_all_rows(list(pairs.split(sizes, dim=0)), group)
    mine = torch.cat(got) if got else pairs[:0]
    u, inv = torch.unique(mine[:, 0], return_inverse=True)
    tot = torch.zeros_like(u).scatter_add_(0, inv, mine[:, 1])
    keep = tot >= min_multiplicity
    u, tot = u[keep], tot[keep]
    # the library: every rank needs all of it
    n_mine = torch.tensor([u.numel()], dtype=torch.int64, device=u.device)
    ns = [torch.zeros_like(n_mine) for _ in range(world)]
    dist.all_gather(ns, n_mine, group=group)
    ns = [int(x) for x in ns]
    nmax = max(max(ns), 1)
    pad = torch.zeros((nmax, 2), dtype=torch.int64, device=u.device)
    pad[: u.numel(), 0], pad[: u.numel(), 1] = u, tot
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    lib = torch.cat([out[r][: ns[r]] for r in range(world)])
    order = torch.argsort(lib[:, 0])
    return lib[order, 0].contiguous(), lib[order, 1].contiguous()


def gather_to_rank0(x: torch.Tensor, group=None):
    """1-D tensors of different lengths -> their concatenation in rank order on rank 0 (None elsewhere);
    tensors travel as tensors (NCCL gather of padded buffers), not as pickled objects."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    n = torch.tensor([x.numel()], dtype=torch.int64, device=x.device)
    ns = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(ns, n, group=group)
    ns = [int(v) for v in ns]
    nmax = max(max(ns), 1)
    pad = torch.zeros(nmax, dtype=x.dtype, device=x.device)
    pad[: x.numel()] = x
    dst = dist.get_global_rank(group, 0) if group is not None and group is not dist.group.WORLD else 0
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == 0 else None
    dist.gather(pad, bufs, dst=dst, group=group)
    if rank != 0:
        return None
    return torch.cat([bufs[r][: ns[r]] for r in range(world)])


class _DeviceArray:
    """CUDA array interface over memory the native library owns (the exchange slab)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class RowExchange:
    """The embedding database of all ranks exchanged through PEER MEMORY (fr_exchange_*, csrc/exchange.cu):
    every rank owns a slab with room for ALL rows, writes its own rows into it, and pulls the other shards
    over NVLink with the copy engines WHILE the search runs (K6 starts on the local shard and waits for a
    shard's "landed" flag only when it gets there; K7 waits for the fp32 rows).  One instance per
    (engine, row counts, D); reused every step.

        x = RowExchange(eng, rows_per_rank, D, group)
        eng.normalize(E, out=x.local())            # this rank's rows straight into the slab
        x.pull()                                   # barrier + queued peer copies (asynchronous)
        idx, dist = eng.knn(x.En, x.Eh, x.norms, k, q0=x.q0, nq=x.nq)
        x.release()                                # nobody overwrites rows a peer still reads
    """

    def __init__(self, eng, rows_per_rank, D: int, group=None):
        import ctypes as C
        self.eng, self.group = eng, group
        self.world = dist.get_world_size(group) if group is not None else 1
        self.rank = dist.get_rank(group) if group is not None else 0
        counts = [int(c) for c in rows_per_rank]
        assert len(counts) == self.world
        self.bounds = np.zeros(self.world + 1, dtype=np.int64)
        np.cumsum(counts, out=self.bounds[1:])
        self.N, self.D, self.Dp = int(self.bounds[-1]), int(D), (int(D) + 63) // 64 * 64
        self.q0, self.nq = int(self.bounds[self.rank]), counts[self.rank]
        L = eng.L
        h = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        from ._lib import check
        check(eng.h, L.fr_exchange_create(eng.h, self.rank, self.world, self.bounds.ctypes.data_as(C.c_void_p), self.D,
                                          self.Dp, C.byref(h), handle), "fr_exchange_create")
        self.h = h
        if self.world > 1:
            mine = torch.tensor(list(bytes(handle)), dtype=torch.uint8, device=f"cuda:{eng.device}")
            allh = torch.empty(64 * self.world, dtype=torch.uint8, device=mine.device)
            dist.all_gather_into_tensor(allh, mine, group=group)
            blob = bytes(allh.cpu().numpy().tobytes())
            check(eng.h, L.fr_exchange_connect(self.h, C.c_char_p(blob)), "fr_exchange_connect")
            dist.barrier(group)
        dev = f"cuda:{eng.device}"
        n = max(self.N, 1)
        self.En = torch.as_tensor(_DeviceArray(L.fr_exchange_rows_f32(self.h), (n, self.D), "<f4"), device=dev)[: self.N]
        self.Eh = torch.as_tensor(_DeviceArray(L.fr_exchange_rows_f16(self.h), (n, self.Dp), "<f2"), device=dev)[: self.N]
        self.norms = torch.as_tensor(_DeviceArray(L.fr_exchange_norms(self.h), (n,), "<f4"), device=dev)[: self.N]

    def local(self):
        a, b = self.q0, self.q0 + self.nq
        return self.En[a:b], self.Eh[a:b], self.norms[a:b]

    def pull(self):
        from ._lib import check
        import ctypes as C
        check(self.eng.h, self.eng.L.fr_exchange_pull(self.h, C.c_void_p(torch.cuda.current_stream().cuda_stream)),
              "fr_exchange_pull")

    def release(self):
        from ._lib import check
        import ctypes as C
        check(self.eng.h, self.eng.L.fr_exchange_release(self.h, C.c_void_p(torch.cuda.current_stream().cuda_stream)),
              "fr_exchange_release")

    @property
    def bytes_pulled_per_step(self) -> int:
        remote = self.N - self.nq
        return remote * (self.D * 4 + self.Dp * 2 + 4)

    def close(self):
        if getattr(self, "h", None):
            torch.cuda.synchronize()
            if self.world > 1:
                dist.barrier(self.group)         # no peer is still reading this slab
            self.En = self.Eh = self.norms = None
            self.eng.L.fr_exchange_destroy(self.h)
            self.h = None


def row_exchange(eng, rows_per_rank, D: int, group=None) -> RowExchange:
    """The engine's exchange for this geometry (created on first use, every rank at the same time)."""
    key = (tuple(int(c) for c in rows_per_rank), int(D))
    cur = getattr(eng, "_row_exchange", None)
    if cur is not None and cur[0] == key:
        return cur[1]
    if cur is not None:
        cur[1].close()
    x = RowExchange(eng, rows_per_rank, D, group)
    eng._row_exchange = (key, x)
    return x


def all_gather_rows(En: torch.Tensor, Eh: torch.Tensor, norms: torch.Tensor, group=None, counts=None,
                    overlap: bool = False):
    """All-gather row shards of unequal height.  Returns (En_all, Eh_all,
    norms_all, q0) where q0 is the global row number of this rank's first row
    (rank-major order).  `counts` = rows per rank when the caller already knows
    them (the read sharding is computed identically on every rank), which saves
    the size exchange and its host synchronisation.  Every rank receives each
    shard directly into its final place (no padding, no compaction copy).
    With overlap=True a fifth value is returned, see below."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if counts is None:
        n_local = torch.tensor([En.shape[0]], dtype=torch.int64, device=En.device)
        cnt = torch.zeros(world, dtype=torch.int64, device=En.device)
        dist.all_gather_into_tensor(cnt, n_local, group=group)
        counts = [int(c) for c in cnt.cpu()]
    counts = [int(c) for c in counts]
    assert counts[rank] == En.shape[0]
    total = sum(counts)
    equal = all(c == counts[0] for c in counts)

    def gather(x):
        out = torch.empty((total,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        if equal:
            dist.all_gather_into_tensor(out, x.contiguous(), group=group)
        else:
            # shards of different heights (reads not divisible by ranks): pad to the tallest, one
            # collective, then copy into place
            nmax = max(counts)
            pad = torch.zeros((nmax,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
            pad[: x.shape[0]] = x
            tmp = torch.empty((world * nmax,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
            if x.is_cuda:
                dist.all_gather_into_tensor(tmp, pad, group=group)
            else:
                dist.all_gather(list(tmp.split(nmax, dim=0)), pad, group=group)   # gloo (CPU tests)
            for r, dst in enumerate(out.split(counts, dim=0)):
                dst.copy_(tmp[r * nmax: r * nmax + counts[r]])
        return out

    q0 = sum(counts[:rank])
    if not (overlap and En.is_cuda):
        return gather(En), gather(Eh), gather(norms), q0
    # The tensor-core pass needs only the fp16 copy and the norms: those are gathered first, on the
    # caller's stream.  The fp32 rows (re-rank only) follow on a side stream; the returned event
    # fires when they are complete (Engine.knn(rows_ready=...) waits for it just before the re-rank).
    Eh_all, norms_all = gather(Eh), gather(norms)
    cur = torch.cuda.current_stream()
    side = _side_stream(En.device)
    side.wait_stream(cur)
    with torch.cuda.stream(side):
        En_all = gather(En)
        ev = torch.cuda.Event()
        ev.record(side)
    En.record_stream(side)
    En_all.record_stream(cur)
    return En_all, Eh_all, norms_all, q0, ev


_side_streams: dict = {}


def _side_stream(dev):
    s = _side_streams.get(dev.index)
    if s is None:
        s = _side_streams[dev.index] = torch.cuda.Stream(device=dev)
    return s
