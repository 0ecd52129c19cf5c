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
    """Contiguous read ranges balanced by bases: returns int64[world+1] read
    boundaries (rank r owns reads [b[r], b[r+1]))."""
    offsets = np.asarray(offsets, dtype=np.int64)
    R = len(offsets) - 1
    total = int(offsets[-1] - offsets[0])
    bounds = np.zeros(world + 1, dtype=np.int64)
    for r in range(1, world):
        target = offsets[0] + (total * r) // world
        bounds[r] = int(np.searchsorted(offsets, target, side="left"))
    bounds[world] = R
    bounds = np.minimum(np.maximum.accumulate(bounds), R)
    return bounds


def all_gather_rows(En: torch.Tensor, Eh: torch.Tensor, norms: torch.Tensor, group=None):
    """All-gather row shards of unequal height.  Returns (En_all, Eh_all,
    norms_all, q0) where q0 is the global row number of this rank's first row
    (rank-major order)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    n_local = torch.tensor([En.shape[0]], dtype=torch.int64, device=En.device)
    counts = torch.zeros(world, dtype=torch.int64, device=En.device)
    dist.all_gather_into_tensor(counts, n_local, group=group)
    counts = [int(c) for c in counts.cpu()]
    nmax = max(counts)

    def gather(x):
        pad = x
        if x.shape[0] < nmax:
            pad = torch.zeros((nmax,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
            pad[: x.shape[0]] = x
        out = torch.empty((world * nmax,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, pad.contiguous(), group=group)
        if all(c == nmax for c in counts):
            return out
        return torch.cat([out[r * nmax: r * nmax + counts[r]] for r in range(world)], dim=0)

    q0 = sum(counts[:rank])
    return gather(En), gather(Eh), gather(norms), q0
