"""Stage 4, drop-in for fedrann/nearest_neighbors.py:13-55.

`NNDescent_ava.get_neighbors` keeps the reference's signature.  The reference
builds an approximate pynndescent index; this backend is EXACT: cosine distance
1 - x.y/(|x||y|) in fp32, k = index_n_neighbors neighbours including the row
itself, ascending distance, ties by lower index, a zero-norm row at distance 1
to everything.  The approximate-index parameters (n_trees, leaf_size, n_iters,
diversify_prob, pruning_degree_multiplier, low_memory, n_jobs, seed, verbose)
are accepted and have no effect on an exact search.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch


@dataclass
class _NearestNeighbors:
    def get_neighbors(self, ref: np.ndarray, que: np.ndarray, n_neighbors: int) -> np.ndarray:
        raise NotImplementedError()


MAX_NEIGHBORS = 128          # FR_MAX_K of csrc/knn_exact.cu: list length of the exact scan and of the fallback
TC_MAX_DIM = 512             # query tile resident in shared memory (csrc/knn_tc.cu TC_MAX_KBLKS * 64)


def check_knn_limits(n_neighbors: int, dim: int) -> None:
    """Limits of the exact backend, checked before any work is done (the reference's pynndescent takes
    any n_neighbors; INTEGRATION.md lists the difference).  Dimensions above 512 are answered by the
    fp32 CUDA-core scan (same lists, ~100x slower than the tensor-core path): warned about, not refused."""
    from .custom_logging import logger
    if int(n_neighbors) > MAX_NEIGHBORS:
        raise ValueError(f"--nndescent-n-neighbors / index_n_neighbors = {n_neighbors}: the exact kNN backend "
                         f"returns at most {MAX_NEIGHBORS} neighbours per row")
    if int(n_neighbors) < 1:
        raise ValueError("n_neighbors must be at least 1")
    if int(dim) > TC_MAX_DIM:
        logger.warning("embedding dimension %d > %d: the kNN stage runs the fp32 scan on the CUDA cores instead "
                       "of the tensor-core path (same neighbour lists, far slower)", dim, TC_MAX_DIM)


def _device_knn(eng, E: torch.Tensor, k: int, q0: int = 0, nq: int | None = None, mode: int = 0):
    En, Eh, norms = eng.normalize(E)
    return eng.knn(En, Eh, norms, k, q0=q0, nq=nq, mode=mode)


class NNDescent_ava(_NearestNeighbors):
    def get_neighbors(
        self,
        data,
        metric="cosine",
        *,
        index_n_neighbors: int = 50,
        n_trees: int | None = 300,
        leaf_size: int | None = 200,
        n_iters: int | None = None,
        diversify_prob: float = 1,
        pruning_degree_multiplier: float = 1.5,
        low_memory: bool = True,
        n_jobs: int | None = 64,
        seed: int | None = 683985,
        verbose: bool = True,
    ):
        if metric != "cosine":
            raise ValueError("fedrann_b200 implements the metric the pipeline uses: 'cosine'")
        from .count_kmers import get_engine
        eng = get_engine()
        if hasattr(data, "toarray"):
            data = data.toarray()
        if isinstance(data, torch.Tensor):
            E = data.to(torch.float32).contiguous().to(f"cuda:{eng.device}")
        else:
            E = eng.dev(np.ascontiguousarray(data, dtype=np.float32))
        n = E.shape[0]
        check_knn_limits(index_n_neighbors, E.shape[1])
        k = min(int(index_n_neighbors), n)
        idx, dist = _device_knn(eng, E, k)
        torch.cuda.synchronize()
        return idx.cpu().numpy(), dist.cpu().numpy()


class ExactNearestNeighbors(_NearestNeighbors):
    """The abstract interface of nearest_neighbors.py:14-19: neighbours of `que` rows among
    `ref` rows.  -> indices (len(que), n_neighbors) into ref."""

    def get_neighbors(self, ref: np.ndarray, que: np.ndarray, n_neighbors: int) -> np.ndarray:
        from .count_kmers import get_engine
        eng = get_engine()
        both = np.concatenate([np.asarray(ref, dtype=np.float32), np.asarray(que, dtype=np.float32)])
        # search the stacked matrix and drop hits that fall in the query block
        k = min(n_neighbors + len(que), len(both))
        idx, _ = _device_knn(eng, eng.dev(both), k, q0=len(ref), nq=len(que))
        idx = idx.cpu().numpy()
        out = np.full((len(que), n_neighbors), -1, dtype=np.int32)
        for i, row in enumerate(idx):
            keep = row[row < len(ref)][:n_neighbors]
            out[i, : len(keep)] = keep
        return out
