"""Host-side logic that runs without a GPU: the projection operand, read
sharding, and the world_size-2 gloo path of the all-gather."""
import os
import sys
from os.path import abspath, dirname, join

import numpy as np
import pytest

from oracle import oracle as O
from helpers import GOLDEN

ROOT = dirname(dirname(abspath(__file__)))


def test_precompute_matches_oracle_and_reference_structure():
    from fedrann_b200.precompute import precompute_matrix
    z = np.load(join(GOLDEN, "project_s0.npz"))
    A = precompute_matrix(int(z["D"]), z["counts"], int(z["n_features"]))
    B = O.precompute_matrix(int(z["D"]), z["counts"], int(z["n_features"]))
    assert np.array_equal(A.indptr, B.indptr) and np.array_equal(A.indices, B.indices)
    assert np.array_equal(A.data.view(np.uint32), B.data.view(np.uint32))
    assert np.array_equal(A.indptr.astype(np.int64), z["P_ref_indptr"])
    assert np.array_equal(A.indices.astype(np.int32), z["P_ref_indices"])


def test_get_precompute_matrix_reads_counter_file(tmp_path):
    from fedrann_b200.precompute import get_precompute_matrix, kmer_count_generator
    p = tmp_path / "fwd_kmer_library.fasta"
    p.write_text(">3\nACGTA\n>7\nCCGTA\n>2\nTTGTA\n")
    assert list(kmer_count_generator(str(p), 3)) == [(0, 3), (3, 3), (1, 7), (4, 7), (2, 2), (5, 2)]
    P, nf = get_precompute_matrix(8, str(p), 6)
    assert nf == 6 and P.shape == (6, 8) and P.dtype == np.float32


def test_shard_reads_balanced_and_contiguous():
    from fedrann_b200.parallel import shard_reads
    rng = np.random.default_rng(0)
    lens = rng.integers(1, 20000, 1000)
    offs = np.concatenate([[0], np.cumsum(lens)])
    for w in (1, 2, 3, 4, 8):
        b = shard_reads(offs, w)
        assert b[0] == 0 and b[-1] == 1000 and np.all(np.diff(b) >= 0)
        assert np.diff(b).max() - np.diff(b).min() <= 1                  # equal read counts (+-1)
        sizes = [offs[b[r + 1]] - offs[b[r]] for r in range(w)]
        assert max(sizes) < 1.3 * (offs[-1] / w)                         # bases follow within noise
    assert list(shard_reads(np.array([0]), 4)) == [0, 0, 0, 0, 0]


def _gloo_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from fedrann_b200.parallel import init_from_env, all_gather_rows
    init_from_env("gloo")
    n = 5 + 3 * rank
    En = torch.full((n, 4), float(rank)) + torch.arange(n)[:, None]
    Eh = En.to(torch.float16)
    norms = torch.arange(n, dtype=torch.float32) + 100 * rank
    A, B, C, q0 = all_gather_rows(En, Eh, norms)
    q.put((rank, A.tolist(), C.tolist(), q0))
    dist.destroy_process_group()


def test_all_gather_rows_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    ps = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=120) for _ in range(2))
    [p.join(60) for p in ps]
    (r0, A0, C0, q00), (r1, A1, C1, q01) = res
    assert A0 == A1 and C0 == C1 and len(A0) == 13
    assert q00 == 0 and q01 == 5
    assert C0[:5] == [0, 1, 2, 3, 4] and C0[5:] == [100 + i for i in range(8)]


def _gloo_count_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from fedrann_b200.parallel import init_from_env, partitioned_kmer_counts, gather_to_rank0
    init_from_env("gloo")
    rng = np.random.default_rng(100 + rank)
    # every rank counted some k-mers of its own reads; many k-mers are seen by both ranks
    codes = np.unique(rng.integers(0, 4000, 1500 + 200 * rank)).astype(np.int64)
    counts = rng.integers(1, 4, len(codes)).astype(np.int64)
    c, n = partitioned_kmer_counts(torch.from_numpy(codes), torch.from_numpy(counts), 3, dist.group.WORLD)
    g = gather_to_rank0(torch.arange(3 + 2 * rank, dtype=torch.int32) + 10 * rank, dist.group.WORLD)
    q.put((rank, codes.tolist(), counts.tolist(), c.tolist(), n.tolist(), None if g is None else g.tolist()))
    dist.destroy_process_group()


def test_partitioned_kmer_counts_world2_gloo():
    """N1 across ranks: hash-partitioned owners, one all-to-all, multiplicity cut at the owner, library
    all-gathered -- same (codes, counts) on every rank as summing everything in one place."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    ps = [ctx.Process(target=_gloo_count_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=180) for _ in range(2))
    [p.join(60) for p in ps]
    total = {}
    for _, codes, counts, _, _, _ in res:
        for c, n in zip(codes, counts):
            total[c] = total.get(c, 0) + n
    want = sorted((c, n) for c, n in total.items() if n >= 3)
    for _, _, _, c, n, _ in res:
        assert list(zip(c, n)) == want
    assert res[0][5] == [0, 1, 2, 10, 11, 12, 13, 14] and res[1][5] is None


def test_phase_schedule_of_the_host_path():
    """pipeline.phase_schedule: first search phase at f* = C / 2W (copy-bound read sets start late, compute-bound
    ones early), later phases where the copy will have got to, pieces of at most 15 % in between."""
    from fedrann_b200.pipeline import phase_schedule
    for n_bases, n_rows, Dp in [(249e6, 50_000, 512), (7.4e9, 992_000, 256), (60e6, 12_000, 512), (1e6, 400, 64),
                                (45e9, 6_000_000, 256), (1, 2, 64)]:
        cuts, ends = phase_schedule(n_bases, n_rows, Dp)
        assert ends[-1] == 1.0 and all(0 < a < b for a, b in zip(ends, ends[1:])) and len(ends) <= 7
        assert all(0 < c < 1 for c in cuts) and cuts == sorted(cuts)
        pts = [0.0] + cuts + [1.0]
        assert max(b - a for a, b in zip(pts, pts[1:])) <= 0.15 + 1e-9
        assert all(any(abs(e - c) < 1e-12 for c in cuts) for e in ends[:-1])     # every phase end is a cut
    _, c2 = phase_schedule(249e6, 50_000, 512)
    assert c2[0] == 0.6 and len(c2) == 3                      # copy-bound: what round 1 had tuned by hand
    _, c3 = phase_schedule(7.4e9, 992_000, 256)
    assert c3[0] < 0.2 and len(c3) >= 4                       # compute-bound: start a sixth into the copy
