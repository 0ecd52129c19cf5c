"""Parity of the CUDA path (through the C ABI) against the oracle and the
reference goldens.  Integer / index work is bit-exact; the projection is
bit-exact (same fp32 accumulation order as scipy); kNN distances agree within
1e-5 (fp32 dot-product summation order) and indices agree except where two
neighbours are closer than that tolerance."""
from os.path import join

import numpy as np
import pytest
import torch

from oracle import oracle as O
from fedrann_b200 import synthetic as S
from helpers import GOLDEN, EXTRACT_CASES, load_extract_golden, reads_to_arrays, csr_rows

pytestmark = pytest.mark.gpu

KNN_ATOL = 1e-5


@pytest.fixture(scope="module")
def eng():
    from fedrann_b200.engine import Engine
    e = Engine(0)
    yield e
    e.close()


def gpu_extract(eng, codes, k, bases, offs, with_freq=True):
    eng.build_table(codes, k)
    b = eng.dev(bases) if len(bases) else torch.zeros(16, dtype=torch.uint8, device="cuda")[:0]
    o = eng.dev(offs)
    df = torch.zeros(max(len(codes), 1), dtype=torch.int32, device="cuda") if with_freq else None
    indptr, indices = eng.extract(b, o, df)
    torch.cuda.synchronize()
    return indptr.cpu().numpy(), indices.cpu().numpy(), (df.cpu().numpy()[: len(codes)] if with_freq else None)


@pytest.mark.parametrize("tag", EXTRACT_CASES)
def test_extract_golden_bit_exact(eng, tag):
    from fedrann_b200 import engine
    g = load_extract_golden(tag)
    codes = engine.parse_library(g["lib_text"], g["k"])
    bases, offs, names = engine.parse_fastx(g["fastx_text"])
    indptr, indices, df = gpu_extract(eng, codes, g["k"], bases, offs)
    assert csr_rows(indptr, indices, names) == g["rows"]
    assert {i: int(c) for i, c in enumerate(df) if c} == g["freq"]


def _check_vs_oracle(eng, lib_text, k, bases, offs):
    from fedrann_b200 import engine
    codes = engine.parse_library(lib_text, k)
    L = O.Library(lib_text, k)
    ip, ix, df = O.extract(L, bases, offs)
    indptr, indices, gdf = gpu_extract(eng, codes, k, bases, offs)
    assert np.array_equal(indptr, ip)
    assert np.array_equal(indices, ix)
    assert np.array_equal(gdf.astype(np.uint64), df)
    return indptr, indices


@pytest.mark.parametrize("k", [1, 2, 5, 13, 16, 21, 31])
def test_extract_random_reads_all_k(eng, k):
    g = S.random_genome(20000, 100 + k)
    rs = S.simulate_reads(g, 6, k, mean_len=2500, len_model="lognormal", min_len=1, sub=0.03, ins=0.01,
                          dele=0.01, n_frac=0.004)
    frac = 1.0 if k < 8 else 0.2
    lib = S.library_from_reads(rs, k, frac, 1, 9)
    _check_vs_oracle(eng, lib.searcher_text(), k, rs.bases, rs.offsets)


def test_extract_ragged_and_empty(eng):
    k = 7
    lib_text = b"AAAAAAA\nACGTACG\nCGTACGT\nTTTTTTT\nGTACGTA\n"
    seqs = [b"", b"A", b"ACGTACGTACGT", b"", b"NNNNNNNN", b"acgtacgNacgtacgt", b"TTTTTTTTTTTTTTTTTTT", b"ACG", b""]
    bases = np.frombuffer(b"".join(seqs), dtype=np.uint8)
    offs = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int64)
    _check_vs_oracle(eng, lib_text, k, bases, offs)
    # no reads at all
    _check_vs_oracle(eng, lib_text, k, np.zeros(0, np.uint8), np.zeros(1, np.int64))
    # empty library
    _check_vs_oracle(eng, b"", k, bases, offs)


def test_extract_long_reads_multi_tile_and_spill(eng):
    # every k-mer of the genome is in the library: a 40 kb read carries ~40k distinct
    # hits, far beyond the shared-memory hit list -> exercises compaction + global spill
    k = 15
    g = S.random_genome(50000, 77)
    rs = S.simulate_reads(g, 8, 3, mean_len=40000, len_model="fixed", sub=0.001)
    lib = S.library_from_genome(g, k, 1.0, 8, 5)
    ip, ix = _check_vs_oracle(eng, lib.searcher_text(), k, rs.bases, rs.offsets)
    assert np.diff(ip).max() > 8192
    # a read of one repeated unit: thousands of duplicate hits, few distinct
    rep = np.frombuffer(b"ACGGTCATTGCAGGCTTAACG" * 3000, dtype=np.uint8)
    _check_vs_oracle(eng, lib.searcher_text() + b"ACGGTCATTGCAGGC\nCGGTCATTGCAGGCT\n", k, rep,
                     np.array([0, len(rep)], dtype=np.int64))


def test_extract_properties_at_scale(eng):
    k = 15
    g = S.random_genome(1_000_000, 1)
    rs = S.simulate_reads(g, 20, 2, mean_len=10000, len_model="lognormal", min_len=1000, sub=0.02, ins=0.02,
                          dele=0.02)
    lib = S.library_from_genome(g, k, 0.05, 20, 602)
    codes = lib.all_codes()
    indptr, indices, df = gpu_extract(eng, codes, k, rs.bases, rs.offsets)
    assert indptr[0] == 0 and np.all(np.diff(indptr) >= 0) and indptr[-1] == len(indices)
    rows = np.repeat(np.arange(rs.n_reads), np.diff(indptr))
    d = np.diff(indices.astype(np.int64))
    same = rows[1:] == rows[:-1]
    assert np.all(d[same] > 0)                       # strictly ascending inside every row
    assert indices.min() >= 0 and indices.max() < len(codes)
    assert int(df.sum()) == len(indices)             # checksum of checksums
    assert np.array_equal(np.bincount(indices, minlength=len(codes)), df)
    # idempotence
    i2, x2, _ = gpu_extract(eng, codes, k, rs.bases, rs.offsets)
    assert np.array_equal(i2, indptr) and np.array_equal(x2, indices)
    # a seeded sample of rows against the oracle
    L = O.Library(lib.searcher_text(), k)
    sel = np.random.default_rng(0).choice(rs.n_reads, 40, replace=False)
    for r in sel:
        b = rs.bases[rs.offsets[r]:rs.offsets[r + 1]]
        ip, ix, _ = O.extract(L, b, np.array([0, len(b)], dtype=np.int64))
        assert np.array_equal(indices[indptr[r]:indptr[r + 1]], ix)


def test_project_bit_exact_vs_golden_and_oracle(eng):
    from fedrann_b200 import engine
    z = np.load(join(GOLDEN, "project_s0.npz"))
    g = load_extract_golden("s0")
    codes = engine.parse_library(g["lib_text"], g["k"])
    bases, offs, names = engine.parse_fastx(g["fastx_text"])
    eng.build_table(codes, g["k"])
    indptr, indices = eng.extract(eng.dev(bases), eng.dev(offs))
    D, F = int(z["D"]), int(z["n_features"])
    P = O.precompute_matrix(D, z["counts"], F)
    E = eng.project(indptr, indices, eng.upload_projection(P), F // 2).cpu().numpy()
    E_or = O.project(indptr.cpu().numpy(), indices.cpu().numpy(), P, F // 2)
    assert np.array_equal(E.view(np.uint32), E_or.view(np.uint32))
    pos = {n: i for i, n in enumerate(names)}
    for row, (n, s) in enumerate(zip([str(x) for x in z["names"]], z["strands"])):
        assert np.array_equal(E[2 * pos[n] + int(s)].view(np.uint32), z["E_ref"][row].view(np.uint32))


@pytest.mark.parametrize("D", [1, 33, 256, 500])
def test_project_dense_rows_and_empty_rows(eng, D):
    import scipy.sparse as sp
    rng = np.random.default_rng(D)
    nk, R = 300, 50
    F = 2 * nk
    # a dense-ish P forces same-column collisions inside one staged group and the sequential path
    dens = 0.9 if D <= 33 else 0.05
    P = sp.random(F, D, density=dens, format="csr", dtype=np.float32, random_state=3)
    P.data = (rng.standard_normal(len(P.data)) * 3).astype(np.float32)
    rows = [np.sort(rng.choice(F, rng.integers(0, 120), replace=False)) for _ in range(R)]
    rows[3] = np.zeros(0, dtype=np.int64)
    indptr = np.concatenate([[0], np.cumsum([len(r) for r in rows])]).astype(np.int64)
    indices = np.concatenate(rows).astype(np.int32)
    E = eng.project(eng.dev(indptr), eng.dev(indices), eng.upload_projection(P), nk).cpu().numpy()
    E_or = O.project(indptr, indices, P, nk)
    assert np.array_equal(E.view(np.uint32), E_or.view(np.uint32))
    assert not E[6].any() and not E[7].any()


def _knn_agree(idx, dist, oidx, odist):
    assert np.allclose(dist, odist, atol=KNN_ATOL, rtol=0)
    bad = idx != oidx
    if bad.any():
        # a mismatch is only legitimate between near-ties
        r, c = np.nonzero(bad)
        for i, j in zip(r, c):
            near = np.abs(odist[i] - odist[i, j]) <= 2 * KNN_ATOL
            assert idx[i, j] in oidx[i][near], (i, j)
    return 1.0 - bad.mean()


@pytest.mark.parametrize("mode", [1, 0])
@pytest.mark.parametrize("N,D,k", [(700, 48, 10), (1500, 500, 50), (300, 64, 20), (130, 256, 128)])
def test_knn_vs_oracle(eng, mode, N, D, k):
    rng = np.random.default_rng(N + D)
    X = rng.standard_normal((N, D)).astype(np.float32)
    X[rng.random((N, D)) < 0.5] = 0       # projected embeddings are half zeros
    X[5] = 0                              # zero-norm row
    X[11] = X[10]                         # exact duplicate rows -> ties
    En, Eh, norms = eng.normalize(eng.dev(X))
    idx, dist = eng.knn(En, Eh, norms, k, mode=mode)
    torch.cuda.synchronize()
    oidx, odist = O.knn(X, k)
    agree = _knn_agree(idx.cpu().numpy(), dist.cpu().numpy(), oidx, odist)
    assert agree > 0.99
    # query slice (what one rank of the sharded search does)
    q0, nq = N // 3, N // 4
    i2, d2 = eng.knn(En, Eh, norms, k, q0=q0, nq=nq, mode=mode)
    assert torch.equal(i2, idx[q0:q0 + nq]) and torch.equal(d2, dist[q0:q0 + nq])


def test_normalize_matches_oracle(eng):
    rng = np.random.default_rng(4)
    X = rng.standard_normal((257, 100)).astype(np.float32)
    X[9] = 0
    En, Eh, norms = eng.normalize(eng.dev(X))
    Xn, on = O.normalize(X)
    assert np.allclose(En.cpu().numpy(), Xn, atol=3e-7)
    assert np.allclose(norms.cpu().numpy(), on, rtol=1e-6)
    assert Eh.shape == (257, 128) and not Eh[:, 100:].any()
    assert np.allclose(Eh[:, :100].float().cpu().numpy(), Xn, atol=1e-3)


@pytest.mark.parametrize("k,fraction", [(15, 0.05), (17, 0.2), (9, 1.0), (31, 0.5), (16, 0.3)])
def test_extract_sampled_library_fast_path(eng, k, fraction):
    """A library declared (and verified) as a hash-threshold sample takes the register-only
    pre-filter; the CSR must equal the generic path's and the oracle's, N/short-read quirks included."""
    g = S.random_genome(40000, k)
    rs = S.simulate_reads(g, 8, k, mean_len=2500, len_model="lognormal", min_len=5, sub=0.02, ins=0.01, dele=0.01,
                          n_frac=0.004)
    lib = S.library_from_reads(rs, k, fraction, 2, 77)
    text = lib.searcher_text()
    from fedrann_b200 import engine
    codes = engine.parse_library(text, k)
    ip, ix, _ = O.extract(O.Library(text, k), rs.bases, rs.offsets)
    b, o = eng.dev(rs.bases), eng.dev(rs.offsets)
    eng.build_table(codes, k)
    assert not eng.sampled
    ip0, ix0 = eng.extract(b, o)
    eng.build_table(codes, k, sampler=(77, fraction))
    assert eng.sampled
    ip1, ix1 = eng.extract(b, o)
    torch.cuda.synchronize()
    assert np.array_equal(ip1.cpu().numpy(), ip) and np.array_equal(ix1.cpu().numpy(), ix)
    assert torch.equal(ip0, ip1) and torch.equal(ix0, ix1)
    # a wrong claim is detected on the device and ignored
    if fraction < 1.0:
        eng.build_table(codes, k, sampler=(78, fraction))
        assert not eng.sampled
        ip2, ix2 = eng.extract(b, o)
        assert torch.equal(ix2, ix1)


def test_host_path_chunked_equals_device_path(eng):
    """run_hot_path_host (reads cross PCIe in pieces, overlapped with extract/project) returns what
    run_hot_path returns on device-resident reads."""
    from fedrann_b200.pipeline import HostBuffers, run_hot_path, run_hot_path_host
    from fedrann_b200.precompute import precompute_matrix
    k, D, knn_k = 15, 64, 10
    g = S.random_genome(80000, 21)
    rs = S.simulate_reads(g, 12, 21, mean_len=3000, len_model="lognormal", min_len=30, sub=0.02, ins=0.01, dele=0.01,
                          n_frac=0.002)
    lib = S.library_from_reads(rs, k, 0.05, 2, 21)
    eng.build_table(lib.all_codes(), k, sampler=(21, 0.05))
    P_dev = eng.upload_projection(precompute_matrix(D, lib.counts, lib.n_features))
    bases_d, offs_d = eng.dev(rs.bases), eng.dev(rs.offsets)
    a = run_hot_path(eng, bases_d, offs_d, P_dev, lib.nk, knn_k)
    hb = HostBuffers.from_device(bases_d, rs.offsets, 2 * rs.n_reads, knn_k)
    b = run_hot_path_host(eng, hb, P_dev, lib.nk, knn_k)          # the default schedule
    assert torch.equal(hb.nbr_idx, a.nbr_idx.cpu()) and torch.equal(hb.nbr_dist, a.nbr_dist.cpu())
    hb.nbr_idx.zero_(); hb.nbr_dist.zero_()
    for chunks, kw in ((1, {}), (3, {}), (8, dict(knn_phase_ends=range(1, 9))), (1000, {}), (8, dict(knn_overlap=False)),
                       (40, dict(knn_phase_ends=range(1, 41))),
                       (8, dict(knn_phase_ends=(3, 6, 7))), (5, dict(async_extract=False)),
                       (4, dict(async_extract=False, knn_overlap=False)), (8, dict(direct_out=False)),
                       ((0.3, 0.5, 0.9), {}), (2, dict(knn_overlap=False, direct_out=False))):
        b = run_hot_path_host(eng, hb, P_dev, lib.nk, knn_k, chunks=chunks, **kw)
        assert torch.equal(a.indptr, b.indptr) and torch.equal(a.indices, b.indices)
        assert torch.equal(a.embedding, b.embedding)
        assert torch.equal(a.nbr_idx.cpu(), b.nbr_idx.cpu()) and torch.equal(a.nbr_dist.cpu(), b.nbr_dist.cpu())
        assert torch.equal(hb.nbr_idx, a.nbr_idx.cpu()) and torch.equal(hb.nbr_dist, a.nbr_dist.cpu())
        hb.nbr_idx.zero_(); hb.nbr_dist.zero_()


def test_knn_uncertifiable_rows_take_exact_fallback(eng):
    """Rows whose k-th neighbour is closer to the next ones than the fp16 bound cannot be
    certified by the tensor-core pass; they are re-done exactly (in parallel over the database
    for the first few, by the scan for the rest) and mode 0 still equals mode 1 bit for bit."""
    g = torch.Generator(device="cuda")
    g.manual_seed(3)
    N, D, k = 6000, 96, 20
    X = torch.randn((N, D), device="cuda", generator=g)
    base = torch.randn((1, D), device="cuda", generator=g)
    X[100:500] = base + 2e-4 * torch.randn((400, D), device="cuda", generator=g)     # 400 near-duplicates
    En, Eh, norms = eng.normalize(X)
    i0, d0 = eng.knn(En, Eh, norms, k, mode=0)
    fb = eng.knn_fallback_rows
    i1, d1 = eng.knn(En, Eh, norms, k, mode=1)
    torch.cuda.synchronize()
    assert fb > 40                       # more than one batch of the parallel fallback
    assert torch.equal(d0, d1) and torch.equal(i0, i1)


def test_knn_rows_ready_event(eng):
    """fr_knn_rows_ready_after: the fp32 rows may still be arriving on another stream; the
    tensor-core pass starts on the fp16 copy and only the re-rank waits for the event."""
    g = torch.Generator(device="cuda")
    g.manual_seed(9)
    X = torch.randn((4000, 128), device="cuda", generator=g)
    En, Eh, norms = eng.normalize(X)
    i_ref, d_ref = eng.knn(En, Eh, norms, 15)
    side = torch.cuda.Stream()
    late = torch.zeros_like(En)
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        torch.cuda._sleep(20_000_000)            # ~10 ms: the rows land well after K6 has started
        late.copy_(En)
        ev = torch.cuda.Event()
        ev.record(side)
    i1, d1 = eng.knn(late, Eh, norms, 15, rows_ready=ev)
    i2, d2 = eng.knn(late, Eh, norms, 15, mode=1)
    torch.cuda.synchronize()
    assert torch.equal(i1, i_ref) and torch.equal(d1, d_ref) and torch.equal(i2, i_ref)


@pytest.mark.parametrize("N,D,k,bounds", [
    (3000, 100, 10, [0, 1, 130, 130, 1700, 2999, 3000]),          # unaligned, an empty phase, one-row phases
    (20000, 256, 50, [0, 5000, 10000, 15000, 20000]),
    (9000, 500, 50, [0, 9000]),                                   # a single phase
    (7000, 64, 20, [0, 0, 3500, 7000, 7000]),                     # empty first and last phases
])
def test_knn_incremental_equals_one_shot(eng, N, D, k, bounds):
    """fr_knn_inc_*: the database arrives in row ranges; rows that are not there yet hold garbage
    (NaN) while the earlier phases run.  Same bits as fr_knn on the complete database."""
    g = torch.Generator(device="cuda")
    g.manual_seed(N + D)
    X = torch.randn((N, D), device="cuda", generator=g)
    X[N // 2] = 0.0                                               # a zero-norm row
    X[7] = X[3]                                                   # exact duplicates (index tie-break)
    En_full, Eh_full, norms_full = eng.normalize(X)
    i_ref, d_ref = eng.knn(En_full, Eh_full, norms_full, k, mode=1)
    En = torch.full_like(En_full, float("nan"))
    Eh = torch.full_like(Eh_full, float("nan"))
    norms = torch.full_like(norms_full, float("nan"))
    eng.knn_inc_begin(En, Eh, norms, k, bounds)
    for j in range(len(bounds) - 1):
        a, b = bounds[j], bounds[j + 1]
        En[a:b], Eh[a:b], norms[a:b] = En_full[a:b], Eh_full[a:b], norms_full[a:b]
        eng.knn_inc_phase(j)
    idx, dist = eng.knn_inc_finish()
    torch.cuda.synchronize()
    assert torch.equal(idx, i_ref) and torch.equal(dist, d_ref)


def test_knn_incremental_argument_errors(eng):
    from fedrann_b200._lib import FedrannError
    X = torch.randn((600, 32), device="cuda")
    En, Eh, norms = eng.normalize(X)
    with pytest.raises(FedrannError):
        eng.knn_inc_begin(En, Eh, norms, 10, [0, 300])            # does not reach n_db
    with pytest.raises(FedrannError):
        eng.knn_inc_begin(En, Eh, norms, 10, [0, 400, 300, 600])  # decreasing
    with pytest.raises(FedrannError):
        eng.knn_inc_begin(En, Eh, norms, 400, [0, 600])           # k outside the tensor-core path
    eng.knn_inc_begin(En, Eh, norms, 10, [0, 300, 600])
    with pytest.raises(FedrannError):
        eng.knn_inc_phase(1)                                      # out of order
    with pytest.raises(FedrannError):
        eng.knn_inc_finish()                                      # phases missing
    eng.knn_inc_phase(0)
    eng.knn_inc_phase(1)
    idx, dist = eng.knn_inc_finish()
    i1, d1 = eng.knn(En, Eh, norms, 10, mode=1)
    assert torch.equal(idx, i1) and torch.equal(dist, d1)
    with pytest.raises(FedrannError):
        eng.knn_inc_phase(0)                                      # nothing in progress


def test_extract_async_equals_sync_and_reports_overflow(eng):
    """fr_extract_async: same CSR as count + fill without a host round trip; a buffer that is too
    small is reported in the status word and nothing is written."""
    k = 13
    g = S.random_genome(30000, 5)
    rs = S.simulate_reads(g, 10, 5, mean_len=2000, len_model="lognormal", min_len=20, sub=0.02, ins=0.01, dele=0.01,
                          n_frac=0.003)
    lib = S.library_from_reads(rs, k, 0.3, 1, 5)
    eng.build_table(lib.all_codes(), k)
    b, o = eng.dev(rs.bases), eng.dev(rs.offsets)
    ip0, ix0 = eng.extract(b, o)
    nnz = ix0.numel()
    status = torch.zeros(2, dtype=torch.int64, device="cuda")
    ix = torch.full((nnz + 100,), -7, dtype=torch.int32, device="cuda")
    ip = eng.extract_async(b, o, ix, status)
    torch.cuda.synchronize()
    assert status.tolist() == [0, nnz]
    assert torch.equal(ip, ip0) and torch.equal(ix[:nnz], ix0) and bool((ix[nnz:] == -7).all())
    # exactly enough room, then one entry short
    ix = torch.full((nnz,), -7, dtype=torch.int32, device="cuda")
    status.zero_()
    eng.extract_async(b, o, ix, status)
    assert status.tolist() == [0, nnz] and torch.equal(ix, ix0)
    ix = torch.full((nnz - 1,), -7, dtype=torch.int32, device="cuda")
    status.zero_()
    eng.extract_async(b, o, ix, status)
    assert status.tolist() == [4, nnz] and bool((ix == -7).all())
    # the flag is sticky across calls on the same status word
    eng.extract_async(b, o, torch.empty(nnz, dtype=torch.int32, device="cuda"), status)
    assert status.tolist()[0] == 4


def test_host_path_redoes_chunks_that_overflow(eng):
    """A dense library (every k-mer sampled: about one hit per base) does not fit the room the
    asynchronous chunks get; the step is redone synchronously and the result is the same."""
    from fedrann_b200.pipeline import HostBuffers, run_hot_path, run_hot_path_host
    from fedrann_b200.precompute import precompute_matrix
    k, D, knn_k = 11, 32, 5
    g = S.random_genome(20000, 8)
    rs = S.simulate_reads(g, 8, 8, mean_len=1500, len_model="lognormal", min_len=50, sub=0.01, ins=0.0, dele=0.0)
    lib = S.library_from_reads(rs, k, 1.0, 1, 8)
    eng.build_table(lib.all_codes(), k)
    P_dev = eng.upload_projection(precompute_matrix(D, lib.counts, lib.n_features))
    bases_d, offs_d = eng.dev(rs.bases), eng.dev(rs.offsets)
    a = run_hot_path(eng, bases_d, offs_d, P_dev, lib.nk, knn_k)
    assert a.indices.numel() > rs.bases.size // 4
    hb = HostBuffers.from_device(bases_d, rs.offsets, 2 * rs.n_reads, knn_k)
    b = run_hot_path_host(eng, hb, P_dev, lib.nk, knn_k, chunks=4)
    assert torch.equal(a.indptr, b.indptr) and torch.equal(a.indices, b.indices)
    assert torch.equal(a.nbr_idx.cpu(), b.nbr_idx.cpu()) and torch.equal(a.nbr_dist.cpu(), b.nbr_dist.cpu())
    assert torch.equal(hb.nbr_idx, a.nbr_idx.cpu())


# ---------------------------------------------------------------------------
# kNN at the sizes of the BASELINE configurations (the regimes the small cases above never reach:
# a database far larger than L2 with the rotated tile order, several candidate buffers per row,
# the bounded merge of K7).  The whole answer cannot be checked by the CPU oracle in seconds, so
# seeded blocks of query rows are re-answered by the exact fp32 scan (fr_knn mode 1, itself pinned
# to the oracle by test_knn_vs_oracle) and must agree bit for bit.
# ---------------------------------------------------------------------------
def _clustered_rows(N, D, seed, n_centres=None, noise=0.35, zero_frac=0.5):
    """Unit-ish rows with real neighbourhoods (reads overlapping the same locus) and the exact zeros
    projected embeddings have."""
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    n_centres = n_centres or max(1, N // 40)
    C = torch.randn((n_centres, D), device="cuda", generator=g)
    a = torch.randint(0, n_centres, (N,), device="cuda", generator=g)
    X = C[a] + noise * torch.randn((N, D), device="cuda", generator=g)
    X[torch.rand((N, D), device="cuda", generator=g) < zero_frac] = 0
    return X


def _check_blocks_vs_scan(eng, En, Eh, norms, k, idx, dist, q0, nq, seed, blocks=4, rows=64):
    rng = np.random.default_rng(seed)
    checked = 0
    for o in sorted(set(int(x) for x in rng.integers(0, max(1, nq - rows + 1), blocks))):
        n = min(rows, nq - o)
        ei, ed = eng.knn(En, Eh, norms, k, q0=q0 + o, nq=n, mode=1)
        assert torch.equal(ei, idx[o:o + n]), "neighbour ids differ from the exact scan at rows %d.." % (q0 + o)
        assert torch.equal(ed.view(torch.int32), dist[o:o + n].view(torch.int32))
        checked += n
    return checked


def test_knn_c2_shape_50k_x_500(eng):
    """BASELINE configs[1] shape: 50 000 rows x D = 500, k = 50 (the reference's defaults -n 500,
    --nndescent-n-neighbors 50, fedrann/__main__.py:124-141) must take the tensor-core path."""
    N, D, k = 50_000, 500, 50
    X = _clustered_rows(N, D, 21)
    X[77] = 0
    En, Eh, norms = eng.normalize(X)
    del X
    idx, dist = eng.knn(En, Eh, norms, k)
    assert eng.knn_last_path == "tcgen05"
    fb = eng.knn_fallback_rows
    assert _check_blocks_vs_scan(eng, En, Eh, norms, k, idx, dist, 0, N, 1) >= 192
    assert fb < N // 100
    # ascending distances, self first for non-zero rows
    assert bool((dist[:, 1:] >= dist[:, :-1]).all())
    nz = norms > 0
    assert bool((idx[nz, 0] == torch.arange(N, device="cuda", dtype=torch.int32)[nz]).float().mean() > 0.999)


def test_knn_1m_x_256_database_slice(eng):
    """chr1-scale database (1 M x 256, BASELINE configs[2]); one rank's query slice of an 8-way split
    (rows [q0, q0 + nq) of the same array): the fp16 copy is 0.5 GB >> L2, every CTA streams it in the
    rotated order, a query tile is shared by several CTAs (several candidate buffers per row)."""
    N, D, k = 1_000_000, 256, 50
    X = _clustered_rows(N, D, 22, n_centres=N // 60)
    En, Eh, norms = eng.normalize(X)
    del X
    q0, nq = 371_234, 12_345          # unaligned on purpose
    idx, dist = eng.knn(En, Eh, norms, k, q0=q0, nq=nq)
    assert eng.knn_last_path == "tcgen05"
    assert _check_blocks_vs_scan(eng, En, Eh, norms, k, idx, dist, q0, nq, 2, blocks=3, rows=48) >= 96
    assert eng.knn_fallback_rows < 50


def test_knn_many_phase_incremental_200k_bounded_merge(eng):
    """A 24-phase incremental search over 200 000 rows: every row ends with dozens of candidate buffers,
    far more keys than K7's merge area (2 x CAP), so the filtered merge (rr_merge_filtered: keys below the
    row's shared threshold - 2 eps are dropped while merging) is what produces the answer."""
    N, D, k = 200_000, 64, 20
    X = _clustered_rows(N, D, 23, n_centres=N // 50, zero_frac=0.3)
    En, Eh, norms = eng.normalize(X)
    del X
    cuts = np.linspace(0, N, 25).astype(np.int64)
    cuts[1:-1] += np.random.default_rng(3).integers(-500, 500, 23)      # unaligned phase bounds
    eng.knn_inc_begin(En, Eh, norms, k, [int(c) for c in cuts])
    for j in range(24):
        eng.knn_inc_phase(j)
    idx, dist = eng.knn_inc_finish()
    fb = eng.knn_fallback_rows
    assert _check_blocks_vs_scan(eng, En, Eh, norms, k, idx, dist, 0, N, 4) >= 192
    i1, d1 = eng.knn(En, Eh, norms, k)                                   # one-shot search: same bits
    assert torch.equal(i1, idx) and torch.equal(d1.view(torch.int32), dist.view(torch.int32))
    assert fb < N // 20


def test_knn_shapes_outside_the_tensor_core_path_are_reported(eng):
    """D > 512 is answered by the fp32 scan (same lists); the engine says so instead of silently
    being ~100x slower, and k > 128 is refused up front."""
    from fedrann_b200.nearest_neighbors import check_knn_limits
    X = _clustered_rows(600, 576, 24)
    En, Eh, norms = eng.normalize(X)
    i0, d0 = eng.knn(En, Eh, norms, 10)
    assert eng.knn_last_path == "scan"
    i1, d1 = eng.knn(En, Eh, norms, 10, mode=1)
    assert torch.equal(i0, i1) and torch.equal(d0, d1)
    with pytest.raises(ValueError):
        check_knn_limits(129, 500)
