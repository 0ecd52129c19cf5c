"""Drop-in surface on the GPU: N1 k-mer counting vs the oracle, the
kmer_searcher executable vs the reference goldens, the reference-named Python
functions, and the `fedrann` CLI end to end vs the oracle pipeline."""
import os
import subprocess
import sys
from os.path import abspath, dirname, isfile, join

import numpy as np
import pytest
import torch

from oracle import oracle as O
from fedrann_b200 import synthetic as S
from helpers import EXTRACT_CASES, load_extract_golden

pytestmark = pytest.mark.gpu
ROOT = dirname(dirname(abspath(__file__)))
KS = join(ROOT, "fedrann_b200", "bin", "kmer_searcher")


def small_reads(seed=602, genome=60000, cov=12, n_frac=0.002):
    g = S.random_genome(genome, seed)
    return S.simulate_reads(g, cov, seed, mean_len=3000, len_model="lognormal", min_len=40, sub=0.02, ins=0.01,
                            dele=0.01, n_frac=n_frac)


@pytest.mark.parametrize("k,fraction,min_mult", [(15, 0.05, 2), (11, 1.0, 1), (17, 0.3, 3), (31, 0.5, 1), (3, 1.0, 1)])
def test_kmer_count_vs_oracle(k, fraction, min_mult):
    from fedrann_b200.count_kmers import get_engine
    eng = get_engine()
    rs = small_reads(seed=k, genome=20000, cov=6)
    want = O.count_kmers(rs.bases, rs.offsets, k, fraction, min_mult, 99)
    codes, counts = eng.count_kmers(eng.dev(rs.bases), eng.dev(rs.offsets), k, fraction, 99, min_mult)
    got = dict(zip(map(int, codes), map(int, counts)))
    assert got == want


def test_kmer_count_grows_table():
    from fedrann_b200.count_kmers import get_engine
    eng = get_engine()
    rs = small_reads(seed=4, genome=30000, cov=4)
    want = O.count_kmers(rs.bases, rs.offsets, 13, 1.0, 1, 1)
    codes, counts = eng.count_kmers(eng.dev(rs.bases), eng.dev(rs.offsets), 13, 1.0, 1, 1, capacity=1024)
    assert dict(zip(map(int, codes), map(int, counts))) == want


@pytest.mark.skipif(not isfile(KS), reason="kmer_searcher not built")
@pytest.mark.parametrize("tag", EXTRACT_CASES)
def test_kmer_searcher_executable_golden(tmp_path, tag):
    g = load_extract_golden(tag)
    lib, fx = tmp_path / "lib.txt", tmp_path / ("reads.fq" if g["fastx_text"][:1] == b"@" else "reads.fa")
    lib.write_bytes(g["lib_text"])
    fx.write_bytes(g["fastx_text"])
    out = tmp_path / "out"
    r = subprocess.run([KS, str(lib), str(fx), str(out), str(g["k"]), "2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    recs = O.read_output_bin(str(out / "output.bin"))
    assert {n: sorted(ix) for n, ix in recs} == {n: sorted(v) for n, v in g["rows"].items()}
    assert [n for n, _ in recs] == g["names"]                    # input order
    assert O.read_frequency_bin(str(out / "kmer_frequency.bin")) == g["freq"]
    # library on stdin (count_kmers.py:137-139 pipes it)
    out2 = tmp_path / "out2"
    r = subprocess.run([KS, "/dev/stdin", str(fx), str(out2), str(g["k"]), "1"], input=g["lib_text"],
                       capture_output=True)
    assert r.returncode == 0
    assert (out2 / "output.bin").read_bytes() == (out / "output.bin").read_bytes()


def _oracle_pipeline(rs, k, fraction, min_mult, seed, D, knn_k):
    lib_d = O.count_kmers(rs.bases, rs.offsets, k, fraction, min_mult, seed)
    codes = np.array(sorted(lib_d), dtype=np.uint64)
    order = np.lexsort((codes, S.splitmix64(codes + np.uint64(seed))))
    codes = codes[order]
    counts = np.array([lib_d[int(c)] for c in codes], dtype=np.uint64)
    lib = S.KmerLibrary(k=k, fwd_codes=codes, counts=counts)
    L = O.Library(lib.searcher_text(), k)
    ip, ix, _ = O.extract(L, rs.bases, rs.offsets)
    E = O.project(ip, ix, O.precompute_matrix(D, counts, lib.n_features), lib.nk)
    oi, od = O.knn(E, min(knn_k, E.shape[0]))
    return lib, ip, ix, E, oi, od


def test_cli_end_to_end(tmp_path):
    rs = small_reads(seed=7)
    k, fraction, min_mult, seed, D, knn_k = 15, 0.05, 2, 602, 64, 10
    fa = tmp_path / "reads.fasta"
    fa.write_bytes(rs.fasta())
    out = tmp_path / "out"
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-m", "fedrann_b200", "-i", str(fa), "-o", str(out), "-k", str(k),
                        "--kmer-sample-fraction", str(fraction), "--kmer-min-multiplicity", str(min_mult),
                        "--seed", str(seed), "-n", str(D), "--nndescent-n-neighbors", str(knn_k),
                        "--save-feature-matrix", "--keep-intermediates"], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr[-3000:]
    lib, ip, ix, E, oi, od = _oracle_pipeline(rs, k, fraction, min_mult, seed, D, knn_k)
    # intermediates: the library file, output.bin
    assert (out / "temp" / "fwd_kmer_library.fasta").read_bytes() == lib.counter_fasta()
    recs = O.read_output_bin(str(out / "temp" / "kmer_searcher" / "output.bin"))
    assert [n for n, _ in recs] == rs.names
    assert np.array_equal(np.concatenate([np.asarray(v, dtype=np.int64) for _, v in recs]), ix.astype(np.int64))
    # feature_matrix.npz: binary (2R x F) matrix, fwd / index-swapped rows
    from scipy.sparse import load_npz
    fm = load_npz(str(out / "feature_matrix.npz")).tocsr()
    assert fm.shape == (2 * rs.n_reads, lib.n_features)
    for r_ in (0, 3, rs.n_reads - 1):
        row = ix[ip[r_]:ip[r_ + 1]].astype(np.int64)
        assert np.array_equal(np.sort(fm[2 * r_].indices), row)
        sw = np.where(row < lib.nk, row + lib.nk, row - lib.nk)
        assert np.array_equal(np.sort(fm[2 * r_ + 1].indices), np.sort(sw))
    # overlaps.tsv: same text as the reference's writer on the oracle's kNN, up to near-ties
    got = (out / "overlaps.tsv").read_text().splitlines()
    names2 = [n for n in rs.names for _ in (0, 1)]
    want = O.overlaps_tsv(oi, od, names2, [0, 1] * rs.n_reads).splitlines()
    assert got[0] == want[0] and len(got) == len(want)
    # the text of every column but the last must agree (up to near-ties); the printed distance
    # differs in its last digits with the fp32 summation order, so it is compared as a number
    same = sum(a.rsplit("\t", 1)[0] == b.rsplit("\t", 1)[0] for a, b in zip(got, want))
    assert same >= 0.99 * len(want)
    gd = np.array([float(x.split("\t")[5]) for x in got[1:]])
    wd = np.array([float(x.split("\t")[5]) for x in want[1:]])
    assert np.allclose(gd, wd, atol=1e-5)
    assert (out / "fedrann.log").exists()


def test_reference_named_functions(tmp_path):
    """run_kmer_searcher -> get_precompute_matrix -> get_feature_matrix -> get_metadata ->
    NNDescent_ava.get_neighbors, called the way run_fedrann_pipeline calls them (__main__.py:321-376)."""
    from fedrann_b200 import global_variables
    from fedrann_b200.count_kmers import run_kmer_searcher
    from fedrann_b200.feature_extraction import get_feature_matrix, get_metadata
    from fedrann_b200.nearest_neighbors import NNDescent_ava
    from fedrann_b200.precompute import get_precompute_matrix
    rs = small_reads(seed=11, genome=40000, cov=10, n_frac=0.0)
    fa = tmp_path / "reads.fa"
    fa.write_bytes(rs.fasta())
    global_variables.temp_dir = str(tmp_path / "temp")
    global_variables.seed = 5
    k, D, knn_k = 13, 48, 8
    ks, n_features, read_count = run_kmer_searcher(str(fa), k, 0.1, 2)
    assert read_count == rs.n_reads
    P, nf = get_precompute_matrix(n_components=D, counter_file=join(global_variables.temp_dir,
                                                                    "fwd_kmer_library.fasta"), n_features=n_features)
    E = get_feature_matrix(ks_file=ks, precompute_matrix=P, kmer_count=n_features, read_count=read_count,
                           chunk_size=1000)
    lib, ip, ix, Eo, oi, od = _oracle_pipeline(rs, k, 0.1, 2, 5, D, knn_k)
    assert n_features == lib.n_features and nf == n_features
    assert E.dtype == np.float32 and np.array_equal(E.view(np.uint32), Eo.view(np.uint32))     # bit-exact
    names, strands = get_metadata(ks, n_features)
    assert names == [n for n in rs.names for _ in (0, 1)] and strands == [0, 1] * rs.n_reads
    idx, dist = NNDescent_ava().get_neighbors(E, metric="cosine", index_n_neighbors=knn_k, n_trees=300,
                                              leaf_size=200, n_iters=None, diversify_prob=1.0,
                                              pruning_degree_multiplier=1.5, low_memory=True, n_jobs=2, seed=1,
                                              verbose=True)
    assert idx.dtype == np.int32 and dist.dtype == np.float32 and idx.shape == (2 * rs.n_reads, knn_k)
    assert np.allclose(dist, od, atol=1e-5) and (idx == oi).mean() > 0.99
    with pytest.raises(ValueError):
        NNDescent_ava().get_neighbors(E, metric="euclidean")


def _run_cli(tmp_path, fasta_bytes, name="r.fa", extra=()):
    fa = tmp_path / name
    fa.write_bytes(fasta_bytes)
    out = tmp_path / "out"
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-m", "fedrann_b200", "-i", str(fa), "-o", str(out), "-k", "11",
                        "--kmer-sample-fraction", "0.5", "-n", "32", "--nndescent-n-neighbors", "5", *extra],
                       capture_output=True, text=True, env=env)
    return r, out


def test_cli_edge_inputs(tmp_path):
    """Ragged and degenerate inputs through the whole CLI: reads shorter than k, N runs, a read
    without any library k-mer, fewer rows than neighbours asked for."""
    rng = np.random.default_rng(0)
    core = "".join("ACGT"[i] for i in rng.integers(0, 4, 400))
    recs = [("a", core[:300]), ("b", core[50:350]), ("short", "ACG"), ("ns", "N" * 40), ("c", core[100:400]),
            ("lower", core[:200].lower()), ("poly", "A" * 100)]
    fasta = "".join(f">{n} desc\n{s}\n" for n, s in recs).encode()
    r, out = _run_cli(tmp_path, fasta)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = (out / "overlaps.tsv").read_text().splitlines()
    assert lines[0].split("\t") == ["query_name", "query_orientation", "target_name", "target_orientation",
                                    "neighbor_rank", "distance"]
    assert not (out / "temp").exists()                                  # removed without --keep-intermediates
    body = [ln.split("\t") for ln in lines[1:]]
    assert 14 * 4 <= len(body) <= 14 * 5                                # self filtered where it is listed
    # overlapping reads find each other first
    first = {(q, qo): t for q, qo, t, to, rk, d in body if rk == "1"}
    assert first[("a", "+")] in ("b", "lower", "c") and first[("b", "+")] in ("a", "c", "lower")
    # a single read: k is clamped to the row count
    (tmp_path / "one").mkdir()
    r, out = _run_cli(tmp_path / "one", b">x\n" + core.encode() + b"\n", extra=("--kmer-min-multiplicity", "1"))
    assert r.returncode == 0, r.stderr[-2000:]
    rows = (out / "overlaps.tsv").read_text().splitlines()[1:]
    assert len(rows) == 2 and {ln.split("\t")[2] for ln in rows} == {"x"}


def test_cli_rejects_unknown_format(tmp_path):
    r, _ = _run_cli(tmp_path, b">x\nACGT\n", name="reads.txt")
    assert r.returncode != 0 and "Unsupported file format" in r.stderr
