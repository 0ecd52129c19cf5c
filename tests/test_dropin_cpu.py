"""Host-side drop-in surface (no GPU): overlaps.tsv writer vs the reference's
pandas path, output.bin reader, CLI flags, kmer_searcher argument errors."""
import os
import subprocess
from os.path import abspath, dirname, isfile, join

import numpy as np
import pytest

from oracle import oracle as O

ROOT = dirname(dirname(abspath(__file__)))
KS = join(ROOT, "fedrann_b200", "bin", "kmer_searcher")


def _tsv(tmp_path, names, idx, dist, **kw):
    from fedrann_b200.engine import write_overlaps_tsv
    p = str(tmp_path / "overlaps.tsv")
    n = write_overlaps_tsv(p, names, idx, dist, **kw)
    return open(p).read(), n


def test_overlaps_tsv_matches_pandas(tmp_path):
    rng = np.random.default_rng(5)
    R, k = 40, 7
    names = ["read%d" % i for i in range(R)]
    idx = rng.integers(0, 2 * R, size=(2 * R, k)).astype(np.int32)
    idx[:, 0] = np.arange(2 * R)                      # self at rank 0, as a kNN graph has it
    idx[5, 0], idx[5, 3] = idx[5, 3], 5               # ... or elsewhere
    dist = rng.random((2 * R, k)).astype(np.float32)
    special = np.array([0.0, 1.0, 2.0, 0.5, 1e-4, 9.9999e-5, 1e-5, 1.1920929e-07, 5.9604645e-08, 0.1, 0.99999994,
                        1 / 3, 1e-45, -5.9604645e-08, 0.00012345679, 1.0000001], dtype=np.float32)
    dist.reshape(-1)[: len(special)] = special
    got, n = _tsv(tmp_path, names, idx, dist)
    read_names = [n_ for n_ in names for _ in (0, 1)]
    strands = [0, 1] * R
    want = O.overlaps_tsv(idx, dist, read_names, strands)
    assert got == want
    assert n == want.count("\n") - 1


def test_overlaps_tsv_random_floats_roundtrip(tmp_path):
    # every float32 bit pattern class the path can emit: shortest repr == numpy's
    rng = np.random.default_rng(11)
    bits = rng.integers(0, 0x40000000, size=4000, dtype=np.int64).astype(np.uint32)   # [0, 2)
    d = bits.view(np.float32).reshape(-1, 4)
    idx = np.tile(np.array([[1, 2, 3, 0]], dtype=np.int32), (d.shape[0], 1))
    idx[:, 3] = np.arange(d.shape[0]) % 4 + 4
    names = ["r%d" % i for i in range(d.shape[0] // 2)]
    got, _ = _tsv(tmp_path, names, idx, d)
    col = [ln.split("\t")[5] for ln in got.splitlines()[1:]]
    exp = [str(x) for i, row in enumerate(d) for j, x in enumerate(row) if idx[i, j] != i]
    assert col == exp


def test_overlaps_tsv_quoting_negative_and_append(tmp_path):
    names = ["a\tb", 'q"x', "plain"]
    idx = np.array([[0, 2], [1, -1], [2, 4], [3, 0], [4, 5], [5, 1]], dtype=np.int32)
    dist = np.linspace(0, 1, 12, dtype=np.float32).reshape(6, 2)
    read_names = [n for n in names for _ in (0, 1)]
    want = O.overlaps_tsv(idx, dist, read_names, [0, 1] * 3)
    got, _ = _tsv(tmp_path, names, idx, dist)
    assert got == want
    # rank-ordered multi-GPU output: two appends give the same file
    from fedrann_b200.engine import write_overlaps_tsv
    p = str(tmp_path / "two.tsv")
    write_overlaps_tsv(p, names, idx[:4], dist[:4], row0=0, n_rows_total=6)
    write_overlaps_tsv(p, names, idx[4:], dist[4:], row0=4, n_rows_total=6, append=True)
    assert open(p).read() == want


def test_overlaps_tsv_empty(tmp_path):
    got, n = _tsv(tmp_path, [], np.zeros((0, 0), np.int32), np.zeros((0, 0), np.float32))
    assert n == 0 and got == O.overlaps_tsv(np.zeros((0, 0), np.int32), np.zeros((0, 0), np.float32), [], [])


def test_output_bin_reader_and_metadata(tmp_path):
    from fedrann_b200.engine import write_output_bin
    from fedrann_b200.feature_extraction import get_metadata, parse_kmer_searcher_output, read_output_bin
    names = ["r1", "second read", "e"]
    indptr = np.array([0, 3, 3, 5], dtype=np.int64)
    indices = np.array([1, 4, 9, 0, 7], dtype=np.int32)
    p = str(tmp_path / "output.bin")
    write_output_bin(p, names, indptr, indices)
    n2, ip, ix = read_output_bin(p)
    assert n2 == names and np.array_equal(ip, indptr) and np.array_equal(ix, indices)
    on, rows = zip(*O.read_output_bin(p)) if O.read_output_bin(p) else ((), ())
    assert list(on) == names
    recs = list(parse_kmer_searcher_output(p, 5))
    assert recs[0] == ("r1", (1, 4, 9), 0)
    assert recs[1] == ("r1", [6, 9, 4], 1)                       # feature_extraction.py:137-139
    assert recs[2] == ("second read", (), 0) and recs[5] == ("e", [5, 2], 1)
    rn, st = get_metadata(p, 10)
    assert rn == ["r1", "r1", "second read", "second read", "e", "e"] and st == [0, 1] * 3


def test_output_bin_reader_errors_and_edge_records(tmp_path):
    """feature_extraction.py:109-117 (the reference reader's ValueErrors), empty file sets, non-ASCII ids."""
    import struct
    from fedrann_b200.feature_extraction import get_metadata, read_output_bin
    p = str(tmp_path / "o.bin")

    def put(b):
        with open(p, "wb") as f:
            f.write(b)

    put(b"KMER\x01\x00\x00")
    with pytest.raises(ValueError, match="文件头不完整"):
        read_output_bin(p)
    put(b"KMRE\x01\x00\x00\x00" + struct.pack("<Q", 0))
    with pytest.raises(ValueError, match="无效的文件格式"):
        read_output_bin(p)
    put(b"KMER\x02\x00\x00\x00" + struct.pack("<Q", 0))
    with pytest.raises(ValueError, match="不支持的版本号: 2"):
        read_output_bin(p)
    put(b"KMER\x01\x00\x00\x00" + struct.pack("<Q", 0))
    n, ip, ix = read_output_bin(p)
    assert n == [] and ip.tolist() == [0] and ix.size == 0 and get_metadata(p, 4) == ([], [])
    rec = struct.pack("<H", 2) + b"ab" + struct.pack("<I", 2) + struct.pack("<2Q", 3, 1 << 40)
    put(b"KMER\x01\x00\x00\x00" + struct.pack("<Q", 2) + rec)          # second record missing
    with pytest.raises(ValueError):
        read_output_bin(p)
    put(b"KMER\x01\x00\x00\x00" + struct.pack("<Q", 1) + rec[:-3])     # indices cut short
    with pytest.raises(ValueError):
        read_output_bin(p)
    odd = struct.pack("<H", 3) + b"a\xffb" + struct.pack("<I", 0)          # not UTF-8: bytes >= 128 become '_'
    utf = struct.pack("<H", 3) + "é1".encode() + struct.pack("<I", 1) + struct.pack("<Q", 7)
    put(b"KMER\x01\x00\x00\x00" + struct.pack("<Q", 3) + rec + odd + utf)
    n, ip, ix = read_output_bin(p)
    assert n == ["ab", "a_b", "é1"] and ip.tolist() == [0, 2, 2, 3] and ix.tolist() == [3, 1 << 40, 7]
    assert [(a, list(b)) for a, b in O.read_output_bin(p)] == [("ab", [3, 1 << 40]), ("a_b", []), ("é1", [7])]


def test_cli_flags_and_defaults():
    from fedrann_b200.__main__ import parse_command_line_arguments
    a = parse_command_line_arguments(["-i", "x.fa", "-o", "out"])
    assert (a.kmer_size, a.kmer_sample_fraction, a.kmer_min_multiplicity, a.threads, a.chunk_size,
            a.embedding_dimension, a.nndescent_n_trees, a.nndescent_n_neighbors, a.seed) == \
           (16, 0.005, 2, 1, 1000, 500, 300, 50, 356115)         # fedrann/__main__.py:69-171
    assert not (a.save_feature_matrix or a.keep_intermediates or a.mprof)
    b = parse_command_line_arguments(["--input", "x.fq.gz", "--output-dir", "o", "-k", "15", "-n", "64",
                                      "--kmer-sample-fraction", "0.05", "--kmer-min-multiplicity", "3",
                                      "--seed", "602", "--save-feature-matrix", "--keep-intermediates"])
    assert (b.kmer_size, b.embedding_dimension, b.kmer_sample_fraction, b.kmer_min_multiplicity, b.seed) == \
           (15, 64, 0.05, 3, 602) and b.save_feature_matrix and b.keep_intermediates
    with pytest.raises(SystemExit):
        parse_command_line_arguments(["-o", "out"])


def test_fastx_formats(tmp_path):
    import gzip
    from fedrann_b200.fastx_io import check_format, load_reads
    assert check_format("a.fasta.gz") == "fasta" and check_format("b.fq") == "fastq"
    with pytest.raises(ValueError):
        check_format("reads.txt")
    p = tmp_path / "r.fastq.gz"
    with gzip.open(p, "wb") as f:
        f.write(b"@r1 extra words\nACGT\n+\nIIII\n@r2\nGG\n+\nII\n")
    rs = load_reads(str(p))
    assert rs.names == ["r1", "r2"] and rs.bases.tobytes() == b"ACGTGG" and list(rs.offsets) == [0, 4, 6]
    q = tmp_path / "m.fa"
    q.write_bytes(b">a desc\nAC\nGT\n>b\nTT\n")
    rs = load_reads(str(q))
    assert rs.names == ["a", "b"] and rs.bases.tobytes() == b"ACGTTT"


@pytest.mark.skipif(not isfile(KS), reason="kmer_searcher not built")
def test_kmer_searcher_argument_errors(tmp_path):
    def run(*a):
        return subprocess.run([KS, *a], capture_output=True, text=True)
    r = run()
    assert r.returncode == 1 and "Usage:" in r.stderr                       # kmer_searcher.cpp:235-238
    r = run("lib", "reads", str(tmp_path / "o"), "32", "1")
    assert r.returncode == 1 and "Invalid k value: 32" in r.stderr          # :240-244
    r = run("lib", "reads", str(tmp_path / "o"), "15", "0")
    assert r.returncode == 1 and "Invalid number of threads: 0" in r.stderr  # :246-250
    r = run("lib", str(tmp_path / "missing.fa"), str(tmp_path / "o"), "15", "1")
    assert r.returncode == 1 and os.path.isdir(tmp_path / "o")              # dir created first (:253-256)


def test_count_kmers_oracle_known_answer():
    # ACGTA, k=3: ACG/CGT are each other's reverse complement (canonical ACG x2), GTA -> canonical GTA? TAC vs GTA
    bases = np.frombuffer(b"ACGTAnACG", dtype=np.uint8)
    d = O.count_kmers(bases, np.array([0, 9]), 3, 1.0, 1, 0)
    enc = lambda s: int("".join(str("ACGT".index(c)) for c in s), 4)
    assert d == {enc("ACG"): 3, enc("GTA"): 1}
    assert O.count_kmers(bases, np.array([0, 9]), 3, 1.0, 2, 0) == {enc("ACG"): 3}
