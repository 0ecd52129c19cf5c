"""CPU-side checks of the C ABI: the library loads, exports every symbol the
header declares, and its host-side parsers/writers agree with the oracle and
the reference goldens (no GPU compute here)."""
import ctypes as C
import os
import re
from os.path import abspath, dirname, join

import numpy as np
import pytest

from oracle import oracle as O
from helpers import EXTRACT_CASES, load_extract_golden, reads_to_arrays

ROOT = dirname(dirname(abspath(__file__)))


def _declared():
    hdr = open(join(ROOT, "include", "fedrann_b200.h")).read()
    return sorted(set(re.findall(r"\b(fr_[a-z0-9_]+)\s*\(", hdr)) - {"fr_ctx"})


def test_library_exports_every_declared_symbol():
    from fedrann_b200 import _lib
    L = C.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 16
    for n in names:
        assert hasattr(L, n), n
    assert sorted(_lib.SYMBOLS) == names
    assert L.fr_version() >= 100


def test_no_device_is_an_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from fedrann_b200 import engine
    with pytest.raises(Exception):
        engine.Engine()


def test_product_never_imports_oracle():
    pkg = join(ROOT, "fedrann_b200")
    pat = re.compile(r"^\s*(from|import)\s+oracle|liboracle|orc_", re.M)
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                assert not pat.search(open(join(dp, f)).read()), join(dp, f)


@pytest.mark.parametrize("tag", EXTRACT_CASES)
def test_host_parsers_match_oracle(tag):
    from fedrann_b200 import engine
    g = load_extract_golden(tag)
    codes = engine.parse_library(g["lib_text"], g["k"])
    lib = O.Library(g["lib_text"], g["k"])
    assert np.array_equal(codes, lib.codes)
    bases, offs, names = engine.parse_fastx(g["fastx_text"])
    recs = O.parse_fastx(g["fastx_text"])
    b2, o2, n2 = reads_to_arrays(recs)
    assert names == n2
    assert np.array_equal(offs, o2)
    assert bases.tobytes() == b2.tobytes()
    assert sorted(names) == sorted(g["names"])


def test_fastx_edge_cases():
    from fedrann_b200 import engine
    cases = [
        b"", b"\n\n", b">a\n", b">\nACGT\n>b x\nAC\n\nGT\n", b"ACGT\n>a\nAC", b">a\tz\nAC\r\nGT\n",
        b"@q1 d\nACGT\n+\n!!!!\n@q2\nAC\n+\n!!\n", b"@q1\nACGT", b"@\nAC\n+\n!!\n", b"\n@q\nAC\n+\n!!\n",
    ]
    for text in cases:
        bases, offs, names = engine.parse_fastx(text)
        recs = O.parse_fastx(text)
        b2, o2, n2 = reads_to_arrays(recs)
        assert names == n2, text
        assert np.array_equal(offs, o2), text
        assert bases.tobytes() == b2.tobytes(), text


def test_output_bin_roundtrip(tmp_path):
    from fedrann_b200 import engine
    names = ["r1", "read two", "z"]
    indptr = np.array([0, 2, 2, 5], dtype=np.int64)
    indices = np.array([3, 9, 0, 4, 2**31 - 1], dtype=np.int32)
    p = str(tmp_path / "output.bin")
    engine.write_output_bin(p, names, indptr, indices)
    recs = O.read_output_bin(p)
    assert recs == [("r1", [3, 9]), ("read two", []), ("z", [0, 4, 2**31 - 1])]
    with pytest.raises(Exception):
        engine.write_output_bin(p, ["bad\x01"], np.array([0, 0]), np.zeros(0, np.int32))
    f = str(tmp_path / "kmer_frequency.bin")
    engine.write_frequency_bin(f, np.array([0, 2, 0, 7], dtype=np.uint32))
    assert O.read_frequency_bin(f) == {1: 2, 3: 7}


def _tricky_fasta(nrec, seed):
    r = np.random.default_rng(seed)
    parts = [b"junk before the first header\n\n"]
    alphabet = np.frombuffer(b"ACGTNacgt>@", dtype=np.uint8)
    for i in range(nrec):
        kind = r.integers(0, 10)
        parts.append(b">\n" if kind == 0 else (b">r%d\tdesc > x\n" % i if kind == 1 else b">r%d some desc\n" % i))
        for _ in range(r.integers(0, 4)):
            line = bytes(r.choice(alphabet, size=int(r.integers(0, 50))))
            if line.startswith(b">"):
                line = b"A" + line
            parts.append(line + (b"\r\n" if r.integers(0, 8) == 0 else b"\n"))
        if r.integers(0, 5) == 0:
            parts.append(b"\n")
    return b"".join(parts)


@pytest.mark.parametrize("threads,min_range", [(1, 1 << 30), (7, 64), (16, 1), (3, 1000)])
def test_fasta_parse_in_parallel_ranges_equals_oracle(monkeypatch, threads, min_range):
    """Large FASTA inputs are parsed by several threads over ranges cut at header lines
    (FR_PARSE_THREADS / FR_PARSE_MIN_RANGE shrink the ranges so that small inputs take that path)."""
    from fedrann_b200 import engine
    monkeypatch.setenv("FR_PARSE_THREADS", str(threads))
    monkeypatch.setenv("FR_PARSE_MIN_RANGE", str(min_range))
    for seed in range(4):
        text = _tricky_fasta(300, seed)
        b2, o2, n2 = reads_to_arrays(O.parse_fastx(text))
        b, o, n = engine.parse_fastx(text)
        assert n == n2 and np.array_equal(o, o2) and b.tobytes() == b2.tobytes()
    # FASTQ whose quality lines begin with '@'
    fq = b"".join(b"@q%d\nACGT\n+\n@@@@\n" % i for i in range(200))
    b2, o2, n2 = reads_to_arrays(O.parse_fastx(fq))
    b, o, n = engine.parse_fastx(fq)
    assert n == n2 and np.array_equal(o, o2) and b.tobytes() == b2.tobytes()


def _tricky_fastq(nrec, seed, odd):
    """Four-line records with quality lines that may begin with '@' or '+'; with `odd` also blank lines, junk
    lines between records, wrapped / truncated records and a missing final newline (inputs on which a guessed
    range start can be wrong: the parser must then give the sequential answer all the same)."""
    r = np.random.default_rng(seed)
    qual = np.frombuffer(b"@+!IJK#5", dtype=np.uint8)
    dna = np.frombuffer(b"ACGTNacgt", dtype=np.uint8)
    parts = []
    for i in range(nrec):
        L = int(r.integers(1, 60))
        seq = bytes(r.choice(dna, size=L))
        q = bytes(r.choice(qual, size=L))
        parts.append(b"@r%d desc %d\n" % (i, L) + seq + b"\n+" + (b"r%d" % i if r.integers(0, 3) == 0 else b"") + b"\n" + q + b"\n")
        if odd:
            kind = r.integers(0, 12)
            if kind == 0:
                parts.append(b"\n")
            elif kind == 1:
                parts.append(b"junk line\n")
            elif kind == 2:
                parts.append(b"@cut%d\nACGT\n" % i)                  # a record without its last two lines
            elif kind == 3:
                parts.append(b"@wrap%d\nACGT\nACGT\n+\n@@@@\n@@@@\n" % i)   # wrapped sequence and quality
            elif kind == 4:
                parts.append(b"@e%d\n\n+\n\n" % i)                 # empty sequence
    text = b"".join(parts)
    return text[:-1] if odd and seed % 2 else text


@pytest.mark.parametrize("threads,min_range", [(1, 1 << 30), (7, 64), (16, 1), (3, 1000)])
@pytest.mark.parametrize("odd", [False, True])
def test_fastq_parse_in_parallel_ranges_equals_oracle(monkeypatch, threads, min_range, odd):
    """FASTQ ranges start at guessed record starts and are used only when every range stops where the next one
    starts; in every case the records are those of the sequential reader (kmer_searcher.cpp:188-194)."""
    from fedrann_b200 import engine
    monkeypatch.setenv("FR_PARSE_THREADS", str(threads))
    monkeypatch.setenv("FR_PARSE_MIN_RANGE", str(min_range))
    for seed in range(6):
        text = _tricky_fastq(250, seed, odd)
        b2, o2, n2 = reads_to_arrays(O.parse_fastx(text))
        b, o, n = engine.parse_fastx(text)
        assert n == n2 and np.array_equal(o, o2) and b.tobytes() == b2.tobytes()
