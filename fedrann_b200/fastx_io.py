"""FASTA/FASTQ(.gz) input for the hot path.  The reference shells out to pigz
and seqkit (fedrann/fastx_io.py:186-226, count_kmers.py:57-70) to obtain an
uncompressed FASTA that kmer_searcher then parses (kmer_searcher.cpp:153-201);
here the file is decompressed in-process and parsed once by the native parser
shared with the kmer_searcher replacement (fr_fastx_parse)."""
from __future__ import annotations

import gzip
from dataclasses import dataclass

import numpy as np

from .engine import parse_fastx

_FASTA_EXT = (".fasta", ".fa")
_FASTQ_EXT = (".fastq", ".fq")


@dataclass
class ReadSet:
    bases: np.ndarray      # uint8 ASCII, all reads concatenated
    offsets: np.ndarray    # int64 [R+1]
    names: list            # R names

    @property
    def n_reads(self) -> int:
        return len(self.offsets) - 1


def check_format(path: str) -> str:
    """count_kmers.py:57-74: '.gz' is stripped, then the extension decides."""
    p = path[:-3] if path.endswith(".gz") else path
    if p.endswith(_FASTQ_EXT):
        return "fastq"
    if p.endswith(_FASTA_EXT):
        return "fasta"
    raise ValueError("Unsupported file format. Please provide a FASTA or FASTQ file.")


def read_text(path: str) -> bytes:
    if path.endswith(".gz"):
        with gzip.open(path, "rb") as f:
            return f.read()
    with open(path, "rb") as f:
        return f.read()


def load_reads(path: str) -> ReadSet:
    """The records kmer_searcher would see.  In the reference a FASTQ input is
    first converted to FASTA by seqkit (count_kmers.py:66-68), so names are cut
    at the first blank in both formats."""
    fmt = check_format(path)
    bases, offsets, names = parse_fastx(read_text(path))
    if fmt == "fastq":
        names = [n.split(" ")[0].split("\t")[0] for n in names]
    return ReadSet(bases, offsets, names)


def reverse_complement_codes(codes: np.ndarray, k: int) -> np.ndarray:
    """Reverse complement of 2-bit k-mer codes (A0 C1 G2 T3, first base most significant)."""
    x = (~codes.astype(np.uint64)) & np.uint64((1 << (2 * k)) - 1)
    out = np.zeros_like(x)
    for _ in range(k):
        out = (out << np.uint64(2)) | (x & np.uint64(3))
        x = x >> np.uint64(2)
    return out


def codes_to_text(codes: np.ndarray, k: int) -> list:
    lut = np.frombuffer(b"ACGT", dtype=np.uint8)
    c = codes.astype(np.uint64)
    shifts = (np.arange(k - 1, -1, -1, dtype=np.uint64) * np.uint64(2))
    mat = lut[((c[:, None] >> shifts[None, :]) & np.uint64(3)).astype(np.int64)]
    return [row.tobytes().decode() for row in mat]
