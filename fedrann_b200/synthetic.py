"""Seeded synthetic read sets and k-mer libraries (harness for tests and bench).

The reference gets its library from jellyfish + awk + seqkit
(fedrann/count_kmers.py:76-133); none of those tools exist here and their
sampling is not reproducible without them, so the harness
builds the library itself: canonical k-mers, a count per k-mer, a seeded
hash-threshold sample, and the reverse complements appended (the layout
count_kmers.py:137 feeds to kmer_searcher: forward list, then RC list).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

_ASCII = np.frombuffer(b"ACGT", dtype=np.uint8)
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x += np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return x


def sample_keep(codes: np.ndarray, seed: int, fraction: float) -> np.ndarray:
    """N1's sampler (csrc/common.cuh fr_sample_keep): a canonical k-mer c is sampled iff
    fmix32(lo(c) ^ hi(c)*0x9E3779B1 ^ seedmix) < floor(fraction * 2^64) >> 32 (everything at fraction 1)."""
    m32 = np.uint64(0xFFFFFFFF)
    thr = min(int(fraction * 2.0 ** 64), 2 ** 64 - 1)
    c = codes.astype(np.uint64)
    if thr == 2 ** 64 - 1:
        return np.ones(len(c), dtype=bool)
    seed &= 2 ** 64 - 1
    seedmix = (seed & 0xFFFFFFFF) ^ (((seed >> 32) * 0x85EBCA77) & 0xFFFFFFFF)
    with np.errstate(over="ignore"):
        x = (c & m32) ^ (((c >> np.uint64(32)) * np.uint64(0x9E3779B1)) & m32) ^ np.uint64(seedmix)
        x ^= x >> np.uint64(16)
        x = (x * np.uint64(0x85EBCA6B)) & m32
        x ^= x >> np.uint64(13)
        x = (x * np.uint64(0xC2B2AE35)) & m32
        x ^= x >> np.uint64(16)
    return x < np.uint64(thr >> 32)


def kmer_codes(codes2: np.ndarray, k: int) -> np.ndarray:
    """All k-mer codes (A0 C1 G2 T3, first base most significant, the encoding
    of kmer_searcher.cpp:138-151) of a 0..3 coded sequence."""
    n = len(codes2) - k + 1
    if n <= 0:
        return np.zeros(0, dtype=np.uint64)
    out = np.zeros(n, dtype=np.uint64)
    for j in range(k):
        out = (out << np.uint64(2)) | codes2[j:j + n].astype(np.uint64)
    return out


def revcomp_codes(codes: np.ndarray, k: int) -> np.ndarray:
    """Reverse complement of 2-bit k-mer codes."""
    x = (~codes.astype(np.uint64)) & np.uint64((1 << (2 * k)) - 1)
    out = np.zeros_like(x)
    for _ in range(k):
        out = (out << np.uint64(2)) | (x & np.uint64(3))
        x = x >> np.uint64(2)
    return out


def codes_to_text(codes: np.ndarray, k: int) -> bytes:
    """One k-mer per line."""
    n = len(codes)
    buf = np.empty((n, k + 1), dtype=np.uint8)
    for j in range(k):
        shift = np.uint64(2 * (k - 1 - j))
        buf[:, j] = _ASCII[((codes >> shift) & np.uint64(3)).astype(np.int64)]
    buf[:, k] = 0x0A
    return buf.tobytes()


@dataclass
class ReadSet:
    bases: np.ndarray      # uint8 ASCII, all reads concatenated
    offsets: np.ndarray    # int64 [R+1]
    names: list

    @property
    def n_reads(self) -> int:
        return len(self.offsets) - 1

    def fasta(self) -> bytes:
        parts = []
        for r, name in enumerate(self.names):
            parts.append(b">" + name.encode() + b"\n")
            parts.append(self.bases[self.offsets[r]:self.offsets[r + 1]].tobytes() + b"\n")
        return b"".join(parts)


@dataclass
class KmerLibrary:
    k: int
    fwd_codes: np.ndarray  # uint64 [nk] canonical k-mers (feature i)
    counts: np.ndarray     # uint64 [nk] multiplicity written to the '>count' headers

    @property
    def nk(self) -> int:
        return len(self.fwd_codes)

    @property
    def n_features(self) -> int:
        return 2 * self.nk

    def all_codes(self) -> np.ndarray:
        """feature index -> k-mer code: forward list then reverse complements."""
        return np.concatenate([self.fwd_codes, revcomp_codes(self.fwd_codes, self.k)])

    def searcher_text(self) -> bytes:
        """What `cat fwd rev | grep -v '^>'` feeds kmer_searcher (count_kmers.py:137)."""
        return codes_to_text(self.all_codes(), self.k)

    def counter_fasta(self) -> bytes:
        """fwd_kmer_library.fasta: '>count' + k-mer lines (precompute.py:44-55)."""
        seqs = codes_to_text(self.fwd_codes, self.k).split(b"\n")[:-1]
        return b"".join(b">%d\n%s\n" % (int(c), s) for c, s in zip(self.counts, seqs))


def random_genome(n: int, seed: int) -> np.ndarray:
    return np.random.default_rng(seed).integers(0, 4, size=n, dtype=np.uint8)


def simulate_reads(genome: np.ndarray, coverage: float, seed: int, *, mean_len: int = 10000,
                   len_model: str = "fixed", sigma: float = 0.5, min_len: int = 1000,
                   sub: float = 0.0, ins: float = 0.0, dele: float = 0.0,
                   n_frac: float = 0.0) -> ReadSet:
    """Reads drawn uniformly from `genome`, random strand, with substitution /
    insertion / deletion errors; `n_frac` sprinkles non-ACGT bases ('N')."""
    rng = np.random.default_rng(seed)
    G = len(genome)
    n_reads = max(1, int(round(G * coverage / mean_len)))
    if len_model == "fixed":
        lens = np.full(n_reads, mean_len, dtype=np.int64)
    elif len_model == "lognormal":
        mu = np.log(mean_len) - sigma * sigma / 2
        lens = np.maximum(min_len, rng.lognormal(mu, sigma, n_reads)).astype(np.int64)
    elif len_model == "normal":
        lens = np.maximum(min_len, rng.normal(mean_len, sigma, n_reads)).astype(np.int64)
    else:
        raise ValueError(len_model)
    lens = np.minimum(lens, G)
    starts = (rng.random(n_reads) * (G - lens + 1)).astype(np.int64)
    strands = rng.integers(0, 2, n_reads)
    offs = np.zeros(n_reads + 1, dtype=np.int64)
    np.cumsum(lens, out=offs[1:])
    codes = np.empty(int(offs[-1]), dtype=np.uint8)
    for r in range(n_reads):
        seg = genome[starts[r]:starts[r] + lens[r]]
        codes[offs[r]:offs[r + 1]] = (3 - seg[::-1]) if strands[r] else seg
    n = len(codes)
    if sub > 0:
        m = rng.random(n) < sub
        codes[m] = (codes[m] + rng.integers(1, 4, int(m.sum()), dtype=np.uint8)) & 3
    if dele > 0 or ins > 0:
        rep = np.ones(n, dtype=np.int64)
        if dele > 0:
            rep[rng.random(n) < dele] = 0
        if ins > 0:
            rep[rng.random(n) < ins] += 1
        new_lens = np.add.reduceat(rep, offs[:-1]) if n_reads else np.zeros(0, np.int64)
        new_lens[lens == 0] = 0
        first = np.zeros(n, dtype=bool)
        src = np.repeat(np.arange(n), rep)
        codes2 = codes[src]
        dup = np.zeros(len(src), dtype=bool)
        dup[1:] = src[1:] == src[:-1]
        codes2[dup] = rng.integers(0, 4, int(dup.sum()), dtype=np.uint8)
        codes = codes2
        del first
        offs = np.zeros(n_reads + 1, dtype=np.int64)
        np.cumsum(new_lens, out=offs[1:])
    bases = _ASCII[codes]
    if n_frac > 0:
        bases = bases.copy()
        bases[rng.random(len(bases)) < n_frac] = ord("N")
    names = ["read%d" % r for r in range(n_reads)]
    return ReadSet(bases=bases, offsets=offs, names=names)


def library_from_reads(reads: ReadSet, k: int, fraction: float, min_mult: int, seed: int) -> KmerLibrary:
    """Exact harness-side stand-in for `jellyfish count -C` + `dump -L` + the awk
    sampler (count_kmers.py:80-123): canonical k-mers of all reads with
    multiplicity >= min_mult, kept when sample_keep(code, seed, fraction).
    Windows containing a non-ACGT base are skipped (jellyfish does the same).
    Small inputs only (np.unique over every k-mer)."""
    lut = np.full(256, 4, dtype=np.uint8)
    for i, ch in enumerate(b"ACGT"):
        lut[ch] = i
        lut[ch + 32] = i
    allc = []
    for r in range(reads.n_reads):
        c = lut[reads.bases[reads.offsets[r]:reads.offsets[r + 1]]]
        if len(c) < k:
            continue
        bad = np.convolve((c > 3).astype(np.int32), np.ones(k, dtype=np.int32), "valid") > 0
        km = kmer_codes(c & 3, k)[~bad]
        allc.append(np.minimum(km, revcomp_codes(km, k)))
    allc = np.concatenate(allc) if allc else np.zeros(0, dtype=np.uint64)
    codes, counts = np.unique(allc, return_counts=True)
    keep = counts >= min_mult
    codes, counts = codes[keep], counts[keep]
    return _sample(codes, counts.astype(np.uint64), k, fraction, seed)


def library_from_genome(genome: np.ndarray, k: int, fraction: float, coverage: float, seed: int) -> KmerLibrary:
    """Large-config shortcut: canonical k-mers of the GENOME (at >= 15x every
    true k-mer clears min-mult 2; error k-mers, which the read-derived library
    would also admit when they recur, are left out), seeded Poisson counts."""
    km = kmer_codes(genome, k)
    codes = np.unique(np.minimum(km, revcomp_codes(km, k)))
    rng = np.random.default_rng(seed ^ 0x5EED)
    counts = np.maximum(2, rng.poisson(coverage, len(codes))).astype(np.uint64)
    return _sample(codes, counts, k, fraction, seed)


def _sample(codes, counts, k, fraction, seed) -> KmerLibrary:
    keep = sample_keep(codes, seed, fraction)
    codes, counts = codes[keep], counts[keep]
    # jellyfish dumps in hash order, not sorted order: shuffle deterministically
    order = np.argsort(splitmix64(codes + np.uint64(seed)), kind="stable")
    return KmerLibrary(k=k, fwd_codes=codes[order], counts=counts[order])
