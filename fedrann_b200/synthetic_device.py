"""Large seeded synthetic read sets generated with torch ops on whatever
device is given (harness for bench.py; the small numpy twin is synthetic.py).

Benchmarks cannot push 0.25-45 Gbases of ASCII through numpy loops or PCIe
(SURVEY.md 7.3 "input feeding"), so the reads of the big configs are built on
the device.  All per-base randomness is a counter hash of the GLOBAL base
position, so a rank that generates only reads [a, b) gets byte-for-byte the
reads the single-GPU run has at those positions: strong-scaling runs shard one
well-defined data set.

The library is what `jellyfish count -C | dump -L min | awk sample` +
`seqkit seq -r -p` produce (fedrann/count_kmers.py:76-133) up to the sampler,
which is a seeded hash threshold here (jellyfish/awk parity is
unpinned): canonical k-mers of all reads with multiplicity >= min_mult whose
hash falls under the sample fraction, in hash order.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from .synthetic import KmerLibrary

_MASK64 = (1 << 64) - 1


def _i64(x: int) -> int:
    x &= _MASK64
    return x - (1 << 64) if x >= (1 << 63) else x


def _lsr(x: torch.Tensor, s: int) -> torch.Tensor:
    """logical shift right of int64 lanes"""
    return (x >> s) & ((1 << (64 - s)) - 1)


def splitmix64_t(x: torch.Tensor) -> torch.Tensor:
    """splitmix64 on int64 lanes (two's-complement wrap == uint64 arithmetic);
    bit-identical to synthetic.splitmix64."""
    x = x + _i64(0x9E3779B97F4A7C15)
    x = (x ^ _lsr(x, 30)) * _i64(0xBF58476D1CE4E5B9)
    x = (x ^ _lsr(x, 27)) * _i64(0x94D049BB133111EB)
    return x ^ _lsr(x, 31)


def _u64_less(a: torch.Tensor, thr: int) -> torch.Tensor:
    """unsigned a < thr for int64-encoded lanes"""
    flip = _i64(1 << 63)
    return (a ^ flip) < _i64(thr ^ (1 << 63))


@dataclass
class ReadPlan:
    """Per-read parameters of the whole data set (host, identical on every rank)."""
    genome_len: int
    seed: int
    lens: np.ndarray     # int64 [R] pre-indel lengths
    starts: np.ndarray   # int64 [R]
    strands: np.ndarray  # int8  [R]
    sub: float
    ins: float
    dele: float

    @property
    def n_reads(self) -> int:
        return len(self.lens)

    @property
    def pre_offsets(self) -> np.ndarray:
        o = np.zeros(self.n_reads + 1, dtype=np.int64)
        np.cumsum(self.lens, out=o[1:])
        return o


def plan_reads(genome_len: int, coverage: float, seed: int, *, mean_len: int, len_model: str = "lognormal",
               sigma: float = 0.5, min_len: int = 1000, sub: float = 0.0, ins: float = 0.0,
               dele: float = 0.0) -> ReadPlan:
    rng = np.random.default_rng(seed)
    n_reads = max(1, int(round(genome_len * coverage / mean_len)))
    if len_model == "fixed":
        lens = np.full(n_reads, mean_len, dtype=np.int64)
    elif len_model == "lognormal":
        mu = np.log(mean_len) - sigma * sigma / 2
        lens = np.maximum(min_len, rng.lognormal(mu, sigma, n_reads)).astype(np.int64)
    elif len_model == "normal":
        lens = np.maximum(min_len, rng.normal(mean_len, sigma, n_reads)).astype(np.int64)
    else:
        raise ValueError(len_model)
    lens = np.minimum(lens, genome_len)
    starts = (rng.random(n_reads) * (genome_len - lens + 1)).astype(np.int64)
    strands = rng.integers(0, 2, n_reads).astype(np.int8)
    return ReadPlan(genome_len, seed, lens, starts, strands, sub, ins, dele)


def make_genome(genome_len: int, seed: int, device) -> torch.Tensor:
    """uint8 codes 0..3; counter-hash based so every rank builds the same genome."""
    out = torch.empty(genome_len, dtype=torch.uint8, device=device)
    step = 1 << 26
    for a in range(0, genome_len, step):
        b = min(genome_len, a + step)
        pos = torch.arange(a, b, dtype=torch.int64, device=device)
        out[a:b] = (_lsr(splitmix64_t(pos ^ _i64(seed * 0x2545F4914F6CDD1D + 0x1234567)), 40) & 3).to(torch.uint8)
    return out


_LUT = None


def _ascii_lut(device):
    return torch.tensor([65, 67, 71, 84], dtype=torch.uint8, device=device)


def generate_reads(plan: ReadPlan, genome: torch.Tensor, r0: int, r1: int, *, chunk_bases: int = 1 << 26):
    """Reads [r0, r1) of the plan -> (bases uint8 ASCII on genome.device, 16-byte
    aligned storage; offsets int64 [r1-r0+1] on the host)."""
    dev = genome.device
    pre = plan.pre_offsets
    lut = _ascii_lut(dev)
    pieces, new_lens = [], []
    thr = lambda p: int(p * (1 << 24))  # noqa: E731
    a = r0
    while a < r1:
        b = a + 1
        while b < r1 and pre[b + 1] - pre[a] <= chunk_bases:
            b += 1
        L = torch.from_numpy(plan.lens[a:b]).to(dev)
        st = torch.from_numpy(plan.starts[a:b]).to(dev)
        sd = torch.from_numpy(plan.strands[a:b].astype(np.int64)).to(dev)
        n = int(pre[b] - pre[a])
        loc = torch.zeros(b - a + 1, dtype=torch.int64, device=dev)
        loc[1:] = torch.cumsum(L, 0)
        rid = torch.repeat_interleave(torch.arange(b - a, device=dev), L, output_size=n)
        j = torch.arange(n, dtype=torch.int64, device=dev) - loc[rid]
        s_r, l_r, d_r = st[rid], L[rid], sd[rid]
        src = torch.where(d_r > 0, s_r + l_r - 1 - j, s_r + j)
        code = genome[src].to(torch.int64)
        code = torch.where(d_r > 0, 3 - code, code)
        del src, s_r, l_r, j
        gpos = torch.arange(int(pre[a]), int(pre[a]) + n, dtype=torch.int64, device=dev)
        h = splitmix64_t(gpos ^ _i64(plan.seed * 0x9E3779B97F4A7C15 + 0xABCDEF))
        if plan.sub > 0:
            m = (h & 0xFFFFFF) < thr(plan.sub)
            code = torch.where(m, (code + 1 + (_lsr(h, 24) & 0xFFFF) % 3) & 3, code)
        if plan.ins > 0 or plan.dele > 0:
            h2 = splitmix64_t(h)
            rep = torch.ones(n, dtype=torch.int64, device=dev)
            if plan.dele > 0:
                rep = rep - ((h2 & 0xFFFFFF) < thr(plan.dele)).to(torch.int64)
            if plan.ins > 0:
                rep = rep + ((_lsr(h2, 24) & 0xFFFFFF) < thr(plan.ins)).to(torch.int64)
            crep = torch.zeros(n + 1, dtype=torch.int64, device=dev)
            crep[1:] = torch.cumsum(rep, 0)
            nl = crep[loc[1:]] - crep[loc[:-1]]
            total = int(crep[-1])
            srcpos = torch.repeat_interleave(torch.arange(n, device=dev), rep, output_size=total)
            out = code[srcpos]
            dup = torch.zeros(total, dtype=torch.bool, device=dev)
            dup[1:] = srcpos[1:] == srcpos[:-1]
            out = torch.where(dup, _lsr(h2[srcpos], 50) & 3, out)
            code = out
            del srcpos, dup, crep, rep, h2
        else:
            nl = L
        pieces.append(lut[code])
        new_lens.append(nl.cpu().numpy())
        del code, h, gpos, rid, d_r
        a = b
    lens = np.concatenate(new_lens) if new_lens else np.zeros(0, np.int64)
    offsets = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(lens, out=offsets[1:])
    total = int(offsets[-1])
    bases = torch.empty(total + 16, dtype=torch.uint8, device=dev)[:total]   # fresh storage: 16-byte aligned
    p = 0
    for t in pieces:
        bases[p:p + t.numel()] = t
        p += t.numel()
    return bases, offsets


def _pair_merge(x: torch.Tensor, width_bits: int, n: int) -> torch.Tensor:
    """x holds w-mers at every position; returns 2w-mers (n fewer trailing positions)."""
    return (x[:-n] << width_bits) | x[n:]


def kmer_codes_t(codes2: torch.Tensor, k: int) -> torch.Tensor:
    """All k-mer codes (first base most significant) of a 0..3 coded int64 sequence."""
    n = codes2.numel() - k + 1
    if n <= 0:
        return codes2.new_zeros(0)
    # binary decomposition of k: build 1,2,4,8,16-mers by doubling, then stitch
    pw = {1: codes2}
    w = 1
    while 2 * w <= k:
        pw[2 * w] = (pw[w][:-w] << (2 * w)) | pw[w][w:]
        w *= 2
    out, have = None, 0
    for w in sorted(pw, reverse=True):
        if k - have >= w:
            seg = pw[w][have:have + n]
            out = seg if out is None else ((out << (2 * w)) | seg)
            have += w
    return out


def revcomp_codes_t(x: torch.Tensor, k: int) -> torch.Tensor:
    """Reverse complement of 2-bit k-mer codes held in int64 lanes."""
    x = (~x) & ((1 << (2 * k)) - 1)
    # reverse the 32 two-bit groups of the 64-bit word, then shift down
    x = (_lsr(x, 2) & 0x3333333333333333) | ((x & 0x3333333333333333) << 2)
    x = (_lsr(x, 4) & 0x0F0F0F0F0F0F0F0F) | ((x & 0x0F0F0F0F0F0F0F0F) << 4)
    x = (_lsr(x, 8) & 0x00FF00FF00FF00FF) | ((x & 0x00FF00FF00FF00FF) << 8)
    x = (_lsr(x, 16) & 0x0000FFFF0000FFFF) | ((x & 0x0000FFFF0000FFFF) << 16)
    x = _lsr(x, 32) | (x << 32)
    return _lsr(x, 64 - 2 * k) if k < 32 else x


def sample_keep_t(c: torch.Tensor, seed: int, fraction: float) -> torch.Tensor:
    """synthetic.sample_keep on int64 lanes (the low 32 bits of an int64 product are exact)."""
    thr = min(int(fraction * 2.0 ** 64), 2 ** 64 - 1)
    if thr == 2 ** 64 - 1:
        return torch.ones_like(c, dtype=torch.bool)
    seed &= 2 ** 64 - 1
    seedmix = (seed & 0xFFFFFFFF) ^ (((seed >> 32) * 0x85EBCA77) & 0xFFFFFFFF)
    m32 = 0xFFFFFFFF
    x = (c & m32) ^ ((_lsr(c, 32) * 0x9E3779B1) & m32) ^ seedmix
    x = x ^ (x >> 16)
    x = (x * 0x85EBCA6B) & m32
    x = x ^ (x >> 13)
    x = (x * 0xC2B2AE35) & m32
    x = x ^ (x >> 16)
    return x < (thr >> 32)


def sampled_kmer_counts(bases: torch.Tensor, offsets: np.ndarray, k: int, fraction: float, seed: int,
                        *, chunk_bases: int = 1 << 25):
    """(codes int64 sorted, counts int64) of the canonical k-mers of these reads
    that pass the seeded hash-threshold sampler.  Windows with a non-ACGT base
    or crossing a read boundary are skipped (what jellyfish does)."""
    dev = bases.device
    lut = torch.full((256,), 4, dtype=torch.int64, device=dev)
    for i, ch in enumerate(b"ACGT"):
        lut[ch] = i
        lut[ch + 32] = i
    thr = min(int(fraction * 2.0 ** 64), 2 ** 64 - 1)
    offs_d = torch.from_numpy(np.ascontiguousarray(offsets)).to(dev)
    total = int(offsets[-1])
    acc = []
    for a in range(0, total, chunk_bases):
        b = min(total, a + chunk_bases + k - 1)
        if b - a < k:
            continue
        c = lut[bases[a:b].to(torch.int64)]
        n = b - a - k + 1
        bad = (c > 3).to(torch.int32)
        cb = torch.zeros(b - a + 1, dtype=torch.int32, device=dev)
        cb[1:] = torch.cumsum(bad, 0)
        ok = (cb[k:k + n] - cb[:n]) == 0
        pos = torch.arange(a, a + n, dtype=torch.int64, device=dev)
        rid = torch.searchsorted(offs_d, pos, right=True) - 1
        ok &= (pos + k) <= offs_d[rid + 1]
        km = kmer_codes_t(c & 3, k)[ok]
        can = torch.minimum(km, revcomp_codes_t(km, k))
        keep = sample_keep_t(can, seed, fraction)
        acc.append(can[keep])
        del c, bad, cb, ok, pos, rid, km, can, keep
    allc = torch.cat(acc) if acc else torch.zeros(0, dtype=torch.int64, device=dev)
    return torch.unique(allc, return_counts=True)


def library_from_counts(codes: torch.Tensor, counts: torch.Tensor, k: int, min_mult: int, seed: int) -> KmerLibrary:
    keep = counts >= min_mult
    codes, counts = codes[keep], counts[keep]
    # jellyfish dumps in hash order, not sorted order (same rule as synthetic._sample)
    hk = splitmix64_t(codes + _i64(seed)) ^ _i64(1 << 63)   # unsigned order as signed
    order = torch.argsort(hk, stable=True)
    codes, counts = codes[order], counts[order]
    return KmerLibrary(k=k, fwd_codes=codes.cpu().numpy().view(np.uint64).copy(),
                       counts=counts.cpu().numpy().astype(np.uint64))
