"""Host-only construction of bench.py's synthetic workloads (TEST / BENCH INFRASTRUCTURE).

The native arm builds its read set, k-mer library and projection on the GPU
(fedrann_b200/synthetic_device.py).  The reference arm must not touch the GPU nor
load libfedrann_b200.so, so the same objects are rebuilt here with host threads
(oracle/workload_gen.c): byte-identical reads, the identical library (codes, counts,
order) and therefore the identical projection matrix.  tests/test_host_cpu.py
pins the equality against the torch generator run on the CPU device.

Nothing in here imports torch or fedrann_b200.engine.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass
from os.path import abspath, dirname, isfile, join

import numpy as np

_HERE = dirname(abspath(__file__))
_LIB_PATH = join(_HERE, "libworkload.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not isfile(_LIB_PATH):
            subprocess.run(["make", "-C", _HERE, "libworkload.so"], check=True, capture_output=True)
        L = C.CDLL(_LIB_PATH)
        vp, i64, u64, u32, i32 = C.c_void_p, C.c_int64, C.c_uint64, C.c_uint32, C.c_int
        L.wl_genome.argtypes = [vp, i64, u64, i32]
        L.wl_genome.restype = None
        model = [vp, vp, vp, vp, vp, u64, u32, u32, u32]
        L.wl_read_lengths.argtypes = model + [i64, i64, vp, i32]
        L.wl_read_lengths.restype = None
        L.wl_read_bases.argtypes = model + [i64, i64, vp, vp, i32]
        L.wl_read_bases.restype = None
        L.wl_count_sampled.argtypes = model + [i64, i64, i32, u32, u32, i32, i32]
        L.wl_count_sampled.restype = i64
        L.wl_count_fetch.argtypes = [vp, vp]
        L.wl_count_fetch.restype = None
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _thr24(p: float) -> int:
    return int(p * (1 << 24)) if p > 0 else 0


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


@dataclass
class HostPlan:
    """Same fields as synthetic_device.ReadPlan (built by the same numpy code there)."""
    genome_len: int
    seed: int
    lens: np.ndarray
    starts: np.ndarray
    strands: np.ndarray
    sub: float
    ins: float
    dele: float

    @property
    def n_reads(self) -> int:
        return len(self.lens)


def plan_reads(genome_len: int, coverage: float, seed: int, *, mean_len: int, len_model: str = "lognormal",
               sigma: float = 0.5, min_len: int = 1000, sub: float = 0.0, ins: float = 0.0,
               dele: float = 0.0) -> HostPlan:
    """synthetic_device.plan_reads restated (pure numpy; the draw order is the contract)."""
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
    return HostPlan(genome_len, seed, lens, starts, strands, sub, ins, dele)


class HostWorkload:
    """genome + plan on the host; reads and the library on demand."""

    def __init__(self, plan, threads: int | None = None):
        self.plan = plan
        self.threads = threads or host_threads()
        self.genome = np.empty(plan.genome_len, dtype=np.uint8)
        lib().wl_genome(_p(self.genome), plan.genome_len, plan.seed & (2 ** 64 - 1), self.threads)
        self.lens = np.ascontiguousarray(plan.lens, dtype=np.int64)
        self.starts = np.ascontiguousarray(plan.starts, dtype=np.int64)
        self.strands = np.ascontiguousarray(plan.strands, dtype=np.int8)
        self.pre = np.zeros(len(self.lens) + 1, dtype=np.int64)
        np.cumsum(self.lens, out=self.pre[1:])

    def _model(self):
        p = self.plan
        return (_p(self.genome), _p(self.lens), _p(self.starts), _p(self.strands), _p(self.pre),
                p.seed & (2 ** 64 - 1), _thr24(p.sub), _thr24(p.ins), _thr24(p.dele))

    def reads(self, r0: int, r1: int):
        """-> (bases uint8 ASCII, offsets int64[r1-r0+1]) of reads [r0, r1)."""
        n = r1 - r0
        lens = np.zeros(max(n, 1), dtype=np.int64)
        lib().wl_read_lengths(*self._model(), r0, r1, _p(lens), self.threads)
        offs = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(lens[:n], out=offs[1:])
        bases = np.empty(max(int(offs[-1]), 1), dtype=np.uint8)
        lib().wl_read_bases(*self._model(), r0, r1, _p(offs), _p(bases), self.threads)
        return bases[: int(offs[-1])], offs

    def sampled_counts(self, k: int, fraction: float, seed: int, r0: int = 0, r1: int | None = None):
        """(codes uint64 ascending, counts int64) of the sampled canonical k-mers of reads [r0, r1)
        (synthetic_device.sampled_kmer_counts)."""
        r1 = self.plan.n_reads if r1 is None else r1
        thr = min(int(fraction * 2.0 ** 64), 2 ** 64 - 1)
        all_pass = 1 if thr == 2 ** 64 - 1 else 0
        s = seed & (2 ** 64 - 1)
        seedmix = (s & 0xFFFFFFFF) ^ (((s >> 32) * 0x85EBCA77) & 0xFFFFFFFF)
        n = lib().wl_count_sampled(*self._model(), r0, r1, k, seedmix, thr >> 32, all_pass, self.threads)
        codes = np.empty(max(n, 1), dtype=np.uint64)
        cnts = np.empty(max(n, 1), dtype=np.uint32)
        lib().wl_count_fetch(_p(codes), _p(cnts))
        codes, cnts = codes[:n], cnts[:n]
        u, inv = np.unique(codes, return_inverse=True)
        tot = np.bincount(inv, weights=cnts.astype(np.float64), minlength=len(u)).astype(np.int64)
        return u, tot


def library_from_counts(codes: np.ndarray, counts: np.ndarray, k: int, min_mult: int, seed: int):
    """synthetic_device.library_from_counts on the host -> fedrann_b200.synthetic.KmerLibrary."""
    from fedrann_b200.synthetic import KmerLibrary, splitmix64
    keep = counts >= min_mult
    codes, counts = codes[keep], counts[keep]
    with np.errstate(over="ignore"):
        order = np.argsort(splitmix64(codes + np.uint64(seed & (2 ** 64 - 1))), kind="stable")
    return KmerLibrary(k=k, fwd_codes=codes[order].copy(), counts=counts[order].astype(np.uint64))
