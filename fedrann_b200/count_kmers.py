"""Stage 1 of the pipeline, drop-in for fedrann/count_kmers.py:52-149.

    run_kmer_searcher(input_path, k, sample_fraction, min_multiplicity)
        -> (output_bin_path, n_features = 2 * n_kmers, read_count)

The reference shells out to pigz, seqkit, jellyfish, awk and the kmer_searcher
executable; here the reads are parsed once, uploaded, and everything runs on
the GPU through the C ABI: canonical k-mer counting + seeded sampling (N1,
fr_kmer_count_*), table build (K2) and extraction (K1+K3).  The same files land
in `global_variables.temp_dir`: fwd_kmer_library.fasta ('>count' + k-mer, read
back by get_precompute_matrix), rev_kmer_library.fasta, kmer_searcher/output.bin
and kmer_searcher/kmer_frequency.bin.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from os.path import join

import numpy as np
import torch

from . import global_variables
from .custom_logging import logger
from .engine import Engine, write_frequency_bin, write_output_bin
from .fastx_io import ReadSet, codes_to_text, load_reads, reverse_complement_codes
from .synthetic import splitmix64


@dataclass
class Session:
    """Device-resident state shared by the stages of one pipeline run (the
    reference passes the same information through files in temp/)."""
    eng: Engine
    reads: ReadSet
    k: int
    nk: int
    indptr: torch.Tensor = None        # int64 [R+1]  this rank's reads
    indices: torch.Tensor = None       # int32 [nnz]
    read_lo: int = 0                   # this rank's read range [read_lo, read_hi)
    read_hi: int = 0
    output_bin: str = ""


_session: Session | None = None


def current_session() -> Session | None:
    return _session


_engine: Engine | None = None


def get_engine() -> Engine:
    """One engine (one CUDA context handle) per process = per GPU."""
    global _engine
    if _engine is None:
        _engine = Engine()
    return _engine


def sample_library(eng: Engine, bases_d: torch.Tensor, offsets_d: torch.Tensor, k: int, sample_fraction: float,
                   min_multiplicity: int, seed: int, group=None):
    """-> (fwd codes uint64, counts uint64) in library order.  With a process group each rank
    counts its own reads and the tables are summed before the multiplicity cut."""
    world = torch.distributed.get_world_size(group) if group is not None else 1
    codes, counts = eng.count_kmers(bases_d, offsets_d, k, sample_fraction, seed,
                                    min_multiplicity if world == 1 else 1)
    counts = counts.astype(np.uint64)
    if world > 1:
        # hash-partitioned merge: one all-to-all of the partial counts to their owners, multiplicity cut there,
        # all-gather of the (small) library
        from .parallel import partitioned_kmer_counts
        c, n = partitioned_kmer_counts(torch.from_numpy(codes.view(np.int64)).to(bases_d.device),
                                       torch.from_numpy(counts.astype(np.int64)).to(bases_d.device),
                                       min_multiplicity, group)
        codes, counts = c.cpu().numpy().view(np.uint64), n.cpu().numpy().astype(np.uint64)
    # jellyfish dumps in its hash order; ours is a seeded hash order (deterministic)
    order = np.lexsort((codes, splitmix64(codes + np.uint64(seed & (2 ** 64 - 1)))))
    return codes[order], counts[order]


def write_library_files(temp_dir: str, codes: np.ndarray, counts: np.ndarray, k: int):
    fwd_path = join(temp_dir, "fwd_kmer_library.fasta")
    rev_path = join(temp_dir, "rev_kmer_library.fasta")
    fwd = codes_to_text(codes, k)
    rev = codes_to_text(reverse_complement_codes(codes, k), k)
    with open(fwd_path, "w") as f:
        f.write("".join(">%d\n%s\n" % (int(c), s) for c, s in zip(counts, fwd)))
    with open(rev_path, "w") as f:                      # seqkit seq -r -p keeps the headers
        f.write("".join(">%d\n%s\n" % (int(c), s) for c, s in zip(counts, rev)))
    return fwd_path, rev_path


def run_kmer_searcher(input_path: str, k: int, sample_fraction: float, min_multiplicity: int = 2, *,
                      group=None, write_files: bool = True):
    global _session
    if not 1 <= k <= 31:
        raise ValueError("k must be in [1, 31]")          # kmer_searcher.cpp:240-244
    temp_dir = global_variables.temp_dir or "."
    os.makedirs(temp_dir, exist_ok=True)
    reads = load_reads(input_path)
    eng = get_engine()
    rank = torch.distributed.get_rank(group) if group is not None else 0
    world = torch.distributed.get_world_size(group) if group is not None else 1
    from .parallel import shard_reads
    bounds = shard_reads(reads.offsets, world)
    r0, r1 = int(bounds[rank]), int(bounds[rank + 1])
    b0, b1 = int(reads.offsets[r0]), int(reads.offsets[r1])
    dev = torch.device("cuda", eng.device)
    bases_d = torch.empty(b1 - b0 + 16, dtype=torch.uint8, device=dev)[: b1 - b0]
    bases_d.copy_(torch.from_numpy(reads.bases[b0:b1].copy()))
    offsets_d = torch.from_numpy(reads.offsets[r0:r1 + 1] - b0).to(dev)

    logger.debug("Counting canonical %d-mers on the GPU (sample fraction %g, min multiplicity %d)",
                 k, sample_fraction, min_multiplicity)
    codes, counts = sample_library(eng, bases_d, offsets_d, k, sample_fraction, min_multiplicity,
                                   global_variables.seed, group)
    kmer_count = len(codes)
    logger.debug(f"Number of kmers in the library: {kmer_count}")
    if write_files and rank == 0:
        write_library_files(temp_dir, codes, counts, k)

    # kmer_searcher.cpp:259-273: index = first-occurrence rank over fwd list + RC list; a
    # palindromic k-mer (even k) appears in both halves and its second copy consumes no index
    all_codes = np.concatenate([codes, reverse_complement_codes(codes, k)])
    _, first = np.unique(all_codes, return_index=True)
    if len(first) != len(all_codes):
        all_codes = all_codes[np.sort(first)]
    eng.build_table(all_codes, k, sampler=(global_variables.seed, sample_fraction))
    docfreq = torch.zeros(max(len(all_codes), 1), dtype=torch.int32, device=dev)
    indptr, indices = eng.extract(bases_d, offsets_d, docfreq)

    out_dir = join(temp_dir, "kmer_searcher")
    output_bin = join(out_dir, "output.bin")
    if world > 1:
        torch.distributed.all_reduce(docfreq, group=group)
    if write_files:
        if world > 1:
            # rows of every rank to rank 0, as tensors: per-read hit counts and the indices, in rank = input order
            from .parallel import gather_to_rank0
            cnt_all = gather_to_rank0(indptr[1:] - indptr[:-1], group)
            ix_all = gather_to_rank0(indices, group)
            if rank == 0:
                ip = np.zeros(cnt_all.numel() + 1, dtype=np.int64)
                np.cumsum(cnt_all.cpu().numpy(), out=ip[1:])
                ix = ix_all.cpu().numpy()
        else:
            ip, ix = indptr.cpu().numpy(), indices.cpu().numpy()
        if rank == 0:
            os.makedirs(out_dir, exist_ok=True)              # kmer_searcher.cpp:253-256
            write_output_bin(output_bin, reads.names, ip, ix)
            write_frequency_bin(join(out_dir, "kmer_frequency.bin"), docfreq.cpu().numpy().view(np.uint32))
    _session = Session(eng=eng, reads=reads, k=k, nk=kmer_count, indptr=indptr, indices=indices,
                       read_lo=r0, read_hi=r1, output_bin=output_bin)
    return output_bin, kmer_count * 2, reads.n_reads
