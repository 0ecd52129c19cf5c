"""Stage 3, drop-in for fedrann/feature_extraction.py:108-140, 216-302:
parse_kmer_searcher_output, get_feature_matrix, get_metadata.

get_feature_matrix applies the precomputed projection on the GPU (K4,
fr_project): E[2r] = sum of P's rows over the read's ascending feature indices,
E[2r+1] the same over the index-swapped (reverse-orientation) row -- bit-identical
to scipy's csr_matmat on the reference's A_chunk.  A read with
no hits gets two all-zero rows (the reference shifts rows there, a defect that
is not reproduced).
"""
from __future__ import annotations

import numpy as np
import torch

from .custom_logging import logger


def read_output_bin(ks_file: str):
    """output.bin (kmer_searcher.cpp:98-130) -> (names, indptr int64[R+1], indices int64[nnz])."""
    with open(ks_file, "rb") as f:
        blob = f.read()
    if len(blob) < 16:
        raise ValueError("文件头不完整")
    if blob[:4] != b"KMER":
        raise ValueError("无效的文件格式")
    if blob[4] != 1:
        raise ValueError(f"不支持的版本号: {blob[4]}")
    total = int(np.frombuffer(blob, dtype="<u8", count=1, offset=8)[0])
    names, counts, starts = [], np.zeros(total, dtype=np.int64), np.zeros(total, dtype=np.int64)
    pos = 16
    mv = memoryview(blob)
    for r in range(total):
        id_len = mv[pos] | (mv[pos + 1] << 8)
        pos += 2
        raw = bytes(mv[pos:pos + id_len])
        try:
            names.append(raw.decode("utf-8"))
        except UnicodeDecodeError:
            names.append("".join(chr(b) if b < 128 else "_" for b in raw))
        pos += id_len
        n = int.from_bytes(mv[pos:pos + 4], "little")
        pos += 4
        counts[r], starts[r] = n, pos
        pos += 8 * n
    indptr = np.zeros(total + 1, dtype=np.int64)
    np.cumsum(counts, out=indptr[1:])
    indices = np.empty(int(indptr[-1]), dtype=np.int64)
    buf = np.frombuffer(blob, dtype=np.uint8)
    for r in range(total):
        if counts[r]:
            indices[indptr[r]:indptr[r + 1]] = buf[starts[r]:starts[r] + 8 * counts[r]].view("<u8")
    return names, indptr, indices


def parse_kmer_searcher_output(ks_file: str, kmer_count):
    """Generator with the reference's record stream (feature_extraction.py:108-140):
    (id, indices, 0) then (id, index-swapped indices, 1) per read; kmer_count = n_kmers."""
    names, indptr, indices = read_output_bin(ks_file)
    for r, name in enumerate(names):
        idx = tuple(int(i) for i in indices[indptr[r]:indptr[r + 1]])
        yield name, idx, 0
        yield name, [i + kmer_count if i < kmer_count else i - kmer_count for i in idx], 1


def get_feature_matrix(ks_file: str, precompute_matrix, kmer_count: int, read_count: int, chunk_size: int = 1000):
    """-> float32 (2 * read_count, D); row 2r forward, 2r+1 reverse orientation of record r.
    kmer_count is n_features (= 2 * n_kmers), as the reference passes it (__main__.py:345)."""
    from .count_kmers import get_engine
    eng = get_engine()
    names, indptr, indices = read_output_bin(ks_file)
    if len(names) != read_count:
        logger.debug("output.bin holds %d records, read_count is %d", len(names), read_count)
    nk = int(kmer_count) // 2
    # the reference's set -> COO -> CSR conversion sorts and de-duplicates columns
    rows = np.repeat(np.arange(len(names), dtype=np.int64), np.diff(indptr))
    key = np.unique(rows * np.int64(max(int(kmer_count), 1) + 1) + indices)
    rows_u = key // np.int64(max(int(kmer_count), 1) + 1)
    cols_u = (key % np.int64(max(int(kmer_count), 1) + 1)).astype(np.int32)
    ip = np.zeros(len(names) + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows_u, minlength=len(names)), out=ip[1:])
    P_dev = eng.upload_projection(precompute_matrix)
    E = eng.project(eng.dev(ip), eng.dev(cols_u), P_dev, nk)
    torch.cuda.synchronize()
    out = np.zeros((2 * read_count, P_dev[3]), dtype=np.float32)
    n = min(out.shape[0], E.shape[0])
    out[:n] = E[:n].cpu().numpy()
    return out


def get_metadata(ks_file: str, kmer_count: int):
    """-> (read_names, strands), two entries per record (feature_extraction.py:295-302)."""
    names, _, _ = read_output_bin(ks_file)
    read_names = [n for n in names for _ in (0, 1)]
    strands = [0, 1] * len(names)
    return read_names, strands
