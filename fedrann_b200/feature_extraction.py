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
    """output.bin (kmer_searcher.cpp:98-130) -> (names, indptr int64[R+1], indices int64[nnz]).
    One native pass over the file image (fr_read_output_bin); the error cases and messages are the
    reference reader's (feature_extraction.py:109-117)."""
    import ctypes as C
    from . import _lib
    L = _lib.lib()
    blob = np.fromfile(ks_file, dtype=np.uint8)
    bp = blob.ctypes.data_as(C.c_void_p)
    nr, nnz, nl, st = C.c_int64(0), C.c_int64(0), C.c_int64(0), C.c_int(0)

    def call(names, noff, indptr, indices):
        _lib.check(None, L.fr_read_output_bin(bp, blob.size, names, noff, indptr, indices, C.byref(nr), C.byref(nnz),
                                              C.byref(nl), C.byref(st)), "fr_read_output_bin")
        if st.value == 1:
            raise ValueError("文件头不完整")
        if st.value == 2:
            raise ValueError("无效的文件格式")
        if st.value == 3:
            raise ValueError(f"不支持的版本号: {nr.value}")
        if st.value == 4:
            raise ValueError("output.bin is truncated")       # the reference dies in struct.unpack here

    call(None, None, None, None)
    name_bytes = np.zeros(max(nl.value, 1), dtype=np.uint8)
    noff = np.zeros(nr.value + 1, dtype=np.int64)
    indptr = np.zeros(nr.value + 1, dtype=np.int64)
    indices = np.empty(nnz.value, dtype=np.int64)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    call(vp(name_bytes), vp(noff), vp(indptr), vp(indices))
    raw = name_bytes.tobytes()
    try:                                    # ASCII ids (the writer refuses anything else): one decode, then slices
        text = raw.decode("ascii")
        names = [text[noff[r]:noff[r + 1]] for r in range(nr.value)]
    except UnicodeDecodeError:
        names = []
        for r in range(nr.value):
            b = raw[noff[r]:noff[r + 1]]
            try:
                names.append(b.decode("utf-8"))
            except UnicodeDecodeError:
                names.append("".join(chr(c) if c < 128 else "_" for c in b))
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
    # The reference builds A_chunk through COO (feature_extraction.py:191-205), which SORTS the columns of a
    # row; a (row, column) pair listed twice would be summed there (int8 value 2).  kmer_searcher writes a set
    # per read (robin_hood::unordered_set, kmer_searcher.cpp:94), so a valid output.bin has no duplicates;
    # a file that does is refused rather than silently de-duplicated.
    rows = np.repeat(np.arange(len(names), dtype=np.int64), np.diff(indptr))
    stride = np.int64(max(int(kmer_count), 1) + 1)
    key = np.sort(rows * stride + indices)
    if len(key) > 1 and bool((key[1:] == key[:-1]).any()):
        raise ValueError("output.bin lists a feature index twice in one record (kmer_searcher never writes that)")
    rows_u = key // stride
    cols_u = (key % stride).astype(np.int32)
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
