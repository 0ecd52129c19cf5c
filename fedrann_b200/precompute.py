"""IDF-weighted very sparse random projection -- host-side set-up of the K4
operand.  Mirrors fedrann/precompute.py:44-115 (same function name, arguments
and return value).  It stays on the host with numpy so that the
`np.random.default_rng(2094)` binomial / choice streams are the reference's
own (D calls, negligible time).

Arithmetic follows the reference's PINNED numpy 1.26.4 (requirements.txt:11),
where `np.float64 scalar * float32 array` stays float32: every entry is
fl32(scale) * (+-1) * fl32(idf[i]) evaluated in fp32.
"""
from __future__ import annotations

import math

import numpy as np
from scipy.sparse import csr_matrix


def kmer_count_generator(filename: str, kmer_count: int):
    """precompute.py:44-55: '>count' header then a k-mer line; yields (i, count)
    and (i + kmer_count, count) for the reverse-complement feature."""
    i = 0
    count = 0
    with open(filename, "r") as f:
        for line in f:
            if line.startswith(">"):
                count = int(line.strip()[1:])
            else:
                yield i, count
                yield i + kmer_count, count
                i += 1


def read_counts(filename: str) -> np.ndarray:
    counts = []
    c = 0
    with open(filename, "r") as f:
        for line in f:
            if line.startswith(">"):
                c = int(line.strip()[1:])
            else:
                counts.append(c)
    return np.asarray(counts, dtype=np.uint64)


def precompute_matrix(n_components: int, counts: np.ndarray, n_features: int, density="auto",
                      seed: int = 2094) -> csr_matrix:
    """counts: multiplicity of each forward-library k-mer.  Returns CSR (F x D) fp32."""
    nk = n_features // 2
    per_feature = np.zeros(n_features, dtype=np.uint64)
    m = min(nk, len(counts))
    per_feature[:m] = counts[:m]
    per_feature[nk:nk + m] = counts[:m]
    idf = np.log(n_features / (per_feature + 1e-12)).astype(np.float32)      # precompute.py:77
    if density == "auto":
        dens = 1 / math.sqrt(n_features)                                     # :81
    else:
        assert isinstance(density, float) and 0 < density <= 1
        dens = density
    rng = np.random.default_rng(seed)                                        # :86
    cols, rows = [], []
    for comp in range(n_components):                                         # :90-96, same draw order
        nnz = rng.binomial(n_features, dens)
        feat = rng.choice(n_features, nnz, replace=False)
        rows.append(feat)
        cols.append(np.full(nnz, comp, dtype=np.int32))
    rows = np.concatenate(rows) if rows else np.zeros(0, dtype=np.int64)
    cols = np.concatenate(cols) if cols else np.zeros(0, dtype=np.int32)
    signs = (rng.binomial(1, 0.5, size=rows.size) * 2 - 1).astype(np.float32)  # :101
    scale32 = np.float32(np.sqrt(1 / dens) / np.sqrt(n_components))             # :107 (fp32 under numpy 1.26)
    vals = (signs * scale32) * idf[rows]                                        # :113, fp32 x fp32
    # CSR by feature row, columns ascending inside a row
    order = np.lexsort((cols, rows))
    rows, cols, vals = rows[order], cols[order], vals[order].astype(np.float32)
    indptr = np.zeros(n_features + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows, minlength=n_features), out=indptr[1:])
    return csr_matrix((vals, cols, indptr), shape=(n_features, n_components), dtype=np.float32)


def get_precompute_matrix(n_components: int, counter_file: str, n_features: int, density="auto",
                          seed: int = 2094):
    """Drop-in for fedrann.precompute.get_precompute_matrix (precompute.py:58-115)."""
    counts = read_counts(counter_file)
    return precompute_matrix(n_components, counts, n_features, density, seed), n_features
