"""ctypes/numpy face of oracle/fedrann_oracle.c plus helpers to drive the
compiled reference binary (oracle/_ref/kmer_searcher).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Reference citations are in
fedrann_oracle.c; the Python-side restatements cite theirs inline.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import struct
import subprocess
from os.path import abspath, dirname, isfile, join

import numpy as np

_HERE = dirname(abspath(__file__))
_LIB_PATH = join(_HERE, "liboracle.so")
REF_BINARY = join(_HERE, "_ref", "kmer_searcher")


def build(force: bool = False) -> None:
    """Compile liboracle.so (and oracle/_ref when /root/reference exists)."""
    if force or not isfile(_LIB_PATH) or (
        os.path.isdir("/root/reference") and not isfile(REF_BINARY)
    ):
        subprocess.run(["make", "-C", _HERE, "all"], check=True, capture_output=True)


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_library_load.restype = C.c_void_p
        L.orc_library_load.argtypes = [C.c_char_p, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_map_free.argtypes = [C.c_void_p]
        L.orc_map_size.restype = C.c_uint64
        L.orc_map_size.argtypes = [C.c_void_p]
        L.orc_kmer_to_int.restype = C.c_uint64
        L.orc_kmer_to_int.argtypes = [C.c_char_p, C.c_int]
        L.orc_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_project.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
        L.orc_normalize.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_knn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                              C.c_int, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Library:
    """k-mer library: code -> first-occurrence index (kmer_searcher.cpp:259-273)."""

    def __init__(self, text: bytes, k: int):
        self.k = k
        ntok_max = len(text) // 2 + 1
        codes = np.zeros(ntok_max, dtype=np.uint64)
        n = C.c_uint64(0)
        self._h = lib().orc_library_load(text, len(text), k, _p(codes), C.byref(n))
        self.n = int(n.value)
        self.codes = codes[: self.n].copy()

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_map_free(self._h)
            self._h = None


def extract(library: Library, bases: np.ndarray, offsets: np.ndarray):
    """-> (indptr int64[R+1], indices int32[nnz] ascending per row, docfreq uint64[F])."""
    bases = np.ascontiguousarray(bases, dtype=np.uint8)
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    R = len(offsets) - 1
    indptr = np.zeros(R + 1, dtype=np.int64)
    docfreq = np.zeros(max(library.n, 1), dtype=np.uint64)
    if bases.size == 0:
        bases = np.zeros(1, dtype=np.uint8)
    lib().orc_extract(library._h, _p(bases), _p(offsets), R, library.k, _p(indptr), None, _p(docfreq))
    indices = np.zeros(max(int(indptr[-1]), 1), dtype=np.int32)
    lib().orc_extract(library._h, _p(bases), _p(offsets), R, library.k, _p(indptr), _p(indices), None)
    return indptr, indices[: int(indptr[-1])], docfreq[: library.n]


def project(indptr, indices, P_csr, nk: int) -> np.ndarray:
    """E (2R x D) fp32; P_csr = scipy CSR (F x D) fp32 (feature_extraction.py:204-213)."""
    P_csr = P_csr.tocsr()
    pi = np.ascontiguousarray(P_csr.indptr, dtype=np.int64)
    pj = np.ascontiguousarray(P_csr.indices, dtype=np.int32)
    pv = np.ascontiguousarray(P_csr.data, dtype=np.float32)
    R = len(indptr) - 1
    D = P_csr.shape[1]
    E = np.zeros((2 * R, D), dtype=np.float32)
    indptr = np.ascontiguousarray(indptr, dtype=np.int64)
    indices = np.ascontiguousarray(indices, dtype=np.int32)
    if indices.size == 0:
        indices = np.zeros(1, dtype=np.int32)
    lib().orc_project(_p(indptr), _p(indices), R, _p(pi), _p(pj), _p(pv), nk, D, _p(E))
    return E


def normalize(X: np.ndarray):
    X = np.ascontiguousarray(X, dtype=np.float32)
    Xn = np.zeros_like(X)
    norms = np.zeros(X.shape[0], dtype=np.float32)
    lib().orc_normalize(_p(X), X.shape[0], X.shape[1], _p(Xn), _p(norms))
    return Xn, norms


def knn(X: np.ndarray, k: int, q0: int = 0, q1: int | None = None):
    """Exact cosine kNN of rows [q0,q1) against all rows -> (idx int32, dist fp32)."""
    Xn, norms = normalize(X)
    N, D = Xn.shape
    q1 = N if q1 is None else q1
    idx = np.zeros((q1 - q0, k), dtype=np.int32)
    dist = np.zeros((q1 - q0, k), dtype=np.float32)
    lib().orc_knn(_p(Xn), _p(norms), N, D, q0, q1, k, _p(idx), _p(dist))
    return idx, dist


# ---------------------------------------------------------------------------
# precompute.py:44-115 restated (IDF x very sparse random projection)
# ---------------------------------------------------------------------------
def read_counter_file(path: str) -> np.ndarray:
    """precompute.py:44-55 -- '>count' header lines followed by one k-mer line;
    returns the counts of the forward library in file order."""
    counts = []
    count = None
    with open(path, "r") as f:
        for line in f:
            if line.startswith(">"):
                count = int(line.strip()[1:])
            else:
                counts.append(count)
    return np.asarray(counts, dtype=np.uint64)


def precompute_matrix(n_components: int, counts: np.ndarray, n_features: int, seed: int = 2094):
    """precompute.py:58-115 with the arithmetic of the PINNED numpy 1.26.4
    (requirements.txt:11): np.float64 scalar x float32 array stays float32, so
    P[i,j] = fl32( fl32(sqrt(1/density)/sqrt(D)) * sign ) * idf32[i] in fp32.
    Returns scipy CSR (n_features x n_components) float32."""
    from scipy.sparse import csr_matrix

    nk = n_features // 2
    result = np.zeros(n_features, dtype=np.uint64)
    result[:nk] = counts[:nk]          # :52  i -> count
    result[nk:2 * nk] = counts[:nk]    # :53-54  i + kmer_count -> same count
    idf = np.log(n_features / (result + 1e-12)).astype(np.float32)  # :77
    density = 1 / math.sqrt(n_features)  # :81
    rng = np.random.default_rng(seed)  # :86
    indices = []
    indptr = [0]
    offset = 0
    for _ in range(n_components):  # :90-96
        n_nonzero_i = rng.binomial(n_features, density)
        indices_i = rng.choice(n_features, n_nonzero_i, replace=False)
        indices.append(indices_i)
        offset += n_nonzero_i
        indptr.append(offset)
    indices = np.concatenate(indices) if indices else np.zeros(0, dtype=np.int64)
    data = rng.binomial(1, 0.5, size=np.size(indices)) * 2 - 1  # :101
    scale32 = np.float32(np.sqrt(1 / density) / np.sqrt(n_components))  # :107 under numpy 1.26
    vals = data.astype(np.float32) * scale32
    comp = csr_matrix((vals, indices, np.asarray(indptr)), shape=(n_components, n_features),
                      dtype=np.float32)
    Pt = comp.T.tocsr()  # (F x D), rows = features
    Pt.sort_indices()
    row_of = np.repeat(np.arange(n_features), np.diff(Pt.indptr))
    Pt.data = (Pt.data * idf[row_of]).astype(np.float32)  # :113, fp32 x fp32
    return Pt


# ---------------------------------------------------------------------------
# output.bin / kmer_frequency.bin codecs (kmer_searcher.cpp:98-130, 203-229;
# reader feature_extraction.py:108-140)
# ---------------------------------------------------------------------------
def read_output_bin(path: str):
    """-> list of (name, sorted list of indices) in file order."""
    recs = []
    with open(path, "rb") as f:
        magic, version, _res, total = struct.unpack("<4sB3sQ", f.read(16))
        assert magic == b"KMER" and version == 1
        for _ in range(total):
            (id_len,) = struct.unpack("<H", f.read(2))
            raw = f.read(id_len)
            try:                                        # feature_extraction.py:121-124
                name = raw.decode("utf-8")
            except UnicodeDecodeError:
                name = "".join(chr(b) if b < 128 else "_" for b in raw)
            (n,) = struct.unpack("<I", f.read(4))
            idx = np.frombuffer(f.read(8 * n), dtype="<u8")
            recs.append((name, sorted(int(x) for x in idx)))
    return recs


def read_frequency_bin(path: str):
    a = np.fromfile(path, dtype="<u8").reshape(-1, 2)
    return {int(i): int(c) for i, c in a}


def run_reference_binary(library_path: str, fastx_path: str, out_dir: str, k: int, threads: int = 1):
    """Runs oracle/_ref/kmer_searcher (the unmodified reference) and returns
    (records dict name -> sorted indices, record order list, freq dict)."""
    if not isfile(REF_BINARY):
        raise FileNotFoundError(REF_BINARY)
    os.makedirs(dirname(out_dir.rstrip("/")) or ".", exist_ok=True)
    subprocess.run([REF_BINARY, library_path, fastx_path, out_dir, str(k), str(threads)], check=True)
    recs = read_output_bin(join(out_dir, "output.bin"))
    freq = read_frequency_bin(join(out_dir, "kmer_frequency.bin"))
    return recs, freq


def parse_fastx(text: bytes):
    """kmer_searcher.cpp:153-201 restated: -> list of (name, sequence bytes).
    FASTA: name = header up to the first space/tab, multi-line sequences are
    concatenated, records with an empty name are dropped.  FASTQ (first line
    starts with '@'): name = whole header after '@', sequence = next line,
    the following two lines are skipped.  Empty lines are skipped."""
    lines = text.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()  # getline does not yield a trailing empty record
    out = []
    if not lines:
        return out
    is_fastq = lines[0][:1] == b"@"
    i = 0
    cur_id, cur_seq = b"", []
    while i < len(lines):
        line = lines[i]
        i += 1
        if not line:
            continue
        if not is_fastq:
            if line[:1] == b">":
                if cur_id:
                    out.append((cur_id.decode(), b"".join(cur_seq)))
                hdr = line[1:]
                for j, ch in enumerate(hdr):
                    if ch in (0x20, 0x09):
                        hdr = hdr[:j]
                        break
                cur_id, cur_seq = hdr, []
            else:
                cur_seq.append(line)
        else:
            if line[:1] == b"@":
                cur_id = line[1:]
                seq = lines[i] if i < len(lines) else b""
                i += 3
                out.append((cur_id.decode(), seq))
    if not is_fastq and cur_id:
        out.append((cur_id.decode(), b"".join(cur_seq)))
    return out


# ---------------------------------------------------------------------------
# multi-core legs for bench.py's CPU baseline (same arithmetic, all host cores)
# ---------------------------------------------------------------------------
def project_threads(indptr, indices, P_csr, nk: int, threads: int) -> np.ndarray:
    """project() over `threads` OS threads (ctypes releases the GIL); rows are
    independent, so the result is bit-identical to project()."""
    from concurrent.futures import ThreadPoolExecutor
    indptr = np.ascontiguousarray(indptr, dtype=np.int64)
    R = len(indptr) - 1
    if threads <= 1 or R < 2 * threads:
        return project(indptr, indices, P_csr, nk)
    P_csr = P_csr.tocsr()
    bounds = np.linspace(0, R, threads + 1).astype(np.int64)

    def work(t):
        a, b = int(bounds[t]), int(bounds[t + 1])
        ip = indptr[a:b + 1] - indptr[a]
        return project(ip, indices[indptr[a]:indptr[b]], P_csr, nk)

    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(work, range(threads)))
    return np.concatenate(parts, axis=0)


def knn_blas(Xn: np.ndarray, norms: np.ndarray, q0: int, q1: int, k: int, block: int = 2048):
    """Exact cosine kNN of unit rows [q0,q1) against all rows with BLAS sgemm
    (all host threads) + partial sort; same definition as knn(): dist = 1 - dot,
    zero-norm rows at distance 1, ties by lower index."""
    N = Xn.shape[0]
    idx = np.zeros((q1 - q0, k), dtype=np.int32)
    dist = np.zeros((q1 - q0, k), dtype=np.float32)
    zero_db = norms == 0
    for a in range(q0, q1, block):
        b = min(q1, a + block)
        d = 1.0 - Xn[a:b] @ Xn.T
        d[:, zero_db] = 1.0
        d[norms[a:b] == 0, :] = 1.0
        kk = min(k, N)
        part = np.argpartition(d, kk - 1, axis=1)[:, :kk]
        pd = np.take_along_axis(d, part, axis=1)
        order = np.lexsort((part, pd), axis=1)
        idx[a - q0:b - q0, :kk] = np.take_along_axis(part, order, axis=1)
        dist[a - q0:b - q0, :kk] = np.take_along_axis(pd, order, axis=1)
    return idx, dist


# ---------------------------------------------------------------------------
# N1 / N2: the callers either side of the path
# ---------------------------------------------------------------------------
def _splitmix64(x):
    x = x.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x += np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return x


def count_kmers(bases: np.ndarray, offsets: np.ndarray, k: int, fraction: float, min_mult: int, seed: int):
    """What `jellyfish count -C` + `jellyfish dump -L min_mult` + a sampler produce
    (fedrann/count_kmers.py:76-123): canonical k-mers (min of k-mer and reverse complement,
    A0 C1 G2 T3) of every read, windows with a non-ACGT base skipped, multiplicity >= min_mult.
    The reference's awk rand() sampler is unpinned; the build's sampler is
    keep(c) <=> fmix32(lo(c) ^ hi(c)*0x9E3779B1 ^ seedmix(seed)) < (fraction * 2^64) >> 32, everything at
    fraction 1 (csrc/common.cuh fr_sample_keep).  -> {code: count} (plain Python loop:
    small inputs only)."""
    code = {65: 0, 67: 1, 71: 2, 84: 3, 97: 0, 99: 1, 103: 2, 116: 3}
    mask = (1 << (2 * k)) - 1
    counts = {}
    for r in range(len(offsets) - 1):
        fw = rc = 0
        run = 0
        for b in bases[offsets[r]:offsets[r + 1]].tolist():
            c = code.get(b)
            if c is None:
                run = 0
                continue
            fw = ((fw << 2) | c) & mask
            rc = (rc >> 2) | ((3 - c) << (2 * (k - 1)))
            run += 1
            if run >= k:
                can = min(fw, rc)
                counts[can] = counts.get(can, 0) + 1
    thr = min(int(fraction * 2.0 ** 64), 2 ** 64 - 1)
    sd = seed & (2 ** 64 - 1)
    seedmix = (sd & 0xFFFFFFFF) ^ (((sd >> 32) * 0x85EBCA77) & 0xFFFFFFFF)

    def keep(c):        # fr_sample_keep (csrc/common.cuh), plain integers
        if thr == 2 ** 64 - 1:
            return True
        x = (c & 0xFFFFFFFF) ^ (((c >> 32) * 0x9E3779B1) & 0xFFFFFFFF) ^ seedmix
        x ^= x >> 16
        x = (x * 0x85EBCA6B) & 0xFFFFFFFF
        x ^= x >> 13
        x = (x * 0xC2B2AE35) & 0xFFFFFFFF
        x ^= x >> 16
        return x < (thr >> 32)

    return {c: n for c, n in counts.items() if n >= min_mult and keep(c)}


def overlaps_tsv(neighbor_matrix: np.ndarray, neighbor_distances: np.ndarray, read_names, strands) -> str:
    """get_output_dataframe (fedrann/__main__.py:261-300) followed by
    df.to_csv(path, sep="\\t", index=False) (:385), returned as text."""
    import io
    import pandas as pd
    query_names, target_names, ranks, distances, qo, to = [], [], [], [], [], []
    for query_index in range(neighbor_matrix.shape[0]):
        for rank, target_index in enumerate(neighbor_matrix[query_index]):
            if target_index == query_index:
                continue
            query_names.append(read_names[query_index])
            qo.append(["+", "-"][strands[query_index]])
            target_names.append(read_names[target_index])
            to.append(["+", "-"][strands[target_index]])
            ranks.append(rank)
            distances.append(neighbor_distances[query_index][rank])
    df = pd.DataFrame({"query_name": query_names, "query_orientation": qo, "target_name": target_names,
                       "target_orientation": to, "neighbor_rank": ranks, "distance": distances})
    if len(df) == 0:
        df = df.astype({"neighbor_rank": "int64", "distance": "float32"})
    buf = io.StringIO()
    df.to_csv(buf, sep="\t", index=False)
    return buf.getvalue()
