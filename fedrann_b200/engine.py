"""Device engine: one `Engine` per GPU (per rank) over the C ABI.

PyTorch is used for device memory, streams and (in parallel.py) NCCL only;
every computation is a kernel of libfedrann_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import FedrannError, check


def _ptr(t):
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# ---------------------------------------------------------------------------
# host-side parsers / writers (native, shared with the kmer_searcher tool)
# ---------------------------------------------------------------------------
def parse_library(text: bytes, k: int) -> np.ndarray:
    """kmer_searcher.cpp:259-273 -> uint64 codes, index = first-occurrence rank."""
    L = _lib.lib()
    cap = len(text) // 2 + 1
    codes = np.zeros(cap, dtype=np.uint64)
    n = C.c_int64(0)
    check(None, L.fr_library_parse(text, len(text), k, codes.ctypes.data_as(C.c_void_p), cap, C.byref(n)),
          "fr_library_parse")
    return codes[: n.value].copy()


def parse_fastx(text: bytes):
    """kmer_searcher.cpp:153-201 -> (bases uint8, offsets int64[R+1], names list)."""
    L = _lib.lib()
    nr, nb, nl = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    check(None, L.fr_fastx_parse(text, len(text), None, None, None, C.byref(nr), C.byref(nb), C.byref(nl)),
          "fr_fastx_parse")
    bases = np.zeros(max(nb.value, 1), dtype=np.uint8)
    offsets = np.zeros(nr.value + 1, dtype=np.int64)
    names = np.zeros(max(nl.value, 1), dtype=np.uint8)
    check(None, L.fr_fastx_parse(text, len(text), bases.ctypes.data_as(C.c_void_p),
                                 offsets.ctypes.data_as(C.c_void_p), names.ctypes.data_as(C.c_void_p),
                                 C.byref(nr), C.byref(nb), C.byref(nl)), "fr_fastx_parse")
    name_list = names[: nl.value].tobytes().decode("utf-8", "replace").split("\n")[:-1] if nl.value else []
    return bases[: nb.value], offsets, name_list


def write_output_bin(path: str, names, indptr: np.ndarray, indices: np.ndarray) -> None:
    """output.bin in the layout of kmer_searcher.cpp:98-130 (input order, ascending indices)."""
    L = _lib.lib()
    enc = [n.encode() for n in names]
    noff = np.zeros(len(enc) + 1, dtype=np.int64)
    np.cumsum([len(e) for e in enc], out=noff[1:])
    blob = b"".join(enc) or b"\0"
    indptr = np.ascontiguousarray(indptr, dtype=np.int64)
    indices = np.ascontiguousarray(indices, dtype=np.int32)
    if indices.size == 0:
        indices = np.zeros(1, dtype=np.int32)
    rc = L.fr_write_output_bin(path.encode(), C.c_char_p(blob), noff.ctypes.data_as(C.c_void_p), len(enc),
                               indptr.ctypes.data_as(C.c_void_p), indices.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise FedrannError("ID contains non-ASCII characters" if rc == -7 else
                           f"cannot open {path}" if rc == -6 else f"fr_write_output_bin: {rc}")


def write_frequency_bin(path: str, docfreq: np.ndarray) -> None:
    L = _lib.lib()
    docfreq = np.ascontiguousarray(docfreq, dtype=np.uint32)
    check(None, L.fr_write_frequency_bin(path.encode(), docfreq.ctypes.data_as(C.c_void_p), len(docfreq)),
          "fr_write_frequency_bin")


def write_overlaps_tsv(path: str, names, idx: np.ndarray, dist: np.ndarray, *, row0: int = 0,
                       n_rows_total: int | None = None, append: bool = False) -> int:
    """overlaps.tsv exactly as get_output_dataframe + to_csv write it (__main__.py:261-300, :385).
    names: one name per READ (rows 2r / 2r+1 are its + / - orientation).  Returns the line count."""
    L = _lib.lib()
    enc = [n.encode() for n in names]
    noff = np.zeros(len(enc) + 1, dtype=np.int64)
    np.cumsum([len(e) for e in enc], out=noff[1:])
    blob = b"".join(enc) or b"\0"
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    dist = np.ascontiguousarray(dist, dtype=np.float32)
    n_rows, k = idx.shape if idx.ndim == 2 else (0, 0)
    total = 2 * len(enc) if n_rows_total is None else n_rows_total
    lines = C.c_int64(0)
    rc = L.fr_write_overlaps_tsv(path.encode(), 1 if append else 0, C.c_char_p(blob),
                                 noff.ctypes.data_as(C.c_void_p), total, row0, n_rows, k,
                                 idx.ctypes.data_as(C.c_void_p), dist.ctypes.data_as(C.c_void_p), C.byref(lines))
    if rc != 0:
        raise FedrannError(f"fr_write_overlaps_tsv: {rc}")
    return lines.value


# ---------------------------------------------------------------------------
class Engine:
    def __init__(self, device: int | None = None):
        if not torch.cuda.is_available():
            raise FedrannError("fedrann_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.L = _lib.lib()
        self.device = torch.cuda.current_device() if device is None else int(device)
        torch.cuda.set_device(self.device)
        h = C.c_void_p()
        rc = self.L.fr_create(self.device, C.byref(h))
        if rc != 0:
            raise FedrannError(f"fr_create failed ({rc}): an sm_100 GPU is required")
        self.h = h
        self.k = None
        self.n_features = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.fr_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self) -> int:
        return int(self.L.fr_launch_count(self.h))

    def profile(self, on: bool) -> None:
        """Bracket every kernel launch with a CUDA event pair (see kernel_times)."""
        check(self.h, self.L.fr_profile_enable(self.h, 1 if on else 0), "fr_profile_enable")

    def kernel_times(self) -> dict:
        """{kernel: (launches, total_ms)} since profile(True); clears the records."""
        buf = C.create_string_buffer(1 << 16)
        n = C.c_int64(0)
        check(self.h, self.L.fr_profile_read(self.h, buf, len(buf), C.byref(n)), "fr_profile_read")
        out = {}
        for line in buf.value.decode().splitlines():
            name, cnt, ms = line.split()
            out[name] = (int(cnt), float(ms))
        return out

    def dev(self, a: np.ndarray, dtype=None) -> torch.Tensor:
        a = np.ascontiguousarray(a)
        if not a.flags.writeable:           # torch.from_numpy refuses to alias read-only memory silently
            a = a.copy()
        t = torch.from_numpy(a)
        if dtype is not None:
            t = t.to(dtype)
        return t.to(f"cuda:{self.device}", non_blocking=True)

    # -- N1 --------------------------------------------------------------
    def count_kmers(self, bases: torch.Tensor, offsets: torch.Tensor, k: int, fraction: float, seed: int,
                    min_multiplicity: int = 1, capacity: int | None = None):
        """Canonical k-mer counts of the sampled k-mers (replaces jellyfish count/dump + the awk
        sampler, count_kmers.py:76-123).  -> (codes uint64 numpy, counts uint32 numpy), unordered."""
        thr = min(int(fraction * 2.0 ** 64), 2 ** 64 - 1)
        n_pos = int(bases.numel())
        if capacity is None:
            capacity = int(min(4 ** min(k, 31), 2.5 * fraction * n_pos + 65536)) + 1024
        R = offsets.numel() - 1
        while True:
            check(self.h, self.L.fr_kmer_count_begin(self.h, k, capacity, _stream()), "fr_kmer_count_begin")
            distinct = C.c_int64(0)
            rc = self.L.fr_kmer_count_add(self.h, _ptr(bases), _ptr(offsets), R, thr, seed & (2 ** 64 - 1),
                                          C.byref(distinct), _stream())
            if rc == -4 and capacity < (1 << 33):      # table full: grow and recount
                capacity *= 4
                continue
            check(self.h, rc, "fr_kmer_count_add")
            break
        n = C.c_int64(0)
        cap = max(distinct.value, 1)
        codes = torch.empty(cap, dtype=torch.int64, device=bases.device)
        counts = torch.empty(cap, dtype=torch.int32, device=bases.device)
        check(self.h, self.L.fr_kmer_count_export(self.h, min_multiplicity, _ptr(codes), _ptr(counts), cap,
                                                  C.byref(n), _stream()), "fr_kmer_count_export")
        m = min(n.value, cap)
        return codes[:m].cpu().numpy().view(np.uint64), counts[:m].cpu().numpy().view(np.uint32)

    # -- K2 --------------------------------------------------------------
    def build_table(self, codes, k: int, sampler=None) -> None:
        """codes: uint64 numpy (or int64-viewed cuda tensor), feature index -> k-mer code.
        sampler = (seed, fraction) declares that the library was sampled by N1's rule
        (fr_sample_keep(canonical, seed, fraction * 2^64), csrc/common.cuh); the claim is verified on the device and
        `self.sampled` tells whether extraction will use it."""
        if isinstance(codes, np.ndarray):
            codes = self.dev(np.ascontiguousarray(codes, dtype=np.uint64).view(np.int64))
        self._codes = codes
        check(self.h, self.L.fr_table_build(self.h, _ptr(codes), codes.numel(), k, _stream()), "fr_table_build")
        self.k = k
        self.n_features = codes.numel()
        self.sampled = False
        if sampler is not None:
            seed, fraction = sampler
            thr = min(int(fraction * 2.0 ** 64), 2 ** 64 - 1)
            rc = self.L.fr_table_set_sampler(self.h, seed & (2 ** 64 - 1), thr, _stream())
            if rc not in (0, -8):                      # -8 = FR_ERR_UNVERIFIED: not such a sample, generic path
                check(self.h, rc, "fr_table_set_sampler")
            self.sampled = rc == 0

    # -- K1 + K3 ---------------------------------------------------------
    def extract(self, bases: torch.Tensor, offsets: torch.Tensor, docfreq: torch.Tensor | None = None):
        """-> (indptr int64[R+1], indices int32[nnz]) on the device; CSR rows in
        input order, columns ascending.  docfreq (int32/uint32 [F]) is added to."""
        assert bases.dtype == torch.uint8 and offsets.dtype == torch.int64
        R = offsets.numel() - 1
        indptr = torch.empty(R + 1, dtype=torch.int64, device=bases.device)
        nnz = C.c_int64(0)
        check(self.h, self.L.fr_extract_count(self.h, _ptr(bases), bases.numel(), _ptr(offsets), R, _ptr(indptr),
                                              _ptr(docfreq), C.byref(nnz), _stream()), "fr_extract_count")
        indices = torch.empty(max(nnz.value, 1), dtype=torch.int32, device=bases.device)
        check(self.h, self.L.fr_extract_fill(self.h, _ptr(indices), _stream()), "fr_extract_fill")
        return indptr, indices[: nnz.value]

    def extract_async(self, bases: torch.Tensor, offsets: torch.Tensor, indices: torch.Tensor, status: torch.Tensor):
        """fr_extract_async: no host synchronisation.  indices: int32 buffer (its size is the capacity);
        status: int64[2] on the device, zeroed by the caller -- after the stream has drained
        status[0] != 0 means the rows did not fit and nothing was written, status[1] = non-zeros.
        -> indptr int64[R+1]."""
        assert bases.dtype == torch.uint8 and offsets.dtype == torch.int64
        assert indices.dtype == torch.int32 and status.dtype == torch.int64 and status.numel() == 2
        R = offsets.numel() - 1
        indptr = torch.empty(R + 1, dtype=torch.int64, device=bases.device)
        check(self.h, self.L.fr_extract_async(self.h, _ptr(bases), bases.numel(), _ptr(offsets), R, _ptr(indptr),
                                              _ptr(indices), indices.numel(), _ptr(status), _stream()),
              "fr_extract_async")
        return indptr

    # -- K4 --------------------------------------------------------------
    def upload_projection(self, P_csr):
        """scipy CSR (F x D) fp32 -> device triple."""
        P_csr = P_csr.tocsr()
        return (self.dev(P_csr.indptr.astype(np.int64)), self.dev(P_csr.indices.astype(np.int32)),
                self.dev(P_csr.data.astype(np.float32)), int(P_csr.shape[1]))

    def project(self, indptr, indices, P_dev, nk: int, out: torch.Tensor | None = None) -> torch.Tensor:
        p_indptr, p_indices, p_data, D = P_dev
        R = indptr.numel() - 1
        E = out if out is not None else torch.empty((2 * R, D), dtype=torch.float32, device=indptr.device)
        if indices.numel() == 0:
            indices = torch.zeros(1, dtype=torch.int32, device=indptr.device)
        if p_indices.numel() == 0:
            p_indices = torch.zeros(1, dtype=torch.int32, device=indptr.device)
            p_data = torch.zeros(1, dtype=torch.float32, device=indptr.device)
        check(self.h, self.L.fr_project(self.h, _ptr(indptr), _ptr(indices), R, _ptr(p_indptr), _ptr(p_indices),
                                        _ptr(p_data), nk, D, _ptr(E), _stream()), "fr_project")
        return E

    # -- K5 --------------------------------------------------------------
    def normalize(self, E: torch.Tensor, out=None):
        """out: (En, Eh, norms) row slices to write into (contiguous, same row count as E)."""
        N, D = E.shape
        Dp = (D + 63) // 64 * 64
        if out is not None:
            En, Eh, norms = out
            assert En.shape == E.shape and Eh.shape == (N, Dp) and norms.shape == (N,)
            assert En.is_contiguous() and Eh.is_contiguous() and norms.is_contiguous()
        else:
            En = torch.empty_like(E)
            Eh = torch.empty((N, Dp), dtype=torch.float16, device=E.device)
            norms = torch.empty(N, dtype=torch.float32, device=E.device)
        if N == 0:
            return En, Eh, norms
        check(self.h, self.L.fr_normalize(self.h, _ptr(E), N, D, Dp, _ptr(En), _ptr(Eh), _ptr(norms), _stream()),
              "fr_normalize")
        return En, Eh, norms

    # -- K6 + K7 ---------------------------------------------------------
    def knn(self, En, Eh, norms, k: int, q0: int = 0, nq: int | None = None, mode: int = 0, rows_ready=None,
            out=None):
        """rows_ready: a torch.cuda.Event after which En is complete (it may still be arriving from
        an all-gather on another stream); the tensor-core pass does not wait for it.
        out: (idx int32 [nq, k], dist fp32 [nq, k]) contiguous tensors to write into -- device tensors,
        or PINNED host tensors (the kernels then write the neighbour lists straight into host memory
        over PCIe while they run, instead of a copy afterwards)."""
        N, D = En.shape
        if rows_ready is not None:
            check(self.h, self.L.fr_knn_rows_ready_after(self.h, C.c_void_p(rows_ready.cuda_event)),
                  "fr_knn_rows_ready_after")
        nq = N - q0 if nq is None else nq
        idx, dist = self._knn_out(out, nq, k, En.device)
        Dp = Eh.shape[1] if Eh is not None else D
        check(self.h, self.L.fr_knn(self.h, _ptr(En), _ptr(Eh), _ptr(norms), N, D, Dp, q0, nq, k, mode,
                                    _ptr(idx), _ptr(dist), _stream()), "fr_knn")
        return idx, dist

    def knn_inc_begin(self, En, Eh, norms, k: int, bounds):
        """Incremental all-vs-all search: En/Eh/norms are the full-size arrays, bounds the cumulative
        row counts of the phases (fr_knn_inc_begin).  Raises FedrannError for shapes the tensor-core
        path does not cover."""
        N, D = En.shape
        b = np.ascontiguousarray(np.asarray(bounds, dtype=np.int64))
        self._inc = (N, k, En.device)
        check(self.h, self.L.fr_knn_inc_begin(self.h, _ptr(En), _ptr(Eh), _ptr(norms), N, D, Eh.shape[1], k,
                                              len(b) - 1, b.ctypes.data_as(C.c_void_p), _stream()), "fr_knn_inc_begin")

    def knn_inc_phase(self, j: int):
        check(self.h, self.L.fr_knn_inc_phase(self.h, j, _stream()), "fr_knn_inc_phase")

    @staticmethod
    def _knn_out(out, nq, k, dev):
        if out is None:
            return (torch.empty((nq, k), dtype=torch.int32, device=dev),
                    torch.empty((nq, k), dtype=torch.float32, device=dev))
        idx, dist = out
        assert idx.dtype == torch.int32 and dist.dtype == torch.float32
        assert tuple(idx.shape) == (nq, k) and tuple(dist.shape) == (nq, k)
        assert idx.is_contiguous() and dist.is_contiguous()
        assert (idx.is_cuda or idx.is_pinned()) and (dist.is_cuda or dist.is_pinned())
        return idx, dist

    def knn_inc_finish(self, out=None):
        N, k, dev = self._inc
        idx, dist = self._knn_out(out, N, k, dev)
        check(self.h, self.L.fr_knn_inc_finish(self.h, _ptr(idx), _ptr(dist), _stream()), "fr_knn_inc_finish")
        return idx, dist

    @staticmethod
    def knn_inc_supported(N: int, D: int, k: int) -> bool:
        """shapes fr_knn_inc_begin accepts (the others go through fr_knn, which has the exact scan)"""
        return D <= 512 and k <= 128 and N >= 2 * k and N < 2 ** 31 - 256

    @property
    def knn_last_path(self) -> str:
        """'tcgen05' or 'scan': the path the last kNN call actually took"""
        return {0: "tcgen05", 1: "scan"}.get(int(self.L.fr_knn_last_path(self.h)), "none")

    @property
    def knn_fallback_rows(self) -> int:
        return int(self.L.fr_knn_last_fallback_rows(self.h))
