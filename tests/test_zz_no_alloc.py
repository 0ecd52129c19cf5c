"""fr_knn_exact_noalloc (csrc/knn_static.cu): the exact kNN scan with its scratch in a static __device__ array.

Collected LAST on purpose.  The GPU test calls the entry point through the C ABI
and compares it with fr_knn mode 1 -- the existing entry point for the same answer -- on two calls in a row that
differ in every dimension, so the second call finds the first call's keys in the shared scratch.  Tolerance: the
one tests/test_gpu_parity.py uses for kNN (distances within 1e-5, indices equal except inside that tolerance).
"""
import ctypes as C
import re
from os.path import abspath, dirname, join

import numpy as np
import pytest

ROOT = dirname(dirname(abspath(__file__)))
KNN_ATOL = 1e-5
MAX_KEYS = 1 << 21


def test_entry_point_is_declared_exported_and_calls_no_allocator():
    from fedrann_b200 import _lib
    assert "fr_knn_exact_noalloc" in _lib.SYMBOLS
    assert hasattr(C.CDLL(_lib.LIB_PATH), "fr_knn_exact_noalloc")
    hdr = open(join(ROOT, "include", "fedrann_b200.h")).read()
    assert "int fr_knn_exact_noalloc(" in hdr and "no counterpart in the reference" in hdr.lower()
    src = open(join(ROOT, "fedrann_b200", "csrc", "knn_static.cu")).read()
    code = re.sub(r"//[^\n]*", "", src)                        # comments may name what the code must not call
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "fr_ensure", "cub::", "thrust::"):
        assert banned not in code, banned
    assert re.search(r"static\s+__device__\s+__align__\(256\)\s+unsigned long long\s+fr_noalloc_keys\[", code)
    assert "cudaGetSymbolAddress(&p, fr_noalloc_keys)" in code


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _both(eng, X, q0, nq, k):
    import torch
    En, Eh, norms = eng.normalize(X)
    ref_idx, ref_dist = eng.knn(En, Eh, norms, k, q0=q0, nq=nq, mode=1)
    idx = torch.full((nq, k), -7, dtype=torch.int32, device="cuda")
    dist = torch.full((nq, k), -7.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    N, D = En.shape
    rc = eng.L.fr_knn_exact_noalloc(eng.h, _ptr(En), _ptr(norms), N, D, q0, nq, k, _ptr(idx), _ptr(dist),
                                    C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, eng.L.fr_last_error(eng.h)
    torch.cuda.synchronize()
    return (idx.cpu().numpy(), dist.cpu().numpy()), (ref_idx.cpu().numpy(), ref_dist.cpu().numpy())


def _assert_same(got, ref):
    (gi, gd), (ri, rd) = got, ref
    assert gi.shape == ri.shape and gd.shape == rd.shape
    assert np.allclose(gd, rd, atol=KNN_ATOL, rtol=0)
    assert ((gi == ri) | (np.abs(gd - rd) <= KNN_ATOL)).all()
    assert (gi == ri).mean() > 0.999
    assert (np.diff(gd, axis=1) >= 0).all()


@pytest.mark.gpu
def test_noalloc_knn_equals_fr_knn_exact_on_two_calls_in_a_row():
    import torch
    from fedrann_b200.engine import Engine
    eng = Engine(0)
    try:
        g = torch.Generator(device="cpu").manual_seed(11)
        # call 1: 256 queries x 3000 rows = 768 000 keys
        X1 = torch.randn(3000, 96, generator=g)
        X1[13] = 0                                             # zero row: distance 1 from everything
        X1[21] = X1[20]                                        # exact tie: lower index first
        got1, ref1 = _both(eng, X1.cuda(), 100, 256, 20)
        _assert_same(got1, ref1)
        # call 2, at once, another shape: 700 x 700 = 490 000 keys over the keys call 1 left behind
        X2 = torch.randn(700, 33, generator=g)
        got2, ref2 = _both(eng, X2.cuda(), 0, 700, 128)
        _assert_same(got2, ref2)
        # and call 1 again: same answer as the first time
        got3, _ = _both(eng, X1.cuda(), 100, 256, 20)
        assert np.array_equal(got3[0], got1[0]) and np.array_equal(got3[1], got1[1])
        # beyond the static scratch: a clear error, nothing launched
        En, Eh, norms = eng.normalize(torch.randn(4096, 16, generator=g).cuda())
        idx = torch.empty((1024, 4), dtype=torch.int32, device="cuda")
        dist = torch.empty((1024, 4), dtype=torch.float32, device="cuda")
        assert 1024 * 4096 > MAX_KEYS
        before = eng.L.fr_launch_count(eng.h)
        rc = eng.L.fr_knn_exact_noalloc(eng.h, _ptr(En), _ptr(norms), 4096, 16, 0, 1024, 4, _ptr(idx), _ptr(dist),
                                        C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == -2 and b"static scratch" in eng.L.fr_last_error(eng.h)
        assert eng.L.fr_launch_count(eng.h) == before
    finally:
        eng.close()
