"""ctypes binding of libfedrann_b200.so (the C ABI declared in include/fedrann_b200.h).

There is no CPU fallback: if the shared library is missing, fails to load, or
no sm_100 device is present, every entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os
from os.path import abspath, dirname, isfile, join

_HERE = dirname(abspath(__file__))
LIB_PATH = join(_HERE, "libfedrann_b200.so")

FR_OK = 0
_ERR = {-1: "FR_ERR_CUDA", -2: "FR_ERR_ARG", -3: "FR_ERR_STATE", -4: "FR_ERR_NOMEM", -5: "FR_ERR_NO_DEVICE",
        -6: "FR_ERR_IO", -7: "FR_ERR_BAD_ID", -8: "FR_ERR_UNVERIFIED"}

SYMBOLS = [
    "fr_version", "fr_create", "fr_destroy", "fr_last_error", "fr_launch_count",
    "fr_profile_enable", "fr_profile_read",
    "fr_library_parse", "fr_fastx_parse", "fr_write_output_bin", "fr_write_frequency_bin", "fr_read_output_bin",
    "fr_table_build", "fr_table_set_sampler", "fr_extract_count", "fr_extract_fill", "fr_extract_async", "fr_project", "fr_normalize",
    "fr_knn", "fr_knn_rows_ready_after", "fr_knn_last_fallback_rows", "fr_knn_last_path",
    "fr_knn_exact_noalloc",
    "fr_knn_inc_begin", "fr_knn_inc_phase", "fr_knn_inc_finish",
    "fr_kmer_count_begin", "fr_kmer_count_add", "fr_kmer_count_export", "fr_write_overlaps_tsv",
    "fr_exchange_create", "fr_exchange_connect", "fr_exchange_rows_f32", "fr_exchange_rows_f16", "fr_exchange_norms",
    "fr_exchange_pull", "fr_exchange_release", "fr_exchange_destroy",
]


class FedrannError(RuntimeError):
    pass


_lib = None


def build(verbose: bool = False) -> None:
    """Compile the CUDA library in-tree (nvcc, sm_100a)."""
    import subprocess
    r = subprocess.run(["make", "-C", join(_HERE, "csrc"), "-j8", "all"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode != 0:
        raise FedrannError("building libfedrann_b200.so failed")


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not isfile(LIB_PATH):
        raise FedrannError(
            f"{LIB_PATH} is missing: build it with `make -C fedrann_b200/csrc` "
            "(there is no CPU fallback for this path)")
    L = C.CDLL(LIB_PATH)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    L.fr_version.restype = i32
    L.fr_create.argtypes = [i32, C.POINTER(vp)]
    L.fr_destroy.argtypes = [vp]
    L.fr_destroy.restype = None
    L.fr_last_error.argtypes = [vp]
    L.fr_last_error.restype = C.c_char_p
    L.fr_launch_count.argtypes = [vp]
    L.fr_launch_count.restype = i64
    L.fr_profile_enable.argtypes = [vp, i32]
    L.fr_profile_read.argtypes = [vp, C.c_char_p, i64, C.POINTER(i64)]
    L.fr_library_parse.argtypes = [C.c_char_p, i64, i32, vp, i64, C.POINTER(i64)]
    L.fr_fastx_parse.argtypes = [C.c_char_p, i64, vp, vp, vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]
    L.fr_write_output_bin.argtypes = [C.c_char_p, vp, vp, i64, vp, vp]
    L.fr_write_frequency_bin.argtypes = [C.c_char_p, vp, i64]
    L.fr_read_output_bin.argtypes = [vp, i64, vp, vp, vp, vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64),
                                     C.POINTER(i32)]
    L.fr_table_build.argtypes = [vp, vp, i64, i32, vp]
    L.fr_extract_count.argtypes = [vp, vp, i64, vp, i64, vp, vp, C.POINTER(i64), vp]
    L.fr_extract_fill.argtypes = [vp, vp, vp]
    L.fr_extract_async.argtypes = [vp, vp, i64, vp, i64, vp, vp, i64, vp, vp]
    L.fr_project.argtypes = [vp, vp, vp, i64, vp, vp, vp, i64, i32, vp, vp]
    L.fr_normalize.argtypes = [vp, vp, i64, i32, i32, vp, vp, vp, vp]
    L.fr_knn.argtypes = [vp, vp, vp, vp, i64, i32, i32, i64, i64, i32, i32, vp, vp, vp]
    L.fr_knn_rows_ready_after.argtypes = [vp, vp]
    L.fr_knn_inc_begin.argtypes = [vp, vp, vp, vp, i64, i32, i32, i32, i32, vp, vp]
    L.fr_knn_inc_phase.argtypes = [vp, i32, vp]
    L.fr_knn_inc_finish.argtypes = [vp, vp, vp, vp]
    L.fr_knn_last_fallback_rows.argtypes = [vp]
    L.fr_knn_last_fallback_rows.restype = i64
    L.fr_knn_last_path.argtypes = [vp]
    L.fr_knn_last_path.restype = i32
    L.fr_knn_exact_noalloc.argtypes = [vp, vp, vp, i64, i32, i64, i64, i32, vp, vp, vp]
    L.fr_knn_exact_noalloc.restype = i32
    u64 = C.c_uint64
    L.fr_table_set_sampler.argtypes = [vp, u64, u64, vp]
    L.fr_kmer_count_begin.argtypes = [vp, i32, i64, vp]
    L.fr_kmer_count_add.argtypes = [vp, vp, vp, i64, u64, u64, C.POINTER(i64), vp]
    L.fr_kmer_count_export.argtypes = [vp, i32, vp, vp, i64, C.POINTER(i64), vp]
    L.fr_write_overlaps_tsv.argtypes = [C.c_char_p, i32, vp, vp, i64, i64, i64, i32, vp, vp, C.POINTER(i64)]
    L.fr_exchange_create.argtypes = [vp, i32, i32, vp, i32, i32, C.POINTER(vp), vp]
    L.fr_exchange_connect.argtypes = [vp, vp]
    for fn in (L.fr_exchange_rows_f32, L.fr_exchange_rows_f16, L.fr_exchange_norms):
        fn.argtypes = [vp]
        fn.restype = vp
    L.fr_exchange_pull.argtypes = [vp, vp]
    L.fr_exchange_release.argtypes = [vp, vp]
    L.fr_exchange_destroy.argtypes = [vp]
    L.fr_exchange_destroy.restype = None
    _lib = L
    return L


def check(ctx, rc: int, what: str) -> None:
    if rc != FR_OK:
        msg = lib().fr_last_error(ctx).decode() if ctx else ""
        raise FedrannError(f"{what}: {_ERR.get(rc, rc)} {msg}")
