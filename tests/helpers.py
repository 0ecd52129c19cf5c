"""Shared test helpers: golden loading and canonical forms."""
from os.path import abspath, dirname, join

import numpy as np

GOLDEN = join(dirname(abspath(__file__)), "golden")
EXTRACT_CASES = ["test1", "test2", "quirk_n", "quirk_short", "quirk_lib", "quirk_fastq", "quirk_k31", "s0"]


def load_extract_golden(tag):
    z = np.load(join(GOLDEN, "extract_%s.npz" % tag))
    names = [str(x) for x in z["names"]]
    rl = z["row_len"]
    ptr = np.concatenate([[0], np.cumsum(rl)])
    rows = {n: list(map(int, z["indices"][ptr[i]:ptr[i + 1]])) for i, n in enumerate(names)}
    freq = dict(zip(map(int, z["freq_index"]), map(int, z["freq_count"])))
    return dict(lib_text=z["lib_text"].tobytes(), fastx_text=z["fastx_text"].tobytes(), k=int(z["k"]),
                names=names, rows=rows, freq=freq)


def reads_to_arrays(recs):
    """[(name, seq bytes)] -> (bases uint8, offsets int64, names)."""
    bases = np.frombuffer(b"".join(s for _, s in recs), dtype=np.uint8)
    offs = np.zeros(len(recs) + 1, dtype=np.int64)
    np.cumsum([len(s) for _, s in recs], out=offs[1:])
    return bases, offs, [n for n, _ in recs]


def csr_rows(indptr, indices, names):
    return {n: list(map(int, indices[indptr[i]:indptr[i + 1]])) for i, n in enumerate(names)}
