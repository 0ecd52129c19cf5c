"""Live differential check of the projection oracle against the REFERENCE's own Python
(fedrann.precompute / fedrann.feature_extraction imported unmodified from /root/reference, as in
make_golden.py), on a seeded synthetic set:  python tests/golden/live_project_check.py SEED
Run as a subprocess by tests/test_oracle_golden.py (the reference forks a multiprocessing pool); exits 0
and prints 'ok ...' when the oracle's E is bit-identical to the reference's and the structure of P matches."""
import struct
import sys
import tempfile
from os.path import abspath, dirname, join

import numpy as np

HERE = dirname(abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as G  # noqa: E402  (puts the repo root on sys.path)
from oracle import oracle as O  # noqa: E402
from fedrann_b200 import synthetic as S  # noqa: E402


def write_output_bin(path, names, indptr, indices):
    with open(path, "wb") as f:
        f.write(b"KMER\x01\x00\x00\x00" + struct.pack("<Q", len(names)))
        for r, n in enumerate(names):
            ix = indices[indptr[r]:indptr[r + 1]]
            f.write(struct.pack("<H", len(n)) + n.encode() + struct.pack("<I", len(ix)) + ix.astype("<u8").tobytes())


def main(seed):
    r = np.random.default_rng(seed)
    k = int(r.choice([11, 13, 15, 17]))
    D = int(r.choice([8, 24, 64]))
    g = S.random_genome(int(r.integers(6000, 20000)), seed)
    rs = S.simulate_reads(g, 8, seed + 1, mean_len=int(r.integers(300, 1500)), len_model="lognormal", min_len=5,
                          sub=0.02, ins=0.01, dele=0.01, n_frac=0.002)
    lib = S.library_from_reads(rs, k, float(r.choice([0.05, 0.2, 0.5])), 2, seed)
    L = O.Library(lib.searcher_text(), k)
    indptr, indices, _ = O.extract(L, rs.bases, rs.offsets)
    keep = np.flatnonzero(np.diff(indptr) > 0)     # the reference shifts rows after an empty record
    names = [rs.names[i] for i in keep]
    ip = np.zeros(len(keep) + 1, dtype=np.int64)
    np.cumsum(np.diff(indptr)[keep], out=ip[1:])
    ix = np.concatenate([indices[indptr[i]:indptr[i + 1]] for i in keep]) if len(keep) else np.zeros(0, np.int32)
    d = tempfile.mkdtemp(prefix="live_project_")
    ob = join(d, "output.bin")
    write_output_bin(ob, names, ip, ix)
    pre, fe, gv = G.import_reference()
    import multiprocessing
    multiprocessing.set_start_method("fork", force=True)
    gv.threads = 2
    counter = join(d, "fwd_kmer_library.fasta")
    open(counter, "wb").write(lib.counter_fasta())
    P_ref, _ = pre.get_precompute_matrix(D, counter, lib.n_features)
    P_ref = P_ref.tocsr(); P_ref.sort_indices()
    P_or = O.precompute_matrix(D, lib.counts, lib.n_features)
    assert np.array_equal(P_ref.indptr, P_or.indptr) and np.array_equal(P_ref.indices, P_or.indices)
    assert np.allclose(P_ref.data, P_or.data, rtol=3e-7, atol=0)          # numpy-2 vs pinned numpy-1.26 last bit
    E_ref = fe.get_feature_matrix(ob, P_or, lib.n_features, len(names), int(r.integers(3, 50)))
    rn, st = fe.get_metadata(ob, lib.n_features)
    E = O.project(ip, ix, P_or, lib.nk)
    assert E.shape == E_ref.shape and np.array_equal(E.view(np.uint32), np.asarray(E_ref, dtype=np.float32).view(np.uint32))
    assert rn == [n for n in names for _ in (0, 1)] and list(st) == [0, 1] * len(names)
    print("ok seed", seed, "k", k, "D", D, "records", len(names), "nnz", len(ix), "F", lib.n_features)


if __name__ == "__main__":
    main(int(sys.argv[1]))
