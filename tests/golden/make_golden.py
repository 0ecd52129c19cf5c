"""Generates tests/golden/*.npz from the REFERENCE itself (run in the build
container only; /root/reference does not exist on the GPU box).

    python tests/golden/make_golden.py

* extract_*.npz   -- outputs of the unmodified reference binary
                     (oracle/_ref/kmer_searcher) on its own fixtures
                     (kmer_searcher/test/data), on the quirk probes of
                     SURVEY.md 3.2 and on a seeded synthetic read set.
* project_s1.npz  -- fedrann.precompute.get_precompute_matrix and
                     fedrann.feature_extraction.get_feature_matrix imported
                     unmodified from /root/reference (modules they import but
                     never use on this path are stubbed), run on the synthetic
                     set above.
"""
import os
import sys
import types
import tempfile
from os.path import abspath, dirname, join

import numpy as np

ROOT = dirname(dirname(dirname(abspath(__file__))))
sys.path.insert(0, ROOT)
REF = "/root/reference"
HERE = dirname(abspath(__file__))

from oracle import oracle as O  # noqa: E402
from fedrann_b200 import synthetic as S  # noqa: E402


def import_reference():
    for name in ["colorama", "Bio", "ahocorasick", "sharedmem", "isal", "pysam",
                 "memory_profiler", "pynndescent", "hnswlib", "xxhash"]:
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                m = types.ModuleType(name)
                sys.modules[name] = m
    cm = sys.modules["colorama"]
    if not hasattr(cm, "Fore"):
        class _Any:
            def __getattr__(self, k):
                return ""
        cm.Fore, cm.Style, cm.init = _Any(), _Any(), (lambda **kw: None)
    if not hasattr(sys.modules["Bio"], "SeqIO"):
        sys.modules["Bio"].SeqIO = types.ModuleType("Bio.SeqIO")
        sys.modules["Bio.SeqIO"] = sys.modules["Bio"].SeqIO
    if not hasattr(sys.modules["isal"], "igzip"):
        sys.modules["isal"].igzip = types.ModuleType("isal.igzip")
        sys.modules["isal.igzip"] = sys.modules["isal"].igzip
    sys.path.insert(0, REF)
    import fedrann.precompute as pre
    import fedrann.feature_extraction as fe
    import fedrann.global_variables as gv
    return pre, fe, gv


def run_binary_case(tag, lib_text, fastx_text, k, fastx_name="reads.fa"):
    d = tempfile.mkdtemp(prefix="golden_")
    lp, rp = join(d, "lib.txt"), join(d, fastx_name)
    open(lp, "wb").write(lib_text)
    open(rp, "wb").write(fastx_text)
    recs, freq = O.run_reference_binary(lp, rp, join(d, "out"), k, 2)
    names = [n for n, _ in recs]
    lens = [len(ix) for _, ix in recs]
    flat = [i for _, ix in recs for i in ix]
    fk = sorted(freq)
    np.savez_compressed(
        join(HERE, "extract_%s.npz" % tag),
        lib_text=np.frombuffer(lib_text, dtype=np.uint8),
        fastx_text=np.frombuffer(fastx_text, dtype=np.uint8),
        k=k, names=np.array(names, dtype=object).astype(str), row_len=np.array(lens, dtype=np.int64),
        indices=np.array(flat, dtype=np.int64),
        freq_index=np.array(fk, dtype=np.int64), freq_count=np.array([freq[i] for i in fk], dtype=np.int64))
    print(tag, "records", len(recs), "nnz", len(flat))
    return d


def main():
    td = join(REF, "kmer_searcher", "test", "data")
    rd = lambda n: open(join(td, n), "rb").read()
    run_binary_case("test1", rd("test1.kmers.txt"), rd("test1.fastq"), 15, "reads.fastq")
    run_binary_case("test2", rd("test2.kmers.txt"), rd("test2.fasta"), 13)
    # quirk probes
    run_binary_case("quirk_n", b"AAC\nTTA\nTAC\nACG\nCGT\n", b">q\nACGNAC\n", 3)
    run_binary_case("quirk_short", b"AAC\nAAA\nTTC\nTAC\n", b">s1\nAC\n>s2\nNC\n>s3\nCN\n>s4\nC\n", 3)
    run_binary_case("quirk_lib", b"ACGT 7\nAACC\nACGT\nGGTT\nACGN\nACG\nacgt\nggtt\n",
                    b">a\nACGTAACCGGTT\n>b\naacc\nggtt\n>c x y\nAC\nGT\n", 4)
    run_binary_case("quirk_fastq", b"ACGTA\nCGTAC\n", b"@r1 extra words\nACGTAC\n+\n!!!!!!\n@r2\nacgta\n+\n!!!!!\n", 5,
                    "reads.fastq")
    run_binary_case("quirk_k31", b"ACGTACGTACGTACGTACGTACGTACGTACG\nTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTT\n",
                    b">a\nACGTACGTACGTACGTACGTACGTACGTACGTNTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTT\n>b\nNT\n", 31)

    # seeded synthetic set S0 (small): reads with all error kinds + N
    g = S.random_genome(60000, 602)
    rs = S.simulate_reads(g, 12, 11, mean_len=3000, len_model="lognormal", min_len=5,
                          sub=0.02, ins=0.01, dele=0.01, n_frac=0.002)
    lib = S.library_from_reads(rs, 15, 0.05, 2, 602)
    d = run_binary_case("s0", lib.searcher_text(), rs.fasta(), 15)

    # projection through the imported reference Python
    pre, fe, gv = import_reference()
    import multiprocessing
    multiprocessing.set_start_method("fork", force=True)
    gv.threads = 2
    counter = join(d, "fwd_kmer_library.fasta")
    open(counter, "wb").write(lib.counter_fasta())
    D = 64
    P_ref, nf = pre.get_precompute_matrix(D, counter, lib.n_features)
    P_ref = P_ref.tocsr()
    P_ref.sort_indices()
    # the oracle's P (pinned numpy-1.26 arithmetic) fed to the reference's own SpGEMM driver
    P_or = O.precompute_matrix(D, lib.counts, lib.n_features)
    # the reference's row-shift defect (feature_extraction.py:187-195) is only
    # harmless when no record is empty: use the canonical ordered output we
    # write ourselves, dropping empty records first
    recs = O.read_output_bin(join(d, "out", "output.bin"))
    assert all(len(ix) > 0 for _, ix in recs), "synthetic set must have no empty rows"
    E_ref = fe.get_feature_matrix(join(d, "out", "output.bin"), P_or, lib.n_features, len(recs), 37)
    names, strands = fe.get_metadata(join(d, "out", "output.bin"), lib.n_features)
    np.savez_compressed(
        join(HERE, "project_s0.npz"),
        D=D, n_features=lib.n_features, counts=lib.counts,
        P_ref_indptr=P_ref.indptr.astype(np.int64), P_ref_indices=P_ref.indices.astype(np.int32),
        P_ref_data64=P_ref.data.astype(np.float64),
        E_ref=E_ref, names=np.array(names).astype(str), strands=np.array(strands, dtype=np.int8))
    print("project_s0", E_ref.shape, "P nnz", P_ref.nnz, "numpy", np.__version__)


if __name__ == "__main__":
    main()
