"""Host-side pieces of bench.py and of the N4 writers (no GPU):

* oracle/workload_gen.c builds byte-identical reads / the identical library as the device
  generator (synthetic_device.py run on the CPU device) -- what lets the reference arm run
  host-only on the same workload;
* the reference arm prints the same `config` as the native arm, uses all host threads under
  torchrun's OMP_NUM_THREADS=1, and maps neither libfedrann_b200.so nor CUDA;
* feature_matrix.npz writer (vectorised, ZIP64, 64-bit indices) and the reference keyword names.
"""
import inspect
import json
import os
import subprocess
import sys
from os.path import abspath, dirname, join

import numpy as np
import pytest

ROOT = dirname(dirname(abspath(__file__)))

CASES = [
    dict(genome=300000, coverage=20, seed=7, mean_len=5000, len_model="lognormal", sigma=0.5, min_len=200,
         sub=0.02, ins=0.02, dele=0.02),
    dict(genome=200000, coverage=10, seed=2, mean_len=3000, len_model="normal", sigma=400, min_len=100,
         sub=0.002, ins=0.0015, dele=0.0015),
    dict(genome=100000, coverage=10, seed=602, mean_len=2000, len_model="fixed", sigma=0, min_len=100,
         sub=0.05, ins=0.0, dele=0.0),
]


@pytest.mark.parametrize("cfg", CASES)
def test_host_workload_equals_device_generator(cfg):
    from fedrann_b200 import synthetic_device as SD
    from oracle import workload as W
    kw = {k: cfg[k] for k in ("mean_len", "len_model", "sigma", "min_len", "sub", "ins", "dele")}
    plan = SD.plan_reads(cfg["genome"], cfg["coverage"], cfg["seed"], **kw)
    hp = W.plan_reads(cfg["genome"], cfg["coverage"], cfg["seed"], **kw)
    assert np.array_equal(plan.lens, hp.lens) and np.array_equal(plan.starts, hp.starts)
    assert np.array_equal(plan.strands, hp.strands)
    g = SD.make_genome(cfg["genome"], cfg["seed"], "cpu")
    hw = W.HostWorkload(hp, threads=3)
    assert np.array_equal(g.numpy(), hw.genome)
    R = plan.n_reads
    b, o = SD.generate_reads(plan, g, 0, R, chunk_bases=1 << 16)
    hb, ho = hw.reads(0, R)
    assert np.array_equal(o, ho) and np.array_equal(b.numpy(), hb)
    b2, o2 = SD.generate_reads(plan, g, 5, R - 3)            # a rank's shard: same bytes as in the whole set
    hb2, ho2 = hw.reads(5, R - 3)
    assert np.array_equal(o2, ho2) and np.array_equal(b2.numpy(), hb2)
    for k, fr in [(15, 0.05), (17, 0.3), (31, 1.0)]:
        c, n = SD.sampled_kmer_counts(b, o, k, fr, cfg["seed"])
        hc, hn = hw.sampled_counts(k, fr, cfg["seed"])
        assert np.array_equal(c.numpy().view(np.uint64), hc) and np.array_equal(n.numpy(), hn)
        l1 = SD.library_from_counts(c, n, k, 2, cfg["seed"])
        l2 = W.library_from_counts(hc, hn, k, 2, cfg["seed"])
        assert np.array_equal(l1.fwd_codes, l2.fwd_codes) and np.array_equal(l1.counts, l2.counts)


def test_config_is_a_function_of_the_command_line_only():
    sys.path.insert(0, ROOT)
    import bench
    for wl in bench.WORKLOADS:
        for n in (1, 8):
            ns_native = type("A", (), dict(workload=wl, gpus=n, knn_mode=0, impl="native"))
            ns_ref = type("A", (), dict(workload=wl, gpus=n, knn_mode=0, impl="reference"))
            assert bench.config_of(ns_native) == bench.config_of(ns_ref)
    c3 = bench.config_of(type("A", (), dict(workload="c3", gpus=1, knn_mode=0)))
    assert c3["workload"].startswith("c3") and c3["reads"] == 496000


def test_reference_arm_is_host_only_and_uses_all_threads(tmp_path):
    """--impl reference under torchrun's environment (OMP_NUM_THREADS=1, RANK set): rank 0 prints the line with
    every host core in use, rank 1 exits silently; the process maps no product library and no CUDA runtime."""
    env = dict(os.environ, OMP_NUM_THREADS="1", RANK="0", WORLD_SIZE="2", LOCAL_RANK="0", CUDA_VISIBLE_DEVICES="")
    code = (
        "import sys, json; sys.argv = ['bench.py', '--impl', 'reference', '--workload', 's1', '--gpus', '2', "
        "'--steps', '1', '--warmup', '0', '--cpu-sample-reads', '300', '--cpu-knn-reads', '100'];\n"
        "import bench\n"
        "rc = bench.main()\n"
        "maps = open('/proc/self/maps').read()\n"
        "import os\n"
        "os.write(2, json.dumps({'fedrann_so': 'libfedrann_b200' in maps, 'cudart': 'libcudart' in maps, "
        "'torch': 'libtorch' in maps, 'omp': os.environ['OMP_NUM_THREADS']}).encode() + b'\\n')\n"
        "sys.exit(rc)\n")
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["gpu_launches"] == 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["value"] > 0
    cores = len(os.sched_getaffinity(0))
    assert line["cpu_baseline"]["cores"] == cores and line["cpu_baseline"]["kind"] in ("reference", "port")
    sys.path.insert(0, ROOT)
    import bench
    assert line["config"] == bench.config_of(type("A", (), dict(workload="s1", gpus=2, knn_mode=0)))
    info = json.loads(r.stderr.strip().splitlines()[-1])
    assert info == {"fedrann_so": False, "cudart": False, "torch": False, "omp": str(cores)}
    # the other ranks do no work
    env["RANK"] = "1"
    r2 = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--workload", "s1", "--gpus", "2"],
                        cwd=ROOT, env=env, capture_output=True, text=True, timeout=120)
    assert r2.returncode == 0 and r2.stdout.strip() == ""


def test_cpu_knn_leg_matches_oracle_definition():
    sys.path.insert(0, ROOT)
    import bench
    from oracle import oracle as O
    rng = np.random.default_rng(5)
    X = rng.standard_normal((700, 48)).astype(np.float32)
    X[13] = 0
    Xn, norms = O.normalize(X)
    bi, bd = bench.knn_blas_threads(Xn.copy(), norms.copy(), 200, 10, 3)
    oi, od = O.knn(X, 10, 0, 200)
    assert np.allclose(bd, od, atol=1e-5)
    assert (bi == oi).mean() > 0.98


def test_feature_matrix_npz_vectorised_writer(tmp_path):
    import scipy.sparse as sp
    from fedrann_b200.__main__ import write_feature_matrix_npz
    rng = np.random.default_rng(0)
    R, nk = 300, 500
    rows = [np.sort(rng.choice(2 * nk, rng.integers(0, 40), replace=False)) for _ in range(R)]
    rows[3] = np.zeros(0, dtype=np.int64)
    rows[7] = np.arange(nk, nk + 9)            # only reverse-half hits
    rows[8] = np.arange(0, 9)                  # only forward-half hits
    indptr = np.zeros(R + 1, dtype=np.int64)
    np.cumsum([len(r) for r in rows], out=indptr[1:])
    indices = np.concatenate(rows).astype(np.int64)
    for dt in (None, np.int64):                # int64 = what 2*nnz >= 2^31 (whole-human scale) selects
        p = str(tmp_path / "f.npz")
        write_feature_matrix_npz(p, indptr, indices, nk, 2 * nk, index_dtype=dt)
        raw = np.load(p)
        assert set(raw.files) == {"indices", "indptr", "format", "shape", "data"}
        assert raw["indices"].dtype == (np.int32 if dt is None else np.int64)
        assert raw["data"].dtype == np.int8 and raw["format"].item() == b"csr"
        m = sp.load_npz(p)
        assert m.shape == (2 * R, 2 * nk) and m.format == "csr"
        for r in range(R):
            assert list(m.indices[m.indptr[2 * r]:m.indptr[2 * r + 1]]) == list(rows[r])
            sw = sorted((i + nk if i < nk else i - nk) for i in rows[r])
            assert list(m.indices[m.indptr[2 * r + 1]:m.indptr[2 * r + 2]]) == sw
    # ZIP64 structures are written even for small members (the 36 GB member of the whole-human scale needs them)
    import zipfile
    with zipfile.ZipFile(p) as zf:
        assert all(zi.extract_version >= 45 for zi in zf.infolist())


def test_index_dtype_rule_switches_to_64_bits():
    from fedrann_b200.__main__ import swapped_rows_csr
    ip = np.array([0, 2], dtype=np.int64)
    ix = np.array([1, 2 ** 30 + 5], dtype=np.int64)
    _, a = swapped_rows_csr(ip, ix, nk=2 ** 30)              # F = 2^31 does not fit int32
    assert a.dtype == np.int64 and list(a) == [1, 2 ** 30 + 5, 5, 2 ** 30 + 1]
    _, b = swapped_rows_csr(ip, np.array([1, 7]), nk=4)
    assert b.dtype == np.int32 and list(b) == [1, 7, 3, 5]


def test_reference_keyword_names_and_knn_limits():
    from fedrann_b200.__main__ import run_fedrann_pipeline
    from fedrann_b200.nearest_neighbors import check_knn_limits
    # fedrann/__main__.py:302-315: the reference's keyword names
    want = {"input_path", "output_dir", "kmer_size", "kmer_sample_fraction", "kmer_min_multiplicity",
            "embedding_dimension", "nndescent_n_trees", "nndescent_n_neighbors", "save_feature_matrix",
            "keep_intermediates", "chunk_size"}
    assert want <= set(inspect.signature(run_fedrann_pipeline).parameters)
    check_knn_limits(50, 500)
    check_knn_limits(128, 256)
    with pytest.raises(ValueError, match="at most 128"):
        check_knn_limits(200, 500)
    with pytest.raises(ValueError):
        check_knn_limits(0, 500)
