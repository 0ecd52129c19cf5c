#!/usr/bin/env python
"""bench.py -- reads/s of the FEDRANN hot path (extract -> project -> kNN).

    python bench.py --gpus N --steps K --warmup W [--workload c3|c2|c4|s1] [--impl reference]

A step is one pass of the hot path over the whole synthetic read set of the
workload.  Default workload c3 = BASELINE.json configs[2] (the largest
configuration that fits one GPU): 248 Mb genome, 30x, ~15 kb HiFi-like reads,
k=17, sample 0.005, min-mult 2, D=256, kNN k=50.  With N ranks the same read
set is sharded by reads (strong scaling); every rank extracts and projects its
shard, the normalised embeddings are exchanged over NVLink and every rank
searches its own rows against the whole database.

value   device-resident inputs, CUDA-event time per step, max over ranks
e2e     through fedrann_b200.pipeline.run_hot_path_host: pinned host reads in,
        neighbour lists out to pinned host memory, copies inside the timed region
roofline        the kernel with the largest share of the step, timed per launch
                with CUDA event pairs inside libfedrann_b200.so (fr_profile_*)
stage_rooflines the extract (hashing; sampled and generic library) and project (SpMM)
                kernels against the HBM roofline (algorithmic bytes of SURVEY.md 8(d))
digest / parity every run proves its own answer: sha256 over all neighbour lists in global
                row order (identical for any N), 256 seeded query rows per rank re-answered by
                the exact fp32 scan (bit-equal), 64 seeded reads re-extracted and re-projected
                by the CPU oracle (bit-equal)
cpu_baseline    the reference's CPU pipeline on the box's host cores, bounded sample
--impl reference   that CPU pipeline alone, same metric/config, host-only (no GPU, no product .so)
"""
from __future__ import annotations

import os
import sys


def _host_cores() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


if "reference" in sys.argv:
    # The CPU arm uses every host core whatever launched it: torch.distributed.run exports
    # OMP_NUM_THREADS=1, which would make BLAS single-threaded (set before numpy loads).
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(_host_cores())

import argparse  # noqa: E402
import hashlib  # noqa: E402
import json  # noqa: E402
import tempfile  # noqa: E402
import threading  # noqa: E402
import time  # noqa: E402
from os.path import abspath, dirname, isfile, join  # noqa: E402

import numpy as np  # noqa: E402

ROOT = dirname(abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # SURVEY.md 8(d)
    "s1": dict(genome=2_000_000, coverage=30, mean_len=10000, len_model="fixed", sigma=0, min_len=1000,
               sub=0.05, ins=0.0, dele=0.0, k=15, fraction=0.05, min_mult=2, D=500, knn_k=50, seed=602,
               cpu_reads=(3000, 3000)),
    "c2": dict(genome=5_000_000, coverage=50, mean_len=10000, len_model="lognormal", sigma=0.5, min_len=1000,
               sub=0.02, ins=0.02, dele=0.02, k=15, fraction=0.05, min_mult=2, D=500, knn_k=50, seed=1,
               cpu_reads=(3000, 3000)),
    "c3": dict(genome=248_000_000, coverage=30, mean_len=15000, len_model="normal", sigma=2000, min_len=1000,
               sub=0.002, ins=0.0015, dele=0.0015, k=17, fraction=0.005, min_mult=2, D=256, knn_k=50, seed=2,
               cpu_reads=(3000, 600)),
    # BASELINE.json configs[3]: 3 M reads x 15 kb from a 3.1 Gb genome (~15x), 6 M embedding rows, kNN k=20
    "c4": dict(genome=3_100_000_000, coverage=14.516129, mean_len=15000, len_model="normal", sigma=2000, min_len=1000,
               sub=0.002, ins=0.0015, dele=0.0015, k=17, fraction=0.005, min_mult=2, D=256, knn_k=20, seed=3,
               cpu_reads=(3000, 100)),
}
METRIC = "reads_per_sec_extract_project_knn"
UNIT = "reads/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--knn-mode", type=int, default=0, help="0 = tensor-core path, 1 = exact fp32 scan")
    ap.add_argument("--cpu-sample-reads", type=int, default=None,
                    help="reads of the CPU extract+project sample (default per workload)")
    ap.add_argument("--cpu-knn-reads", type=int, default=None,
                    help="reads whose rows the CPU kNN leg searches (default per workload)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the digest and the parity block")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--ncu-traffic", type=float, default=None,
                    help="dram bytes per launch of the dominant kernel, measured by `ncu --set full` on this "
                         "very command (never a committed constant); roofline.traffic is null without it")
    return ap.parse_args()


def config_of(args) -> dict:
    """Static description of the run: a function of the command line only, so that both arms
    (and every N) print the same dictionary for the same flags."""
    w = WORKLOADS[args.workload]
    reads = max(1, int(round(w["genome"] * w["coverage"] / w["mean_len"])))
    return {"workload": "%s: synthetic %.1f Mb genome, %.3gx, ~%d kb reads, k=%d, sample %.3g, min-mult %d, D=%d, "
                        "kNN k=%d (exact cosine)" % (args.workload, w["genome"] / 1e6, w["coverage"],
                                                     w["mean_len"] // 1000, w["k"], w["fraction"], w["min_mult"],
                                                     w["D"], w["knn_k"]),
            "reads": reads, "embedding_rows": 2 * reads,
            "parallelism": "reads sharded over %d GPU(s), embedding rows exchanged over NVLink, query rows "
                           "sharded" % args.gpus,
            "l2": "inputs larger than L2, streamed once per step; 256 MiB L2 flush between steps",
            "knn_mode": "tcgen05+rerank" if args.knn_mode == 0 else "fp32 scan",
            "library": "canonical k-mers with count >= min-mult over all reads, seeded hash-threshold sample "
                       "(N1 rule), reverse complements appended"}


# ---------------------------------------------------------------------------
class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (NVML; same
    fields as the nvidia-smi clocks line of the profiling recipe)."""

    def __init__(self, index: int, period: float = 0.02):
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._h = None

    _NAMES = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
              0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
              0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                for bit, name in self._NAMES.items():
                    if r & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        if self._h is not None:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(1.0)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def physical_gpu_index(local: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local])
        except Exception:
            return local
    return local


def measured_peaks():
    p = join(ROOT, "MEASURED_PEAKS.json")
    if isfile(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tc_burst=d["bf16_tflops"], tc_sustained=d["bf16_tflops_sustained"],
                    src="measured")
    return dict(hbm=6650.0, tc_burst=1590.0, tc_sustained=1400.0, src="fallback")


# ---------------------------------------------------------------------------
# workload (native arm: built on the device)
# ---------------------------------------------------------------------------
class Workload:
    """The whole read set's plan + this rank's shard on the device."""

    def __init__(self, name: str, rank: int, world: int, device):
        import torch
        from fedrann_b200 import synthetic_device as SD
        from fedrann_b200.parallel import shard_reads
        from fedrann_b200.precompute import precompute_matrix
        w = WORKLOADS[name]
        self.name, self.cfg = name, w
        self.plan = SD.plan_reads(w["genome"], w["coverage"], w["seed"], mean_len=w["mean_len"],
                                  len_model=w["len_model"], sigma=w["sigma"], min_len=w["min_len"],
                                  sub=w["sub"], ins=w["ins"], dele=w["dele"])
        R = self.plan.n_reads
        bounds = shard_reads(self.plan.pre_offsets, world)
        self.r0, self.r1 = int(bounds[rank]), int(bounds[rank + 1])
        self.rows_per_rank = [2 * int(bounds[i + 1] - bounds[i]) for i in range(world)]   # known on every rank
        genome = SD.make_genome(w["genome"], w["seed"], device)
        self.bases, self.offsets = SD.generate_reads(self.plan, genome, self.r0, self.r1)
        del genome
        torch.cuda.empty_cache()
        self.n_reads_total, self.n_reads_local = R, self.r1 - self.r0
        codes, counts = SD.sampled_kmer_counts(self.bases, self.offsets, w["k"], w["fraction"], w["seed"])
        if world > 1:       # N1 across ranks: hash-partitioned owners, one all-to-all (parallel.partitioned_kmer_counts)
            import torch.distributed as dist
            from fedrann_b200.parallel import partitioned_kmer_counts
            codes, counts = partitioned_kmer_counts(codes, counts, w["min_mult"], dist.group.WORLD)
        self.lib = SD.library_from_counts(codes, counts, w["k"], w["min_mult"], w["seed"])
        del codes, counts
        torch.cuda.empty_cache()
        self.P = precompute_matrix(w["D"], self.lib.counts, self.lib.n_features)


def library_text_file(lib, path: str) -> None:
    with open(path, "wb") as f:
        f.write(lib.searcher_text())


# ---------------------------------------------------------------------------
# the reference's CPU pipeline on a bounded sample (cpu_baseline / --impl reference)
# ---------------------------------------------------------------------------
def knn_blas_threads(db: np.ndarray, dbn: np.ndarray, q1: int, k: int, cores: int):
    """Exact cosine kNN of unit rows [0, q1) against all rows: BLAS sgemm (all host threads) per block of
    query rows, then the per-row partial sort spread over the same threads (numpy releases the GIL)."""
    from concurrent.futures import ThreadPoolExecutor
    N = db.shape[0]
    kk = min(k, N)
    idx = np.zeros((q1, kk), dtype=np.int32)
    dist = np.zeros((q1, kk), dtype=np.float32)
    zero_db = np.flatnonzero(dbn == 0)
    block = int(max(cores, min(2048, (1 << 28) // max(N, 1))))       # <= 1 GiB of distances at a time

    def select(args):
        d, a0, lo, hi = args
        sub = d[lo:hi]
        part = np.argpartition(sub, kk - 1, axis=1)[:, :kk]
        pd = np.take_along_axis(sub, part, axis=1)
        order = np.lexsort((part, pd), axis=1)
        idx[a0 + lo:a0 + hi] = np.take_along_axis(part, order, axis=1)
        dist[a0 + lo:a0 + hi] = np.take_along_axis(pd, order, axis=1)

    with ThreadPoolExecutor(cores) as ex:
        for a in range(0, q1, block):
            b = min(q1, a + block)
            d = db[a:b] @ db.T
            np.subtract(1.0, d, out=d)
            if len(zero_db):
                d[:, zero_db] = 1.0
            d[dbn[a:b] == 0, :] = 1.0
            cuts = np.linspace(0, b - a, min(cores, b - a) + 1).astype(int)
            list(ex.map(select, [(d, a, int(cuts[i]), int(cuts[i + 1])) for i in range(len(cuts) - 1)]))
    return idx, dist


class CpuPipeline:
    """kmer_searcher (unmodified reference binary, oracle/_ref, all host threads)
    -> the oracle's restatement of get_feature_matrix (all host threads) ->
    exact cosine kNN of the first `knn_reads` reads' rows against a database of the
    workload's full size (BLAS sgemm + threaded partial sort; the reference's
    pynndescent backend is not installed here).  Host-only: nothing in here touches
    the GPU or libfedrann_b200.so."""

    def __init__(self, lib, P, bases_host: np.ndarray, offsets: np.ndarray, n_db_rows: int, knn_reads: int,
                 k_nn: int, cores: int, seed: int):
        from oracle import oracle as O
        self.O, self.lib, self.P, self.k_nn, self.cores = O, lib, P, k_nn, cores
        self.bases, self.offsets = bases_host, offsets
        self.n_reads = len(offsets) - 1
        self.knn_reads = min(knn_reads, self.n_reads)
        D = P.shape[1]
        # the database: the sample's own rows (written every step) + seeded Gaussian unit rows for the
        # reads outside the sample -- a brute-force search costs the same whatever the rows hold
        rng = np.random.default_rng(seed ^ 0xDB)
        n_db_rows = max(n_db_rows, 2 * self.knn_reads)
        self.db = np.empty((n_db_rows, D), dtype=np.float32)
        step = max(1, (1 << 26) // D)
        for a in range(0, n_db_rows, step):
            blk = rng.standard_normal((min(step, n_db_rows - a), D), dtype=np.float32)
            blk /= np.linalg.norm(blk, axis=1, keepdims=True)
            self.db[a:a + len(blk)] = blk
        self.dbn = np.ones(n_db_rows, dtype=np.float32)
        self.tmp = tempfile.mkdtemp(prefix="fedrann_cpu_")
        self.lib_path = join(self.tmp, "lib.txt")
        library_text_file(lib, self.lib_path)
        self.fa_path = join(self.tmp, "reads.fa")
        self.names = ["read%d" % i for i in range(self.n_reads)]
        with open(self.fa_path, "wb") as f:
            for i, n in enumerate(self.names):
                f.write(b">" + n.encode() + b"\n")
                f.write(bases_host[offsets[i]:offsets[i + 1]].tobytes())
                f.write(b"\n")
        self.kind = "reference" if isfile(O.REF_BINARY) else "port"
        self.lib_or = None if self.kind == "reference" else O.Library(lib.searcher_text(), lib.k)
        self.stage_s = {}
        self.last = None

    def step(self) -> float:
        O = self.O
        t0 = time.perf_counter()
        if self.kind == "reference":
            recs, _ = O.run_reference_binary(self.lib_path, self.fa_path, join(self.tmp, "out"), self.lib.k,
                                             self.cores)
            by = dict(recs)
            rows = [by[n] for n in self.names]
            indptr = np.zeros(len(rows) + 1, dtype=np.int64)
            np.cumsum([len(r) for r in rows], out=indptr[1:])
            indices = np.fromiter((x for r in rows for x in r), dtype=np.int32, count=int(indptr[-1]))
        else:
            indptr, indices, _ = O.extract(self.lib_or, self.bases, self.offsets)
        t1 = time.perf_counter()
        E = O.project_threads(indptr, indices, self.P, self.lib.nk, self.cores)
        t2 = time.perf_counter()
        nq = 2 * self.knn_reads
        Xn, norms = O.normalize(E[:nq])
        self.db[:nq], self.dbn[:nq] = Xn, norms
        try:
            from threadpoolctl import threadpool_limits
            with threadpool_limits(limits=self.cores):
                knn_blas_threads(self.db, self.dbn, nq, min(self.k_nn, self.db.shape[0]), self.cores)
        except ImportError:
            knn_blas_threads(self.db, self.dbn, nq, min(self.k_nn, self.db.shape[0]), self.cores)
        t3 = time.perf_counter()
        self.stage_s = {"extract_s": t1 - t0, "project_s": t2 - t1, "knn_s": t3 - t2}
        self.last = (indptr, indices, E)
        return t3 - t0

    def reads_per_sec(self) -> float:
        """seconds per read = (extract + project) / sample reads + kNN / kNN reads"""
        s = self.stage_s
        per_read = (s["extract_s"] + s["project_s"]) / self.n_reads + s["knn_s"] / max(1, self.knn_reads)
        return 1.0 / per_read

    def close(self):
        import shutil
        shutil.rmtree(self.tmp, ignore_errors=True)


def run_cpu(args, w, lib, P, bases_h, offsets, n_reads_total, steps, warmup):
    """Times the CPU pipeline on a bounded sample: extract + project on the first Rs reads, the kNN
    leg on the first Rk of them against a database of the workload's size."""
    cores = _host_cores()
    Rs = len(offsets) - 1
    Rk = min(args.cpu_knn_reads or w["cpu_reads"][1], Rs)
    cpu = CpuPipeline(lib, P, bases_h, offsets, 2 * n_reads_total, Rk, w["knn_k"], cores, w["seed"])
    try:
        for _ in range(warmup):
            cpu.step()
        ts, vals, stages = [], [], []
        for _ in range(max(1, steps)):
            ts.append(cpu.step())
            vals.append(cpu.reads_per_sec())
            stages.append(dict(cpu.stage_s))
    finally:
        cpu.close()
    sec = float(np.mean(ts))
    value = float(np.mean(vals))
    st = {k: float(np.mean([s[k] for s in stages])) for k in stages[0]}
    sample = ("extract+project on the first %d of %d reads (%.1f Mb): %s kmer_searcher %d threads %.2fs (includes "
              "loading the %d-k-mer library), projection (C restatement, %d threads) %.2fs; exact cosine kNN of the "
              "first %d reads' %d rows vs a %d-row database (sample rows + seeded Gaussian unit rows; BLAS sgemm + "
              "threaded partial sort, %d threads; pynndescent not installed) %.2fs; reads/s = 1 / (seconds per read "
              "summed over the stages)"
              % (Rs, n_reads_total, offsets[-1] / 1e6, "reference" if cpu.kind == "reference" else "ported", cores,
                 st["extract_s"], lib.n_features, cores, st["project_s"], Rk, 2 * Rk, cpu.db.shape[0], cores,
                 st["knn_s"]))
    return {"value": value, "ms_per_step": sec * 1e3,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": cpu.kind, "sample": sample,
                             "stage_s": st}}


def reference_arm(args):
    """Host-only: the workload (genome, sample reads, the library counted over ALL reads, P) is rebuilt
    with host threads (oracle/workload_gen.c: byte-identical to the device generator) and the CPU
    pipeline is timed.  No torch.cuda, no libfedrann_b200.so."""
    from oracle import workload as W
    from fedrann_b200.precompute import precompute_matrix
    w = WORKLOADS[args.workload]
    t0 = time.perf_counter()
    plan = W.plan_reads(w["genome"], w["coverage"], w["seed"], mean_len=w["mean_len"], len_model=w["len_model"],
                        sigma=w["sigma"], min_len=w["min_len"], sub=w["sub"], ins=w["ins"], dele=w["dele"])
    hw = W.HostWorkload(plan)
    codes, counts = hw.sampled_counts(w["k"], w["fraction"], w["seed"])
    lib = W.library_from_counts(codes, counts, w["k"], w["min_mult"], w["seed"])
    P = precompute_matrix(w["D"], lib.counts, lib.n_features)
    Rs = min(args.cpu_sample_reads or w["cpu_reads"][0], plan.n_reads)
    bases_h, offs = hw.reads(0, Rs)
    setup_s = time.perf_counter() - t0
    out = run_cpu(args, w, lib, P, bases_h, offs, plan.n_reads, steps=args.steps, warmup=args.warmup)
    out["cpu_baseline"]["setup_s"] = setup_s
    out["cpu_baseline"]["n_features"] = lib.n_features
    return out


# ---------------------------------------------------------------------------
_REAL_STDOUT = None


_T0 = time.perf_counter()


def progress(msg: str) -> None:
    """One line per phase on stderr (with the seconds since start): a run that stalls says where."""
    if int(os.environ.get("RANK", "0")) == 0:
        sys.stderr.write("[bench %7.1fs] %s\n" % (time.perf_counter() - _T0, msg))
        sys.stderr.flush()


def emit(line: dict) -> None:
    """The ONE JSON line goes to the real stdout; everything else a native library prints there
    (NCCL's version banner) was redirected to stderr at start-up."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    config = config_of(args)

    if args.impl == "reference":
        # host-only arm: rank 0 alone runs it, no process group, no GPU
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        out = reference_arm(args)
        line = {"metric": METRIC, "value": out["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": out["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "impl": "reference", "config": config, "cpu_baseline": out["cpu_baseline"],
                "e2e": {"value": out["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
        return 0

    import torch
    import torch.distributed as dist
    from fedrann_b200.parallel import init_from_env
    rank, world, local = init_from_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the native arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from fedrann_b200.engine import Engine
    from fedrann_b200.pipeline import run_hot_path, run_hot_path_host, HostBuffers, StageTimes

    w = WORKLOADS[args.workload]
    group = dist.group.WORLD if world > 1 else None
    progress("building workload %s on %d rank(s)" % (args.workload, world))
    wl = Workload(args.workload, rank, world, dev)
    assert wl.n_reads_total == config["reads"]
    progress("workload ready: %d reads on this rank, %d features" % (wl.n_reads_local, wl.lib.n_features))
    eng = Engine(local)
    # the library is a hash-threshold sample (what N1 produces): declared to the engine, which verifies it
    eng.build_table(wl.lib.all_codes(), w["k"], sampler=(w["seed"], w["fraction"]))
    P_dev = eng.upload_projection(wl.P)
    offsets_d = torch.from_numpy(wl.offsets).to(dev)
    nk, k_nn = wl.lib.nk, w["knn_k"]

    def device_step(timers=None):
        return run_hot_path(eng, wl.bases, offsets_d, P_dev, nk, k_nn, mode=args.knn_mode, group=group,
                            timers=timers, rows_per_rank=wl.rows_per_rank)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed_steps(fn, n):
        """n steps, one CUDA event pair per step on the launching stream, L2 flushed in between."""
        evs = []
        for _ in range(n):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in evs]

    n_warm = max(args.warmup, 3)
    progress("table and projection uploaded; warm-up")
    sync_all()      # ranks enter the first exchange together (its flag barrier has a 60 s watchdog)
    for i in range(n_warm):
        res = device_step()
        if i == 0:
            torch.cuda.synchronize()
            progress("first step done")
    sync_all()
    progress("warm-up done; timing %d steps" % args.steps)
    clocks = ClockSampler(physical_gpu_index(local))
    l0 = eng.launches
    eng.profile(True)
    stage = StageTimes(True)
    clocks.start()
    sync_all()
    per_step = timed_steps(lambda: device_step(stage), args.steps)
    sync_all()
    clk = clocks.stop()
    launches = eng.launches - l0
    ktimes = eng.kernel_times()
    eng.profile(False)
    fb_device = eng.knn_fallback_rows          # of the last device-resident step
    stage_ms = {k: v / args.steps for k, v in stage.totals().items()}
    total_ms = float(sum(per_step))
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = wl.n_reads_total / (ms_per_step / 1e3)
    res = device_step()
    torch.cuda.synchronize()
    progress("device-resident steps done: %.2f ms per step" % ms_per_step)

    # ---- the run proves its own answer (untimed)
    parity = None
    if not args.no_parity:
        parity = parity_block(args, w, wl, eng, res, rank, world, group, dev)

    progress("parity block done" if parity is not None else "parity block skipped")
    # ---- e2e: pinned host buffers -> public API -> pinned host results
    e2e = None
    fb_e2e = None
    if not args.no_e2e:
        hb = HostBuffers.from_device(wl.bases, wl.offsets, 2 * wl.n_reads_local, min(k_nn, 2 * wl.n_reads_total))

        def host_step():
            return run_hot_path_host(eng, hb, P_dev, nk, k_nn, mode=args.knn_mode, group=group,
                                     rows_per_rank=wl.rows_per_rank)
        for _ in range(2):
            host_step()
        sync_all()
        e2e_steps = timed_steps(host_step, args.steps)
        sync_all()
        fb_e2e = eng.knn_fallback_rows
        t = torch.tensor([float(sum(e2e_steps))], dtype=torch.float64, device=dev)
        hd = torch.tensor([float(hb.h2d_bytes), float(hb.d2h_bytes)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(hd, op=dist.ReduceOp.SUM)
        e2e = {"value": wl.n_reads_total / (float(t.item()) / args.steps / 1e3), "unit": UNIT,
               "h2d_bytes_per_step": int(hd[0].item()), "d2h_bytes_per_step": int(hd[1].item())}
        if parity is not None:      # the host path must give the very same lists
            same = bool(torch.equal(hb.nbr_idx.to(dev), res.nbr_idx) and
                        torch.equal(hb.nbr_dist.to(dev).view(torch.int32), res.nbr_dist.view(torch.int32)))
            flag = torch.tensor([1 if same else 0], dtype=torch.int64, device=dev)
            if world > 1:
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            parity["e2e_lists_equal_device_lists"] = bool(flag.item())
            parity["parity_ok"] = bool(parity["parity_ok"] and flag.item())
        del hb

    progress("host-path steps done")
    # ---- roofline of the dominant kernel (largest share of the device-timed step on this rank)
    peaks = measured_peaks()
    roof = roofline(ktimes, args.steps, wl, res, w, world, peaks, args.ncu_traffic, ms_per_step)
    # the hashing and SpMM stages against the HBM roofline as well (north star: "achieved HBM GB/s for the
    # hashing and SpMM stages"), same definition of algorithmic bytes; never allowed to break the line
    stage_roofs = {}
    for kn in ("fr_extract_kernel", "fr_project_kernel"):
        try:
            stage_roofs[kn] = roofline(ktimes, args.steps, wl, res, w, world, peaks, None, ms_per_step, name=kn)
        except Exception as e:      # noqa: BLE001
            stage_roofs[kn] = {"error": repr(e)}
    # the same extraction for a GENERIC library (no sampler declared: what bin/kmer_searcher and any
    # library produced by jellyfish | awk (fedrann/count_kmers.py:104-123) takes), timed on its own
    try:
        eng.build_table(wl.lib.all_codes(), w["k"])
        for _ in range(2):
            eng.extract(wl.bases, offsets_d)
        torch.cuda.synchronize()
        eng.profile(True)
        n_gen = 3
        for _ in range(n_gen):
            flush.zero_()
            ip_g, ix_g = eng.extract(wl.bases, offsets_d)
        torch.cuda.synchronize()
        kt_g = eng.kernel_times()
        eng.profile(False)
        stage_roofs["fr_extract_kernel_generic"] = roofline(kt_g, n_gen, wl, res, w, world, peaks, None, ms_per_step,
                                                            name="fr_extract_kernel")
        stage_roofs["fr_extract_kernel_generic"]["same_rows_as_sampled_path"] = bool(
            torch.equal(ip_g, res.indptr) and torch.equal(ix_g, res.indices))
        del ip_g, ix_g
    except Exception as e:      # noqa: BLE001
        stage_roofs["fr_extract_kernel_generic"] = {"error": repr(e)}

    progress("rooflines done")
    # ---- CPU baseline (rank 0, N=1 only): same host code as the reference arm, on this run's library
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        Rs = min(args.cpu_sample_reads or w["cpu_reads"][0], wl.n_reads_local)
        off = wl.offsets[: Rs + 1].copy()
        bases_h = wl.bases[: int(off[-1])].cpu().numpy()
        del flush
        torch.cuda.empty_cache()
        cpu = run_cpu(args, w, wl.lib, wl.P, bases_h, off, wl.n_reads_total, steps=1, warmup=0)["cpu_baseline"]

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": n_warm, "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "u64 hashing / f32 projection / f16 tensor-core "
                "distance tile with f32 accumulate + f32 re-rank" if args.knn_mode == 0 else
                "u64 hashing / f32 projection / f32 kNN",
                "data": "synthetic", "config": config, "clocks": clk,
                "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "stage_rooflines": stage_roofs,
                "cpu_baseline": cpu, "stage_ms": stage_ms,
                "kernels_ms_per_step": {k: v[1] / args.steps for k, v in ktimes.items()},
                "knn_fallback_rows": {"device_path_last_step": fb_device, "e2e_path_last_step": fb_e2e},
                "exchange": exchange_info(eng, world),
                "workload_stats": {"n_features": wl.lib.n_features, "reads_this_rank": wl.n_reads_local,
                                   "bases_this_rank": int(wl.bases.numel()), "nnz_this_rank": int(res.indices.numel())}}
        if parity is not None:
            line.update(parity)
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


# ---------------------------------------------------------------------------
def exchange_info(eng, world):
    """How the ranks' embedding rows reached each other in this run (None on one GPU)."""
    if world == 1:
        return None
    from fedrann_b200.pipeline import EXCHANGE
    x = getattr(eng, "_row_exchange", None)
    if EXCHANGE == "peer" and x:
        return {"transport": "peer memory over NVLink (CUDA IPC slabs, copy-engine pulls in ring order, shard flags "
                             "consumed by the distance kernel; csrc/exchange.cu)",
                "bytes_pulled_per_rank_per_step": int(x[1].bytes_pulled_per_step)}
    return {"transport": "nccl all_gather_into_tensor x3"}


def gather_rows_rank0(x, counts, rank, world, group):
    """Row shards of every rank -> all rows in global order (valid on every rank)."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return x
    nmax = max(counts)
    pad = torch.zeros((nmax,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    pad[: x.shape[0]] = x
    out = torch.empty((world * nmax,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * nmax: r * nmax + counts[r]] for r in range(world)])


def parity_block(args, w, wl, eng, res, rank, world, group, dev):
    """digest: sha256 over (neighbour idx, dist) of ALL rows in global row order -- the same string for
    any number of ranks.  knn rows: 4 seeded blocks of 64 of this rank's query rows re-answered by the
    exact fp32 scan (fr_knn mode 1), compared bit for bit.  reads: 64 seeded reads of rank 0 re-extracted
    and re-projected by the CPU oracle (oracle/fedrann_oracle.c), compared bit for bit."""
    import torch
    import torch.distributed as dist
    out = {}
    counts = wl.rows_per_rank
    idx_all = gather_rows_rank0(res.nbr_idx, counts, rank, world, group)
    dist_all = gather_rows_rank0(res.nbr_dist, counts, rank, world, group)
    if rank == 0:
        h = hashlib.sha256()
        h.update(idx_all.cpu().numpy().tobytes())
        h.update(dist_all.cpu().numpy().tobytes())
        out["digest"] = h.hexdigest()
        out["digest_of"] = "int32 idx [%d, %d] then fp32 dist, global row order" % tuple(idx_all.shape)
    del idx_all, dist_all
    ok = True
    # ---- kNN rows vs the exact scan
    nq = res.nbr_idx.shape[0]
    k_eff = res.nbr_idx.shape[1]
    rng = np.random.default_rng(1234 + rank)
    n_checked = 0
    knn_bad = 0
    if args.knn_mode == 0 and nq > 0:
        En, Eh, norms = res.db
        blk = min(64, nq)
        for o in sorted(set(int(x) for x in rng.integers(0, max(1, nq - blk + 1), 4))):
            ei, ed = eng.knn(En, Eh, norms, k_eff, q0=res.row_offset + o, nq=blk, mode=1)
            same_i = torch.equal(ei, res.nbr_idx[o:o + blk])
            same_d = torch.equal(ed.view(torch.int32), res.nbr_dist[o:o + blk].view(torch.int32))
            n_checked += blk
            if not (same_i and same_d):
                knn_bad += int(((ei != res.nbr_idx[o:o + blk]).any(1) |
                                (ed.view(torch.int32) != res.nbr_dist[o:o + blk].view(torch.int32)).any(1)).sum())
        ok = ok and knn_bad == 0
    # ---- reads vs the CPU oracle (rank 0)
    reads_checked = 0
    if rank == 0 and wl.n_reads_local > 0:
        from oracle import oracle as O
        R = wl.n_reads_local
        sel = sorted(set(int(x) for x in np.random.default_rng(4321).integers(0, R, 64)))
        offs = wl.offsets
        pieces = [wl.bases[int(offs[r]):int(offs[r + 1])].cpu().numpy() for r in sel]
        so = np.zeros(len(sel) + 1, dtype=np.int64)
        np.cumsum([len(p) for p in pieces], out=so[1:])
        L = O.Library(wl.lib.searcher_text(), w["k"])
        ip, ix, _ = O.extract(L, np.concatenate(pieces), so)
        g_ip = res.indptr.cpu().numpy()
        g_ix = res.indices
        E_or = O.project(ip, ix, wl.P, wl.lib.nk)
        for i, r in enumerate(sel):
            row = g_ix[int(g_ip[r]):int(g_ip[r + 1])].cpu().numpy()
            same = np.array_equal(row, ix[ip[i]:ip[i + 1]])
            e_gpu = res.embedding[2 * r:2 * r + 2].cpu().numpy()
            same = same and np.array_equal(e_gpu.view(np.uint32), E_or[2 * i:2 * i + 2].view(np.uint32))
            ok = ok and same
            reads_checked += 1
        del L
    flag = torch.tensor([1 if ok else 0, n_checked, knn_bad], dtype=torch.int64, device=dev)
    if world > 1:
        red = flag.clone()
        dist.all_reduce(red[:1], op=dist.ReduceOp.MIN)
        dist.all_reduce(flag[1:], op=dist.ReduceOp.SUM)
        flag[0] = red[0]
    out["parity_rows_checked"] = {"knn_rows_vs_exact_scan": int(flag[1].item()), "knn_rows_differing": int(flag[2].item()),
                                  "reads_vs_cpu_oracle": reads_checked}
    out["parity_ok"] = bool(flag[0].item())
    return out


def roofline(ktimes, steps, wl, res, w, world, peaks, ncu_traffic, ms_per_step, name=None):
    """Algorithmic bytes/flops per launch (DESIGN.md 'Kernels') / measured launch duration, for the
    kernel with the largest share of the step or for the one named."""
    if not ktimes or (name is not None and name not in ktimes):
        return None
    if name is None:
        name = max(ktimes, key=lambda k: ktimes[k][1])
    n_launch, tot_ms = ktimes[name]
    ms = tot_ms / n_launch
    R = wl.n_reads_local
    n_bases = int(wl.bases.numel())
    nnz = int(res.indices.numel())
    k = w["k"]
    N_db, nq, D = 2 * wl.n_reads_total, 2 * R, w["D"]
    step_s = tot_ms / steps / 1e3       # this kernel's time per step (all its launches)
    if name == "fr_extract_kernel":
        # per read: L ASCII + 8 B per probe + 4 B per hit + 8 B indptr
        lookups = n_bases - (k - 1) * R
        units = n_bases + 8 * lookups + 4 * nnz + 8 * R
        return _roof("hbm", units / 1e9 / step_s, peaks["hbm"], "GB/s", name, ms, n_launch // steps, ncu_traffic,
                     peaks["src"], tot_ms / steps / ms_per_step)
    if name == "fr_project_kernel":
        nnzP = int(wl.P.nnz)
        touched = 2 * nnz * (nnzP / max(1, wl.lib.n_features))
        units = 4 * nnz + 16 * 2 * nnz + 8 * touched + 2 * D * 4 * R
        return _roof("hbm", units / 1e9 / step_s, peaks["hbm"], "GB/s", name, ms, n_launch // steps, ncu_traffic,
                     peaks["src"], tot_ms / steps / ms_per_step)
    if name.startswith("fr_knn"):
        flops = 2.0 * nq * N_db * D      # true (unpadded) D, per step (all launches of the kernel)
        if name == "fr_knn_tc_kernel":
            r = _roof("tensor", flops / 1e12 / step_s, peaks["tc_burst"], "TFLOP/s", name, ms,
                      n_launch // steps, ncu_traffic, peaks["src"] + " burst (cuBLAS bf16, best of 10)",
                      tot_ms / steps / ms_per_step)
            # the kernel runs for seconds back to back in this bench: the power-capped figure beside it
            r["peak_sustained"] = peaks["tc_sustained"]
            r["frac_of_sustained"] = r["achieved"] / peaks["tc_sustained"]
            return r
        # the fp32 scan runs on the CUDA cores; report it against the tensor peak it is meant to be replaced by
        return _roof("tensor", flops / 1e12 / step_s, peaks["tc_burst"], "TFLOP/s", name, ms,
                     n_launch // steps, ncu_traffic, peaks["src"] + " burst (kernel runs on CUDA cores)",
                     tot_ms / steps / ms_per_step)
    units = 0.0
    return _roof("hbm", units, peaks["hbm"], "GB/s", name, ms, n_launch // steps, ncu_traffic, peaks["src"],
                 tot_ms / steps / ms_per_step)


def _roof(bound, achieved, peak, unit, kernel, ms, per_step, traffic, src, share):
    # traffic: dram__bytes_read.sum + dram__bytes_write.sum per launch from an ncu capture of THIS command
    # (--ncu-traffic); null when no capture was made for the run
    return {"bound": bound, "achieved": achieved, "peak": peak, "unit": unit, "frac": achieved / peak,
            "traffic": traffic, "traffic_gbps": (traffic / 1e9 / (ms / 1e3)) if traffic and ms > 0 else None,
            "kernel": kernel, "kernel_ms": ms, "launches_per_step": per_step,
            "share_of_step": share, "peak_source": src}


if __name__ == "__main__":
    sys.exit(main())
