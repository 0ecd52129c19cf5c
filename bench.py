#!/usr/bin/env python
"""bench.py -- reads/s of the FEDRANN hot path (extract -> project -> kNN).

    python bench.py --gpus N --steps K --warmup W [--workload c2|c3|s1] [--impl reference]

A step is one pass of the hot path over the whole synthetic read set of the
workload (default c2 = BASELINE.json configs[1]: 5 Mb genome, 50x, ~10 kb ONT
reads, k=15, sample 0.05, min-mult 2, D=500, kNN k=50).  With N ranks the same
read set is sharded by reads (strong scaling); every rank extracts and
projects its shard, the normalised embeddings are all-gathered and every rank
searches its own rows against the whole database.

value   device-resident inputs, CUDA-event time per step, max over ranks
e2e     through fedrann_b200.pipeline.run_hot_path_host: pinned host reads in,
        neighbour lists out to pinned host memory, copies inside the timed region
roofline        the kernel with the largest share of the step, timed per launch
                with CUDA event pairs inside libfedrann_b200.so (fr_profile_*)
stage_rooflines the extract (hashing) and project (SpMM) kernels against the HBM roofline,
                same algorithmic-byte definitions
cpu_baseline    the reference's CPU pipeline on the box's host cores, bounded sample
--impl reference   that CPU pipeline alone, same metric/config
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import threading
import time
from os.path import abspath, dirname, isfile, join

import numpy as np

ROOT = dirname(abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # SURVEY.md 8(d)
    "s1": dict(genome=2_000_000, coverage=30, mean_len=10000, len_model="fixed", sigma=0, min_len=1000,
               sub=0.05, ins=0.0, dele=0.0, k=15, fraction=0.05, min_mult=2, D=500, knn_k=50, seed=602),
    "c2": dict(genome=5_000_000, coverage=50, mean_len=10000, len_model="lognormal", sigma=0.5, min_len=1000,
               sub=0.02, ins=0.02, dele=0.02, k=15, fraction=0.05, min_mult=2, D=500, knn_k=50, seed=1),
    "c3": dict(genome=248_000_000, coverage=30, mean_len=15000, len_model="normal", sigma=2000, min_len=1000,
               sub=0.002, ins=0.0015, dele=0.0015, k=17, fraction=0.005, min_mult=2, D=256, knn_k=50, seed=2),
}
METRIC = "reads_per_sec_extract_project_knn"
UNIT = "reads/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--knn-mode", type=int, default=0, help="0 = tensor-core path, 1 = exact fp32 scan")
    ap.add_argument("--cpu-sample-reads", type=int, default=3000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ncu-traffic", type=float, default=None,
                    help="dram bytes per launch of the dominant kernel from an ncu --set full capture")
    return ap.parse_args()


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
# workload
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
        self.n_reads_total, self.n_reads_local = R, self.r1 - self.r0
        codes, counts = SD.sampled_kmer_counts(self.bases, self.offsets, w["k"], w["fraction"], w["seed"])
        codes, counts = SD.merge_counts(codes, counts)
        self.lib = SD.library_from_counts(codes, counts, w["k"], w["min_mult"], w["seed"])
        del codes, counts
        torch.cuda.empty_cache()
        self.P = precompute_matrix(w["D"], self.lib.counts, self.lib.n_features)
        self.total_bases = None


def library_text_file(lib, path: str) -> None:
    with open(path, "wb") as f:
        f.write(lib.searcher_text())


# ---------------------------------------------------------------------------
# the reference's CPU pipeline on a bounded sample (cpu_baseline / --impl reference)
# ---------------------------------------------------------------------------
class CpuPipeline:
    """kmer_searcher (unmodified reference binary, oracle/_ref, all host threads)
    -> the oracle's restatement of get_feature_matrix (all host threads) ->
    exact cosine kNN of the sample's rows against the whole database (BLAS, all
    host threads; the reference's pynndescent backend is not installed here)."""

    def __init__(self, lib, P, bases_host: np.ndarray, offsets: np.ndarray, db_unit: np.ndarray,
                 db_norms: np.ndarray, k_nn: int, cores: int):
        from oracle import oracle as O
        self.O, self.lib, self.P, self.k_nn, self.cores = O, lib, P, k_nn, cores
        self.bases, self.offsets = bases_host, offsets
        self.db, self.db_norms = db_unit, db_norms
        self.tmp = tempfile.mkdtemp(prefix="fedrann_cpu_")
        self.lib_path = join(self.tmp, "lib.txt")
        library_text_file(lib, self.lib_path)
        self.fa_path = join(self.tmp, "reads.fa")
        self.names = ["read%d" % i for i in range(len(offsets) - 1)]
        with open(self.fa_path, "wb") as f:
            for i, n in enumerate(self.names):
                f.write(b">" + n.encode() + b"\n")
                f.write(bases_host[offsets[i]:offsets[i + 1]].tobytes())
                f.write(b"\n")
        self.kind = "reference" if isfile(O.REF_BINARY) else "port"
        self.lib_or = None if self.kind == "reference" else O.Library(lib.searcher_text(), lib.k)
        self.stage_s = {}

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
        Xn, norms = O.normalize(E)
        # database = the full (GPU-computed, parity-checked) embedding set with the sample's rows in place
        db, dbn = self.db, self.db_norms
        n = Xn.shape[0]
        db[:n], dbn[:n] = Xn, norms
        O.knn_blas(db, dbn, 0, n, min(self.k_nn, db.shape[0]))
        t3 = time.perf_counter()
        self.stage_s = {"extract_s": t1 - t0, "project_s": t2 - t1, "knn_s": t3 - t2}
        return t3 - t0

    def close(self):
        import shutil
        shutil.rmtree(self.tmp, ignore_errors=True)


# ---------------------------------------------------------------------------
_REAL_STDOUT = None


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
    import torch
    import torch.distributed as dist
    from fedrann_b200.parallel import init_from_env

    if args.impl == "reference":
        # host-only arm: rank 0 alone runs it, no process group
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        rank, world, local = 0, 1, int(os.environ.get("LOCAL_RANK", "0"))
    else:
        rank, world, local = init_from_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the native arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from fedrann_b200.engine import Engine
    from fedrann_b200.pipeline import run_hot_path, run_hot_path_host, HostBuffers, StageTimes

    w = WORKLOADS[args.workload]
    group = dist.group.WORLD if world > 1 else None
    world_eff, rank_eff = world, rank
    # (the reference arm builds the same workload, library and embedding database on the GPU as
    # untimed set-up, then times host code only)
    wl = Workload(args.workload, rank_eff, world_eff, dev)
    eng = Engine(local)
    # the library is a hash-threshold sample (what N1 produces): declared to the engine, which verifies it
    eng.build_table(wl.lib.all_codes(), w["k"], sampler=(w["seed"], w["fraction"]))
    P_dev = eng.upload_projection(wl.P)
    offsets_d = torch.from_numpy(wl.offsets).to(dev)
    nk, k_nn = wl.lib.nk, w["knn_k"]
    config = {"workload": "%s: synthetic %.1f Mb genome, %dx, ~%d kb reads, k=%d, sample %.3g, min-mult %d, D=%d, "
                          "kNN k=%d (exact cosine)" % (args.workload, w["genome"] / 1e6, w["coverage"],
                                                       w["mean_len"] // 1000, w["k"], w["fraction"], w["min_mult"],
                                                       w["D"], k_nn),
              "reads": wl.n_reads_total, "embedding_rows": 2 * wl.n_reads_total, "n_features": wl.lib.n_features,
              "parallelism": "reads sharded over %d GPU(s), one all-gather of embeddings" % world_eff,
              "l2": "inputs (%.0f MB of reads per rank) streamed once per step; 256 MiB L2 flush between steps"
                    % (wl.bases.numel() / 1e6),
              "knn_mode": "tcgen05+rerank" if args.knn_mode == 0 else "fp32 scan",
              "library": "hash-threshold sample, verified on device" if eng.sampled else "arbitrary"}

    def device_step(timers=None):
        return run_hot_path(eng, wl.bases, offsets_d, P_dev, nk, k_nn, mode=args.knn_mode, group=group,
                            timers=timers, rows_per_rank=wl.rows_per_rank)

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        res = device_step()
        torch.cuda.synchronize()
        out = run_cpu(args, wl, eng, res, w, k_nn, steps=args.steps, warmup=args.warmup)
        line = {"metric": METRIC, "value": out["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": out["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "impl": "reference", "config": config, "cpu_baseline": out["cpu_baseline"],
                "e2e": {"value": out["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
        return 0

    # ------------------------------------------------------------------ native arm
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

    for _ in range(max(args.warmup, 3)):
        res = device_step()
    sync_all()
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
    stage_ms = {k: v / args.steps for k, v in stage.totals().items()}
    total_ms = float(sum(per_step))
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = wl.n_reads_total / (ms_per_step / 1e3)

    # ---- e2e: pinned host buffers -> public API -> pinned host results
    hb = HostBuffers.from_device(wl.bases, wl.offsets, 2 * wl.n_reads_local, min(k_nn, 2 * wl.n_reads_total))
    for _ in range(2):
        run_hot_path_host(eng, hb, P_dev, nk, k_nn, mode=args.knn_mode, group=group, rows_per_rank=wl.rows_per_rank)
    sync_all()
    e2e_steps = timed_steps(lambda: run_hot_path_host(eng, hb, P_dev, nk, k_nn, mode=args.knn_mode, group=group, rows_per_rank=wl.rows_per_rank),
                            args.steps)
    sync_all()
    t = torch.tensor([float(sum(e2e_steps))], dtype=torch.float64, device=dev)
    hd = torch.tensor([float(hb.h2d_bytes), float(hb.d2h_bytes)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(hd, op=dist.ReduceOp.SUM)
    e2e_value = wl.n_reads_total / (float(t.item()) / args.steps / 1e3)

    # ---- roofline of the dominant kernel (largest share of the device-timed step on this rank)
    peaks = measured_peaks()
    traffic = args.ncu_traffic
    if traffic is None and args.workload == "c2" and world == 1:
        # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel from the committed
        # `ncu --set full` capture of this workload (profiles/ncu_traffic.json)
        tp = join(ROOT, "profiles", "ncu_traffic.json")
        if isfile(tp) and ktimes:
            top = max(ktimes, key=lambda k: ktimes[k][1])
            traffic = json.load(open(tp)).get(top, {}).get("dram_bytes_per_launch")
    roof = roofline(ktimes, args.steps, wl, res, w, world_eff, peaks, traffic, ms_per_step)
    # the hashing and SpMM stages against the HBM roofline as well (north star: "achieved HBM GB/s for the
    # hashing and SpMM stages"), same definition of algorithmic bytes; never allowed to break the line
    stage_roofs = {}
    for kn in ("fr_extract_kernel", "fr_project_kernel"):
        try:
            tr = None
            tp = join(ROOT, "profiles", "ncu_traffic.json")
            if args.workload == "c2" and world == 1 and isfile(tp):
                tr = json.load(open(tp)).get(kn, {}).get("dram_bytes_per_launch")
            stage_roofs[kn] = roofline(ktimes, args.steps, wl, res, w, world_eff, peaks, tr, ms_per_step, name=kn)
        except Exception as e:      # noqa: BLE001
            stage_roofs[kn] = {"error": repr(e)}

    # ---- CPU baseline (rank 0, N=1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = run_cpu(args, wl, eng, res, w, k_nn, steps=1, warmup=0)["cpu_baseline"]

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "u64 hashing / f32 projection / f16 tensor-core "
                "distance tile with f32 accumulate + f32 re-rank" if args.knn_mode == 0 else
                "u64 hashing / f32 projection / f32 kNN",
                "data": "synthetic", "config": config, "clocks": clk,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(hd[0].item()),
                        "d2h_bytes_per_step": int(hd[1].item())},
                "gpu_launches": int(launches), "roofline": roof, "stage_rooflines": stage_roofs,
                "cpu_baseline": cpu,
                "stage_ms": stage_ms,
                "kernels_ms_per_step": {k: v[1] / args.steps for k, v in ktimes.items()},
                "knn_fallback_rows": eng.knn_fallback_rows}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


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
    in_step = ms_per_step > 1000.0   # a kernel inside a seconds-long step: sustained peak applies
    if name == "fr_extract_kernel":
        # per read: L ASCII + 8 B per probe + 4 B per hit + 8 B indptr
        lookups = n_bases - (k - 1) * R
        units = n_bases + 8 * lookups + 4 * nnz + 8 * R
        return _roof("hbm", units / 1e9 / (ms / 1e3), peaks["hbm"], "GB/s", name, ms, n_launch // steps, ncu_traffic,
                     peaks["src"], tot_ms / steps / ms_per_step)
    if name == "fr_project_kernel":
        nnzP = int(wl.P.nnz)
        touched = 2 * nnz * (nnzP / max(1, wl.lib.n_features))
        units = 4 * nnz + 16 * 2 * nnz + 8 * touched + 2 * D * 4 * R
        return _roof("hbm", units / 1e9 / (ms / 1e3), peaks["hbm"], "GB/s", name, ms, n_launch // steps, ncu_traffic,
                     peaks["src"], tot_ms / steps / ms_per_step)
    if name.startswith("fr_knn"):
        flops = 2.0 * nq * N_db * D      # true (unpadded) D
        per_launch = flops * steps / n_launch
        if name == "fr_knn_tc_kernel":
            pk = peaks["tc_sustained"] if in_step else peaks["tc_burst"]
            return _roof("tensor", per_launch / 1e12 / (ms / 1e3), pk, "TFLOP/s", name, ms, n_launch // steps,
                         ncu_traffic, peaks["src"] + (" sustained" if in_step else " burst"),
                         tot_ms / steps / ms_per_step)
        # the fp32 scan runs on the CUDA cores; report it against the tensor peak it is meant to be replaced by
        return _roof("tensor", per_launch / 1e12 / (ms / 1e3), peaks["tc_burst"], "TFLOP/s", name, ms,
                     n_launch // steps, ncu_traffic, peaks["src"] + " burst (kernel runs on CUDA cores)",
                     tot_ms / steps / ms_per_step)
    units = 0.0
    return _roof("hbm", units, peaks["hbm"], "GB/s", name, ms, n_launch // steps, ncu_traffic, peaks["src"],
                 tot_ms / steps / ms_per_step)


def _roof(bound, achieved, peak, unit, kernel, ms, per_step, traffic, src, share):
    # traffic_gbps: the kernel's measured DRAM traffic (ncu capture) over its live launch duration
    return {"bound": bound, "achieved": achieved, "peak": peak, "unit": unit, "frac": achieved / peak,
            "traffic": traffic, "traffic_gbps": (traffic / 1e9 / (ms / 1e3)) if traffic and ms > 0 else None,
            "kernel": kernel, "kernel_ms": ms, "launches_per_step": per_step,
            "share_of_step": share, "peak_source": src}


def run_cpu(args, wl, eng, res, w, k_nn, steps, warmup):
    """Times the CPU pipeline on the first `cpu_sample_reads` reads of the workload."""
    import torch
    cores = os.cpu_count() or 1
    Rs = min(args.cpu_sample_reads, wl.n_reads_local)
    off = wl.offsets[: Rs + 1].copy()
    bases_h = wl.bases[: int(off[-1])].cpu().numpy()
    En, _, norms = eng.normalize(res.embedding)
    db = En.cpu().numpy().copy()
    dbn = norms.cpu().numpy().copy()
    cpu = CpuPipeline(wl.lib, wl.P, bases_h, off, db, dbn, k_nn, cores)
    try:
        for _ in range(warmup):
            cpu.step()
        ts = [cpu.step() for _ in range(max(1, steps))]
    finally:
        cpu.close()
    sec = float(np.mean(ts))
    value = Rs / sec
    sample = ("first %d of %d reads (%.1f Mb): %s kmer_searcher %d threads %.2fs, projection (C restatement, %d "
              "threads) %.2fs, exact cosine kNN of the sample's %d rows vs all %d rows (BLAS sgemm + partial sort, "
              "%d threads; pynndescent not installed) %.2fs"
              % (Rs, wl.n_reads_total, off[-1] / 1e6, "reference" if cpu.kind == "reference" else "ported", cores,
                 cpu.stage_s["extract_s"], cores, cpu.stage_s["project_s"], 2 * Rs, db.shape[0], cores,
                 cpu.stage_s["knn_s"]))
    return {"value": value, "ms_per_step": sec * 1e3,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": cpu.kind, "sample": sample,
                             "stage_s": cpu.stage_s}}


if __name__ == "__main__":
    sys.exit(main())
