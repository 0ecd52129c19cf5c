"""`fedrann` command line, drop-in for fedrann/__main__.py:69-171, 302-457:
same flags and defaults, same outputs (fedrann.log, overlaps.tsv,
feature_matrix.npz with --save-feature-matrix, temp/ kept with
--keep-intermediates).  Run under torchrun to shard the reads over the GPUs of
one box (one process per GPU); rank 0 owns the log and the library files, every
rank appends its rows of overlaps.tsv in rank order.
"""
from __future__ import annotations

import argparse
import functools
import os
import time
from os.path import abspath, join
from shutil import rmtree

import numpy as np

from . import global_variables
from .custom_logging import add_log_file, logger

__version__ = "1.0.0+b200"
__description__ = "FEDRANN (B200-native hot path): k-mer features -> sparse random projection -> exact kNN"


def parse_command_line_arguments(argv=None):
    parser = argparse.ArgumentParser(prog="fedrann", description=__description__,
                                     formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    parser.add_argument("-i", "--input", type=str, required=True, help="Path to the input FASTQ/FASTA file.")
    parser.add_argument("-o", "--output-dir", type=str, required=True, help="Directory to save output files.")
    parser.add_argument("-k", "--kmer-size", type=int, required=False, default=16,
                        help="K-mer size for feature extraction.")
    parser.add_argument("--kmer-sample-fraction", type=float, required=False, default=0.005,
                        help="Percentage of k-mer used to build feature matrix.")
    parser.add_argument("--kmer-min-multiplicity", type=int, required=False, default=2,
                        help="Minimum allowed frequency of a k-mer in all reads.")
    parser.add_argument("--threads", type=int, required=False, default=1)
    parser.add_argument("--chunk-size", type=int, required=False, default=1000)
    parser.add_argument("-n", "--embedding-dimension", type=int, required=False, default=500)
    parser.add_argument("--nndescent-n-trees", type=int, default=300,
                        help="Number of trees to use in NNDescent (accepted; the exact search has no trees).")
    parser.add_argument("--nndescent-n-neighbors", type=int, default=50,
                        help="Number of neighbors (including the read itself) reported per read orientation.")
    parser.add_argument("--seed", type=int, required=False, default=356115, help="Random seed for reproducibility.")
    parser.add_argument("--save-feature-matrix", action="store_true", default=False,
                        help="Save the feature matrix to a file.")
    parser.add_argument("--keep-intermediates", action="store_true", default=False,
                        help="Do not remove intermediate files")
    parser.add_argument("--mprof", action="store_true", default=False, help="Record memory usage.")
    return parser.parse_args(argv)


def swapped_rows_csr(indptr: np.ndarray, indices: np.ndarray, nk: int, index_dtype=None):
    """CSR of the (2R x F) binary feature matrix from the R extracted rows: row 2r = the read's
    ascending hits, row 2r+1 = the index-swapped (reverse-orientation) set in ascending order, which
    is `[hits >= nk] - nk` then `[hits < nk] + nk` (feature_extraction.py:137-139) -- no sort.
    Vectorised over all non-zeros (3 M reads at the whole-human scale).  Indices are int32 unless
    2 * nnz (or F) needs 64 bits, like scipy's own index-dtype rule."""
    indptr = np.asarray(indptr, dtype=np.int64)
    R = len(indptr) - 1
    cnt = np.diff(indptr)
    nnz = int(indptr[-1])
    ip2 = np.zeros(2 * R + 1, dtype=np.int64)
    np.cumsum(np.repeat(cnt, 2), out=ip2[1:])
    if index_dtype is None:
        index_dtype = np.int32 if max(2 * nnz, 2 * int(nk)) < 2 ** 31 else np.int64
    ix2 = np.empty(2 * nnz, dtype=index_dtype)
    if nnz:
        indices = np.asarray(indices[:nnz])
        rows = np.repeat(np.arange(R, dtype=np.int64), cnt)
        j = np.arange(nnz, dtype=np.int64) - indptr[rows]            # position inside the row
        lo = indices < nk
        n_lo = np.bincount(rows[lo], minlength=R).astype(np.int64)
        n_hi = cnt - n_lo
        ix2[ip2[2 * rows] + j] = indices                              # forward row
        # the row is ascending: its first n_lo entries are < nk, the rest >= nk
        dst = ip2[2 * rows + 1] + np.where(lo, n_hi[rows] + j, j - n_lo[rows])
        ix2[dst] = np.where(lo, indices.astype(np.int64) + nk, indices.astype(np.int64) - nk)
    return ip2, ix2


def write_feature_matrix_npz(path: str, indptr: np.ndarray, indices: np.ndarray, nk: int, n_features: int,
                             index_dtype=None) -> None:
    """feature_matrix.npz in scipy.sparse.save_npz's layout (members format, shape, data, indices,
    indptr; loadable with scipy.sparse.load_npz) of the binary (2R x n_features) k-mer feature
    matrix, row 2r forward / 2r+1 reverse orientation (README.md:66, 201-206).
    Written member by member straight into a ZIP64 archive (more than 2^31 non-zeros at the
    whole-human scale: 64-bit indices and members beyond 4 GiB), compressed like save_npz."""
    import zipfile
    from numpy.lib import format as npy_format
    ip2, ix2 = swapped_rows_csr(indptr, indices, nk, index_dtype)
    R2 = len(ip2) - 1
    if ix2.dtype == np.int32:
        ip2 = ip2.astype(np.int32)
    members = [("indices", ix2), ("indptr", ip2), ("format", np.array("csr".encode("ascii"))),
               ("shape", np.array((R2, int(n_features)), dtype=np.int64)),
               ("data", np.ones(len(ix2), dtype=np.int8))]
    with zipfile.ZipFile(path, mode="w", compression=zipfile.ZIP_DEFLATED, allowZip64=True) as zf:
        for name, arr in members:
            with zf.open(name + ".npy", mode="w", force_zip64=True) as fh:
                npy_format.write_array(fh, np.asanyarray(arr), allow_pickle=False)


def run_fedrann_pipeline(*, input_path: str, output_dir: str, kmer_size: int, kmer_sample_fraction: float,
                         kmer_min_multiplicity: int, embedding_dimension: int, nndescent_n_trees: int,
                         nndescent_n_neighbors: int, save_feature_matrix: bool, keep_intermediates: bool,
                         chunk_size: int, group=None):
    """run_fedrann_pipeline (__main__.py:302-392) with device-resident intermediates."""
    from .nearest_neighbors import check_knn_limits
    check_knn_limits(nndescent_n_neighbors, embedding_dimension)
    import torch
    import torch.distributed as dist
    from .count_kmers import current_session, run_kmer_searcher
    from .engine import write_overlaps_tsv
    from .parallel import all_gather_rows
    from .precompute import get_precompute_matrix

    rank = dist.get_rank(group) if group is not None else 0
    world = dist.get_world_size(group) if group is not None else 1
    t0 = time.perf_counter()
    logger.info("--- 1. Counter kmers ---")
    ks_path, n_features, read_count = run_kmer_searcher(
        input_path=input_path, k=kmer_size, sample_fraction=kmer_sample_fraction,
        min_multiplicity=kmer_min_multiplicity, group=group,
        write_files=True)
    ses = current_session()
    eng = ses.eng
    logger.debug(f"kmer_searcher n_features: {n_features}")
    if world > 1:
        dist.barrier(group)          # rank 0 has written fwd_kmer_library.fasta

    logger.info("--- 2. Generate dimension reduction and IDF matrix ---")
    fwd_kmer_library_path = join(global_variables.temp_dir, "fwd_kmer_library.fasta")
    precompute_mat, n_features = get_precompute_matrix(n_components=embedding_dimension,
                                                       counter_file=fwd_kmer_library_path, n_features=n_features)

    logger.info("--- 3. Generate feature matrix ---")
    E = eng.project(ses.indptr, ses.indices, eng.upload_projection(precompute_mat), ses.nk)

    logger.info("--- 4. Nearest Neighbors Search ---")
    logger.info("Exact cosine kNN on the tensor cores (n_neighbors = %d; n_trees = %d is not used)",
                nndescent_n_neighbors, nndescent_n_trees)
    from .pipeline import EXCHANGE, _rows_per_rank
    xch = None
    if world > 1 and EXCHANGE == "peer":
        from .parallel import row_exchange
        xch = row_exchange(eng, _rows_per_rank(E.shape[0], group, None), E.shape[1], group)
    En, Eh, norms = eng.normalize(E, out=xch.local() if xch is not None else None)
    q0, nq = 0, En.shape[0]
    if xch is not None:           # peer-memory exchange over NVLink, consumed by the search as it lands
        xch.pull()
        En, Eh, norms, q0 = xch.En, xch.Eh, xch.norms, xch.q0
    elif world > 1:
        En, Eh, norms, q0 = all_gather_rows(En, Eh, norms, group)
    k = min(nndescent_n_neighbors, En.shape[0])
    idx, dstn = eng.knn(En, Eh, norms, k, q0=q0, nq=nq) if nq > 0 and k > 0 else (
        torch.zeros((0, k), dtype=torch.int32), torch.zeros((0, k), dtype=torch.float32))
    if xch is not None:
        xch.release()
    idx_h, dist_h = idx.cpu().numpy(), dstn.cpu().numpy()
    if xch is not None:
        xch.close()
        eng._row_exchange = None

    nbr_output_file = join(output_dir, "overlaps.tsv")
    logger.debug("Saving overlap table to %s", nbr_output_file)
    for r in range(world):                 # rows in rank order = input order
        if r == rank:
            write_overlaps_tsv(nbr_output_file, ses.reads.names, idx_h, dist_h, row0=q0,
                               n_rows_total=2 * read_count, append=rank > 0)
        if world > 1:
            dist.barrier(group)

    if save_feature_matrix and rank == 0:
        from .feature_extraction import read_output_bin
        _, ip, ix = read_output_bin(ks_path)
        write_feature_matrix_npz(join(output_dir, "feature_matrix.npz"), ip, ix, ses.nk, n_features)
    if not keep_intermediates and rank == 0:
        logger.debug("Removing intermediate files")
        rmtree(global_variables.temp_dir)
    logger.info("Pipeline completed in %.2f s (%d reads, %d GPU(s)).", time.perf_counter() - t0, read_count, world)


def main(argv=None):
    args = parse_command_line_arguments(argv)
    global_variables.threads = args.threads
    global_variables.seed = args.seed
    from .parallel import init_from_env
    rank, world, _local = init_from_env()
    import torch.distributed as dist
    group = dist.group.WORLD if world > 1 else None

    output_dir = abspath(args.output_dir)
    os.makedirs(output_dir, exist_ok=True)
    global_variables.output_dir = output_dir
    if rank == 0:
        add_log_file(join(output_dir, "fedrann.log"))
    temp_dir = join(output_dir, "temp")
    os.makedirs(temp_dir, exist_ok=True)
    global_variables.temp_dir = temp_dir

    logger.info(f"FEDRANN version: {__version__}")
    logger.debug(f"Input file: {args.input}")
    logger.debug(f"Output directory: {output_dir}")
    logger.debug(f"Parameters: {args}")

    f = functools.partial(
        run_fedrann_pipeline, input_path=args.input, output_dir=output_dir, kmer_size=args.kmer_size,
        kmer_sample_fraction=args.kmer_sample_fraction, kmer_min_multiplicity=args.kmer_min_multiplicity,
        embedding_dimension=args.embedding_dimension, nndescent_n_trees=args.nndescent_n_trees,
        nndescent_n_neighbors=args.nndescent_n_neighbors, keep_intermediates=args.keep_intermediates,
        save_feature_matrix=args.save_feature_matrix, chunk_size=args.chunk_size, group=group)
    if args.mprof and rank == 0:
        # __main__.py:432-454: an mprof-format trace of the host process (psutil RSS, 1 s interval)
        import threading
        import psutil
        mprof_dir = join(output_dir, "mprof")
        os.makedirs(mprof_dir, exist_ok=True)
        samples, stop = [], threading.Event()
        proc = psutil.Process()

        def sampler():
            while not stop.is_set():
                samples.append((proc.memory_info().rss / 2 ** 20, time.time()))
                stop.wait(1.0)
        th = threading.Thread(target=sampler, daemon=True)
        th.start()
        try:
            f()
        finally:
            stop.set()
            th.join(2.0)
            with open(join(mprof_dir, "memory_profile.dat"), "wt") as fs:
                fs.write("MT 1.0\n")
                for mem, ts in samples:
                    fs.write(f"MEM {mem:.6f} {ts:.6f}\n")
    else:
        f()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
