"""Multi-GPU determinism check ("identical output for G in {1,2,4,8}"), run under torchrun:
every rank answers the WHOLE read set alone on its GPU, then the ranks answer it together (reads sharded,
embedding rows exchanged, query rows sharded) and each compares its rows bit for bit -- device-resident path
and host path, several steps in a row (buffer reuse across steps: the exchange's barriers and epochs).
    torchrun --nproc-per-node N scripts/multi_check.py [--reads-mod M] [--out file]
FEDRANN_EXCHANGE=peer|nccl selects the transport."""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrann_b200 import synthetic as S  # noqa: E402
from fedrann_b200.engine import Engine  # noqa: E402
from fedrann_b200.parallel import init_from_env, shard_reads  # noqa: E402
from fedrann_b200.pipeline import EXCHANGE, HostBuffers, run_hot_path, run_hot_path_host  # noqa: E402
from fedrann_b200.precompute import precompute_matrix  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome", type=int, default=400_000)
    ap.add_argument("--coverage", type=float, default=30)
    ap.add_argument("--D", type=int, default=192)
    ap.add_argument("--k", type=int, default=17)
    ap.add_argument("--knn-k", type=int, default=20)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, local = init_from_env()
    torch.cuda.set_device(local)
    group = dist.group.WORLD if world > 1 else None
    g = S.random_genome(a.genome, 11)
    rs = S.simulate_reads(g, a.coverage, 5, mean_len=6000, len_model="lognormal", min_len=300, sub=0.01, ins=0.005,
                          dele=0.005, n_frac=0.0005)
    R = rs.n_reads                      # not a multiple of the rank count in general: uneven shards
    lib = S.library_from_reads(rs, a.k, 0.05, 2, 5)
    P = precompute_matrix(a.D, lib.counts, lib.n_features)
    eng = Engine(local)
    eng.build_table(lib.all_codes(), a.k, sampler=(5, 0.05))
    P_dev = eng.upload_projection(P)
    # --- alone
    full = run_hot_path(eng, eng.dev(rs.bases), eng.dev(rs.offsets), P_dev, lib.nk, a.knn_k)
    torch.cuda.synchronize()
    # --- together
    b = shard_reads(rs.offsets, world)
    r0, r1 = int(b[rank]), int(b[rank + 1])
    rows = [2 * int(b[i + 1] - b[i]) for i in range(world)]
    o = rs.offsets[r0:r1 + 1] - rs.offsets[r0]
    bases_d = eng.dev(rs.bases[rs.offsets[r0]:rs.offsets[r1]].copy())
    offs_d = eng.dev(o)
    ok, why = True, []
    for step in range(a.steps):
        res = run_hot_path(eng, bases_d, offs_d, P_dev, lib.nk, a.knn_k, group=group, rows_per_rank=rows)
        torch.cuda.synchronize()
        same = (res.row_offset == 2 * r0 and torch.equal(res.nbr_idx, full.nbr_idx[2 * r0:2 * r1]) and
                torch.equal(res.nbr_dist.view(torch.int32), full.nbr_dist[2 * r0:2 * r1].view(torch.int32)) and
                torch.equal(res.embedding, full.embedding[2 * r0:2 * r1]))
        if not same:
            ok = False
            why.append("device path step %d" % step)
    hb = HostBuffers.from_device(bases_d, o, 2 * (r1 - r0), min(a.knn_k, 2 * R))
    for step in range(2):
        hres = run_hot_path_host(eng, hb, P_dev, lib.nk, a.knn_k, group=group, rows_per_rank=rows)
        same = (torch.equal(hb.nbr_idx.to(bases_d.device), full.nbr_idx[2 * r0:2 * r1]) and
                torch.equal(hb.nbr_dist.to(bases_d.device).view(torch.int32), full.nbr_dist[2 * r0:2 * r1].view(torch.int32)))
        if not same:
            ok = False
            why.append("host path step %d" % step)
        del hres
    flag = torch.tensor([1 if ok else 0], device=bases_d.device)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    line = {"world": world, "exchange": EXCHANGE if world > 1 else "none", "reads": R, "rows_per_rank": rows,
            "ok": bool(flag.item()), "rank": rank, "why": why, "fallback_rows": eng.knn_fallback_rows}
    if rank == 0:
        print(json.dumps(line), flush=True)
        if a.out:
            open(a.out, "w").write(json.dumps(line))
    if not ok:
        sys.stderr.write("rank %d: %s\n" % (rank, why))
    x = getattr(eng, "_row_exchange", None)
    if x:
        x[1].close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0 if flag.item() else 1


if __name__ == "__main__":
    sys.exit(main())
