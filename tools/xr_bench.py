#!/usr/bin/env python
"""Latency of the level exchange alone (h2ot_allreduce_level back to back, no histogram kernels in between, ranks in lockstep):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/xr_bench.py [--reps 50]
Run once as is (peer-memory exchange) and once with H2OT_XREDUCE=0 (NCCL). The sums are re-reduced every call; only time is read."""
import argparse
import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=50)
    ap.add_argument("--configs", default="1:1024,4:256,32:256")
    args = ap.parse_args()
    world, rank = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P = importlib.import_module("h2o-3_b200")
    uid = torch.zeros(P.COMM_ID_BYTES, dtype=torch.uint8)
    if rank == 0:
        uid = torch.from_numpy(P.comm_unique_id().copy())
    uid = uid.cuda()
    dist.broadcast(uid, 0)
    uid = uid.cpu().numpy()
    n, c = 200_000, 28
    rng = np.random.default_rng(rank)
    cols = [rng.random(n).astype(np.float32).astype(np.float64) for _ in range(c)]
    y = (rng.random(n) < 0.5).astype(np.float64)
    e = P.Engine(nrows=n, ncols=c, device=local, row_offset=rank * n, nrows_global=world * n)
    for j in range(c):
        e.load_column(j, cols[j])
    e.load_response(y)
    e.finalize_frame()
    e.comm_init(world, rank, uid)
    e.init_model([0.0])
    e.set_work((y - 0.5).astype(np.float32).astype(np.float64))
    e.begin_tree(0, 0, 1.0, 0.5)
    for cfg in args.configs.split(","):
        L, nb = (int(t) for t in cfg.split(":"))
        leaf0 = 0 if L == 1 else 1000
        dn = None
        if L > 1:
            e.set_nids((leaf0 + rng.integers(0, L, n)).astype(np.int32))
            dn = (P.DecidedNode * L)()
            for l in range(L):
                dn[l] = P.DecidedNode(nid=leaf0 + l, pid=-1, col=-1, equal=0, nasplit=0, splat=float("nan"), child=(C.c_int32 * 2)(-1, -1))
        else:
            e.set_nids(np.zeros(n, dtype=np.int32))
        descs = (P.HistDesc * (L * c))()
        for l in range(L):
            for j in range(c):
                descs[l * c + j] = P.make_desc(0.0, 1.0000001, nb, order=j)
        e.build_level(descs, leaf_offset=leaf0, decided=dn)
        for _ in range(5):
            e.allreduce_level()
        dist.barrier()
        torch.cuda.synchronize()
        e.stats_reset()
        for _ in range(args.reps):
            e.allreduce_level()
        s = e.stats()
        if rank == 0:
            words = L * c * (nb + 1) * 3 + L + L * c * 2
            print("world %d  L=%2d nbins=%4d  payload %.2f MB : %.1f us per exchange (%s)" % (
                world, L, nb, words * 8 / 1e6, 1e3 * s.ms_exchange / args.reps, "NCCL" if os.environ.get("H2OT_XREDUCE") == "0" else "peer memory"), flush=True)
    e.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
