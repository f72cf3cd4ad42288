"""Multi-GPU parity check (run under torchrun on N GPUs of one box; not collected by pytest):
   python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_check.py
Every rank builds the same GBM on its row shard (NCCL allreduce of the integer histograms per level); rank 0 also builds
it alone on the full frame. Trees must be bit-identical across ranks AND to the single-GPU build; predictions of each
shard must equal the matching slice of the single-GPU predictions."""
import hashlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import h2ot_lib  # noqa: E402


def tree_bytes(gm, K=1):
    out = []
    for t in range(gm.ntrees):
        for k in range(K):
            nodes, ls = gm.tree(t, k)
            out.append(bytes(nodes))
            out.append(ls.to_bytes(4, "little"))
    return b"".join(out)


def main():
    world = int(os.environ["WORLD_SIZE"])
    rank = int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P = h2ot_lib.pkg()
    uid = torch.zeros(P.COMM_ID_BYTES, dtype=torch.uint8)
    if rank == 0:
        uid = torch.from_numpy(P.comm_unique_id().copy())
    uid = uid.cuda()
    dist.broadcast(uid, 0)
    uid = uid.cpu().numpy()
    ok = True
    for case, kw in (("bernoulli+sampling", dict(distribution="bernoulli", ntrees=4, max_depth=5, nbins=20, sample_rate=0.7, col_sample_rate=0.8, seed=11)),
                     ("gaussian+weights", dict(distribution="gaussian", nclass=1, ntrees=3, max_depth=4, nbins=32, seed=5)),
                     ("multinomial", dict(distribution="multinomial", nclass=3, ntrees=2, max_depth=4, nbins=20, seed=7)),
                     ("drf mtries+oob", dict(_cls="DRF", distribution="bernoulli", ntrees=3, max_depth=8, mtries=2, sample_rate=0.63, seed=9)),
                     ("random histograms", dict(distribution="bernoulli", ntrees=3, max_depth=4, nbins=20, histogram_type="Random", seed=3))):
        kw = dict(kw)
        Model = getattr(P, kw.pop("_cls", "GBM"))
        n = 120001
        X, y, cats = h2ot_lib.make_frame(n=n, c=6, seed=3)
        w = None
        K = 1
        if case.startswith("gaussian"):
            y = (np.nan_to_num(X[:, 0]) * 2 + np.nan_to_num(X[:, 2]) ** 2).astype(np.float32).astype(np.float64)
            w = np.random.default_rng(1).integers(0, 3, n).astype(np.float64)
        if case == "multinomial":
            y = (np.argmax(np.nan_to_num(X[:, :3]), axis=1)).astype(np.float64)
            K = 3
        r0, r1 = (n * rank) // world, (n * (rank + 1)) // world
        gm = Model(device=local, **kw)
        gm.load([X[r0:r1, c] for c in range(X.shape[1])], y[r0:r1], weights=None if w is None else w[r0:r1], categorical=cats,
                row_offset=r0, nrows_global=n, comm=(world, rank, uid))
        gm.train()
        mine = tree_bytes(gm, K)
        sizes = [None] * world
        dist.all_gather_object(sizes, (len(mine), hashlib.sha256(mine).hexdigest()))
        same = all(s == sizes[0] for s in sizes)
        preds = [gm.predictions(k) for k in range(K)]
        single_ok = True
        if rank == 0:
            g1 = Model(device=local, **kw)
            g1.load([X[:, c] for c in range(X.shape[1])], y, weights=w, categorical=cats)
            g1.train()
            single_ok = tree_bytes(g1, K) == mine
            for k in range(K):
                single_ok = single_ok and np.array_equal(g1.predictions(k)[r0:r1], preds[k])
            g1.close()
        flag = torch.tensor([1 if (same and single_ok) else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print("%-20s world=%d identical_across_ranks=%s identical_to_single_gpu=%s" % (case, world, same, single_ok), flush=True)
        ok = ok and bool(flag.item())
        gm.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MGPU_CHECK", "PASS" if ok else "FAIL", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
