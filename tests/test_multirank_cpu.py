"""World-size-2 checks of the host-side sharding logic on CPU (gloo): the row shards the bench/driver cut are exactly the
slices of the global frame, the sampler keys rows by GLOBAL index, and the rendezvous used for the NCCL id works."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import h2ot_lib

ROOT = h2ot_lib.ROOT

WORKER = r'''
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, __ROOT__)
sys.path.insert(0, os.path.join(__ROOT__, "tests"))
import bench
import oracle_lib as O
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
N = 50001
r0, r1 = (N * rank) // world, (N * (rank + 1)) // world
cols, y, _ = bench.make_frame(N, r0, r1, pinned=False)
# id broadcast as bench.py / mgpu_check.py do it (payload stands in for ncclUniqueId)
uid = torch.zeros(128, dtype=torch.uint8)
if rank == 0:
    uid = torch.arange(128, dtype=torch.uint8)
dist.broadcast(uid, 0)
assert uid.tolist() == list(range(128))
# every rank's shard must be the slice of the frame rank 0 would build alone
full_cols, full_y, _ = bench.make_frame(N, 0, N, pinned=False)
for c in range(bench.NCOLS):
    assert np.array_equal(cols[c], full_cols[c][r0:r1]), c
assert np.array_equal(y, full_y[r0:r1])
# shards tile the row range
spans = [None] * world
dist.all_gather_object(spans, (r0, r1))
assert spans[0][0] == 0 and spans[-1][1] == N and all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
# the sampler is keyed by the global row index: shard-wise draws == slices of the global draw (T/Sample.java:30)
seed, rate = 0x5DEECE66D, 0.63
mine = np.zeros(r1 - r0, dtype=np.uint8)
O.lib().orc_sample(seed, rate, r0, r1 - r0, O.ptr(mine))
allr = np.zeros(N, dtype=np.uint8)
O.lib().orc_sample(seed, rate, 0, N, O.ptr(allr))
assert np.array_equal(mine, allr[r0:r1])
# max / sum merges used for rollups and the residual bound
t = torch.tensor([float(np.max(np.abs(y - 0.5))), float(len(y))], dtype=torch.float64)
mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
assert sm[1].item() == N and mx[0].item() == float(np.max(np.abs(full_y - 0.5)))
dist.barrier()
dist.destroy_process_group()
sys.stdout.write("RANK_OK_%d\n" % rank); sys.stdout.flush()
'''


def test_two_rank_sharding_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.replace("__ROOT__", repr(ROOT)))
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", str(script)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "RANK_OK_0" in r.stdout and "RANK_OK_1" in r.stdout


def test_bench_reference_arm_line():
    """bench.py --impl reference: one JSON line with the contract keys, timed on the oracle port (bounded sample)."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--rows", "60000", "--ref-rows", "60000",
                        "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "hist_build_rows_x_cols_per_s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["value"] == line["value"]


def test_bench_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--rows", "20000", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""
