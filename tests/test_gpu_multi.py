"""Row-sharded multi-GPU builds against the single-GPU build, bit for bit -- the reference's own protocol for its
distributed mode (h2o-algos/src/test/java/hex/tree/gbm/GBMTest.java:1085-1143: 1 node vs 5 nodes, same model).

Spawns tests/mgpu_check.py under torch.distributed.run on 2 GPUs (and on all of them when the box has more), once with
the peer-memory exchange of the level histograms (default) and once with the NCCL all-reduce (H2OT_XREDUCE=0). Every rank's
trees must equal every other rank's and the trees of a single-GPU build of the whole frame; every shard's predictions
must equal the matching slice of the single-GPU predictions. Skipped, with the reason printed, on a one-GPU box.
"""
import os
import subprocess
import sys

import pytest

import h2ot_lib

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def _worlds():
    n = h2ot_lib.pkg().device_count() if h2ot_lib.has_gpu() else 0
    w = [2] if n >= 2 else []
    if n > 2:
        w.append(n)
    return n, w


@pytest.mark.parametrize("exchange", ["peer-memory", "nccl"])
def test_sharded_build_is_bit_identical_to_single_gpu(exchange):
    n, worlds = _worlds()
    if not worlds:
        pytest.skip("multi-GPU parity needs >= 2 GPUs on the box; this one has %d" % n)
    for i, world in enumerate(worlds):
        env = dict(os.environ)
        env.pop("H2OT_XREDUCE", None)
        if exchange == "nccl":
            env["H2OT_XREDUCE"] = "0"
        port = 29600 + 10 * i + (1 if exchange == "nccl" else 0)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
               "--master-port", str(port), os.path.join(HERE, "mgpu_check.py")]
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
        sys.stdout.write(r.stdout[-4000:])
        assert r.returncode == 0, "world %d (%s): rc %d\n%s\n%s" % (world, exchange, r.returncode, r.stdout[-3000:], r.stderr[-3000:])
        assert "MGPU_CHECK PASS" in r.stdout, r.stdout[-3000:]
