"""h2ot_hist_column_noalloc — the entry point for embedders that cannot allocate device memory once running (a process
whose GPU memory belongs to its own pool, a latency-critical loop where cudaMalloc's device-wide synchronisation is not
acceptable, a plug-in host that forbids allocation after start-up). Its accumulators sit in a fixed device array inside
the library, shared by every call on that GPU, so the test makes two calls in a row with different inputs: the second
only comes out right if the call clears what the first one left behind.

Checked through the C ABI against the existing entry points (h2ot_build_level + h2ot_finish_level) on the same column,
with the tolerance the other parity tests use (w exact, wY / wYY within 1e-6 of the column scale).
This file sorts last on purpose: a run that stops at the first failure has already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

import h2ot_lib

pytestmark = pytest.mark.gpu


def _existing_entry_point(P, x, work, splat, descs2):
    """Level 1 under a root split `x >= splat` (NAs left), via h2ot_build_level / h2ot_finish_level."""
    n = len(x)
    e = P.Engine(nrows=n, ncols=1)
    e.load_column(0, x.astype(np.float64))
    e.load_response(np.zeros(n))
    e.finalize_frame()
    e.init_model([0.0])
    e.set_work(work.astype(np.float64))
    e.begin_tree(0, 0, 1.0, float(np.max(np.abs(work))))
    root = (P.HistDesc * 1)()
    root[0] = P.make_desc(-8.0, 8.0, 16)
    e.build_level(root)
    e.finish_level(want_vals=False)
    dn = (P.DecidedNode * 1)()
    dn[0] = P.DecidedNode(nid=0, pid=-1, col=0, equal=0, nasplit=P.NA_LEFT, splat=float(splat), child=(C.c_int32 * 2)(1, 2))
    e.build_level(descs2, leaf_offset=1, decided=dn)
    r = e.finish_level()
    assert r["has_vals"].all()
    e.close()
    return r["vals"]


def test_noalloc_histogram_matches_build_level_twice_in_a_row():
    import torch

    P = h2ot_lib.pkg()
    rng = np.random.default_rng(17)
    n = 200003
    splat = np.float32(0.125)
    descs = (P.HistDesc * 2)()
    descs[0] = P.make_desc(-8.0, 0.125, 128)
    descs[1] = P.make_desc(0.125, 8.0, 256)
    nvals = (128 + 1 + 256 + 1) * 3
    d_out = torch.full((nvals,), -1.0, dtype=torch.float64, device="cuda")
    for call in range(2):
        x = rng.normal(size=n).astype(np.float32)
        x[rng.random(n) < 0.03] = np.nan
        work = ((rng.random(n) - 0.5) * (1 + call)).astype(np.float32)  # second call: other data, other scale
        leaf = np.where(np.isnan(x), 0, (x >= splat).astype(np.int64)).astype(np.uint16)
        leaf[::1001] = 0xFFFF  # rows that take no part
        # the existing path has no "skip" input: drop those rows there by replaying it without them
        keep = leaf != 0xFFFF
        want = _existing_entry_point(P, x[keep], work[keep], splat, descs)
        d_x = torch.from_numpy(x).cuda()
        d_w = torch.from_numpy(work).cuda()
        d_l = torch.from_numpy(leaf.view(np.int16)).cuda()
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        P.hist_column_noalloc(d_x.data_ptr(), d_w.data_ptr(), d_l.data_ptr(), n, descs, float(np.max(np.abs(work[keep]))),
                              d_out.data_ptr())
        assert torch.cuda.memory_allocated() == before
        got = d_out.cpu().numpy()
        assert got.shape == want.shape
        assert np.array_equal(got[0::3], want[0::3]), "call %d: bin counts differ" % call
        for j in (1, 2):
            scale = float(np.max(np.abs(want[j::3])))
            assert float(np.max(np.abs(got[j::3] - want[j::3]))) <= 1e-6 * scale, (call, j)


def test_noalloc_rejects_sizes_beyond_its_scratch():
    import torch

    P = h2ot_lib.pkg()
    d = torch.zeros(16, dtype=torch.float32, device="cuda")
    out = torch.zeros(16, dtype=torch.float64, device="cuda")
    too_many = (P.HistDesc * 2)()
    too_many[0] = P.make_desc(0.0, 1.0, 40000)
    too_many[1] = P.make_desc(0.0, 1.0, 40000)
    with pytest.raises(P.H2OTError) as ei:
        P.hist_column_noalloc(d.data_ptr(), d.data_ptr(), 0, 16, too_many, 1.0, out.data_ptr())
    assert ei.value.code == -4 and "65536" in str(ei.value)
    too_many_leaves = (P.HistDesc * 65)()
    for i in range(65):
        too_many_leaves[i] = P.make_desc(0.0, 1.0, 4)
    with pytest.raises(P.H2OTError):
        P.hist_column_noalloc(d.data_ptr(), d.data_ptr(), 0, 16, too_many_leaves, 1.0, out.data_ptr())
