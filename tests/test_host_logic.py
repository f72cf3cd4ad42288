"""CPU tests of the host driver's bookkeeping (h2o-3_b200/csrc/host_gbm.cpp) against the oracle's restatement of the same
reference code — no device involved: the DHistogram constructor rule, find_maxEx, the PCG generator and the per-node mtries
column draws. (Everything else the driver does needs the engine and is covered by the -m gpu parity tests.)"""
import ctypes as C

import numpy as np
import pytest

import h2ot_lib
import oracle_lib as O


def test_dhistogram_ctor_rule_matches_oracle():
    P = h2ot_lib.pkg()
    L = P.load()
    rng = np.random.default_rng(0)
    cases = [(20, 1024, 0, 0.0, 1.0), (1024, 1024, 1, -3.0, 40.0), (20, 1024, 1, 0.0, 5000.0), (20, 1024, 2, 0.0, 12.0),
             (20, 1024, 2, 0.0, 5000.0), (256, 1024, 0, 1.0, 6.900000000000001), (2, 2, 0, 0.0, 2.0), (20, 1024, 1, 0.0, 20.0),
             (20, 1024, 1, 0.5, 21.0), (64, 64, 2, 0.0, 64.0), (64, 64, 2, 0.0, 65.0)]
    for _ in range(300):
        isint = int(rng.integers(0, 3))
        lo = float(np.float32(rng.normal() * 10.0 ** int(rng.integers(-3, 6))))
        span = float(abs(rng.normal()) * 10.0 ** int(rng.integers(-6, 6))) + 1e-9
        if isint:
            lo, span = float(np.floor(lo)), float(np.ceil(span) + 1)
        cases.append((int(rng.choice([2, 13, 20, 64, 256, 1024])), int(rng.choice([16, 64, 1024])), isint, lo, lo + span))
    for nbins, ncats, isint, mn, mx in cases:
        nb, st = C.c_int32(), C.c_double()
        ok = L.h2ot_hostcheck_hist(nbins, ncats, isint, mn, mx, C.byref(nb), C.byref(st))
        h = O.lib().orc_hist_new(nbins, ncats, isint, mn, mx, 0, 0, 0.0, 0, 0)
        assert bool(ok) == bool(h), (nbins, ncats, isint, mn, mx)
        if h:
            assert (nb.value, st.value) == (O.lib().orc_hist_nbin(h), O.lib().orc_hist_step(h)), (nbins, ncats, isint, mn, mx)
            O.lib().orc_hist_free(h)
    for v, isint in [(1.0, 0), (1.0, 1), (6.9, 0), (1e300, 0), (np.finfo(np.float64).max, 0), (0.0, 1), (-5.5, 0), (2.0 ** 60, 1)]:
        assert L.h2ot_hostcheck_find_max_ex(v, isint) == O.lib().orc_find_maxEx(v, isint)


def test_pcg_matches_oracle():
    L = h2ot_lib.pkg().load()
    for seed in (0, 1, 1234, -77, 2 ** 62 + 12345):
        for nth in (0, 1, 5):
            assert L.h2ot_hostcheck_pcg_next_long(seed, nth) == O.lib().orc_pcg_next_long(seed, nth)
        for bound in (1, 2, 7, 28, 1000, 2 ** 31 - 1):
            a = np.zeros(64, dtype=np.int32)
            b = np.zeros(64, dtype=np.int32)
            L.h2ot_hostcheck_pcg_bounded(seed, bound, 64, a.ctypes.data_as(C.c_void_p))
            O.lib().orc_pcg_ints(seed, bound, 64, O.ptr(b))
            assert np.array_equal(a, b), (seed, bound)


def test_per_node_column_draws_match_oracle():
    """UndecidedNode.scoreCols / DTree ctor column sampling (T/DTree.java:60-86, 462-499): same draws, same order."""
    L = h2ot_lib.pkg().load()
    rng = np.random.default_rng(4)
    for ncols, mtry, mpt, change, depth in [(28, 5, 28, 1.0, 0), (28, 28, 28, 1.0, 3), (10, 3, 7, 1.0, 2), (54, 7, 40, 0.9, 4),
                                            (8, 1, 8, 1.0, 0), (2000, 44, 2000, 1.0, 1), (12, 12, 6, 1.1, 2)]:
        present = (rng.random(ncols) < 0.9).astype(np.uint8)
        present[0] = 1
        n_nodes = 9
        a = np.zeros((n_nodes, ncols), dtype=np.int32)
        an = np.zeros(n_nodes, dtype=np.int32)
        b = np.zeros((n_nodes, ncols), dtype=np.int32)
        bn = np.zeros(n_nodes, dtype=np.int32)
        L.h2ot_hostcheck_score_cols(ncols, mtry, mpt, 987654321, change, depth, present.ctypes.data_as(C.c_void_p), n_nodes,
                                    a.ctypes.data_as(C.c_void_p), an.ctypes.data_as(C.c_void_p))
        O.lib().orc_tree_score_cols(ncols, mtry, mpt, 987654321, change, depth, O.ptr(present), n_nodes, O.ptr(b), O.ptr(bn))
        assert np.array_equal(an, bn) and np.array_equal(a, b), (ncols, mtry, mpt)


@pytest.mark.parametrize("world,n,n_sums", [(2, 11, 7), (3, 100, 64), (8, 8, 8), (8, 86_101, 86_000), (5, 1000, 0), (16, 4099, 4099)])
def test_pushed_line_exchange_layout(world, n, n_sums):
    """The slot arithmetic of xr_ll_kernel (csrc/k_xreduce.cuh), restated: word i belongs to owner o(i) = the slice [n p / W, n (p+1) / W)
    it falls in; rank r pushes it to the owner's reduce-scatter area at line r * ceil(n / W) + (i - lo(o)); the owner publishes the
    reduced word at line i of every peer's all-gather area. No two writers may share a line, every line read must have been
    written, everything must fit the area (n + W lines), and the result must be the sum (words < n_sums) / max (the rest) over ranks."""
    rng = np.random.default_rng(world * 1000 + n)
    data = rng.integers(0, 1 << 40, size=(world, n), dtype=np.uint64)
    lo = [n * p // world for p in range(world + 1)]
    sl = (n + world - 1) // world
    cap_lines = n + world
    rs = [dict() for _ in range(world)]   # per owner: line -> value
    ag = [dict() for _ in range(world)]   # per reader: line -> value
    owner = np.searchsorted(np.array(lo[1:world]), np.arange(n), side="right")  # sum over q of (i >= lo[q]), as the kernel counts
    for i in range(n):
        assert lo[owner[i]] <= i < lo[owner[i] + 1]
    # A: push
    for r in range(world):
        for i in range(n):
            o = int(owner[i])
            if o == r:
                continue
            line = r * sl + (i - lo[o])
            assert 0 <= line < cap_lines
            assert line not in rs[o], "two senders share a reduce-scatter line"
            rs[o][line] = data[r, i]
    # B: reduce + publish
    out = [data[r].copy() for r in range(world)]
    for r in range(world):
        for i in range(lo[r], lo[r + 1]):
            acc = int(data[r, i])
            for p in range(world):
                if p == r:
                    continue
                v = int(rs[r][p * sl + (i - lo[r])])
                acc = (acc + v) & 0xFFFFFFFFFFFFFFFF if i < n_sums else max(acc, v)
            out[r][i] = acc
            for p in range(world):
                if p != r:
                    assert i < cap_lines and i not in ag[p]
                    ag[p][i] = acc
    # C: gather
    for r in range(world):
        for i in range(n):
            if not (lo[r] <= i < lo[r + 1]):
                out[r][i] = ag[r][i]
    want = np.where(np.arange(n) < n_sums, data.sum(axis=0, dtype=np.uint64), data.max(axis=0))
    for r in range(world):
        assert np.array_equal(out[r], want), r
