"""GPU parity tests of the per-level engine entry points (through the C ABI) against the CPU oracle."""
import ctypes as C

import numpy as np
import pytest

import h2ot_lib
import oracle_lib as O

pytestmark = pytest.mark.gpu
NaN = float("nan")


def _engine(X, y=None, w=None, family="gaussian"):
    P = h2ot_lib.pkg()
    n, c = X.shape
    e = P.Engine(nrows=n, ncols=c, family=family, has_weights=w is not None)
    for j in range(c):
        e.load_column(j, X[:, j])
    e.load_response(np.zeros(n) if y is None else y)
    if w is not None:
        e.load_weights(w)
    e.finalize_frame()
    e.init_model([0.0])
    return P, e


def _oracle_hist(desc_args, cs, ys, ws=None, rows=None, int_path=False):
    h = O.Hist(*desc_args).init()
    if int_path:
        ci = np.where(np.isnan(cs), -2 ** 31, cs).astype(np.int32)
        h.update_int(ws, ci, ys, rows)
    else:
        h.update(ws, cs, ys, rows)
    h.reduce_precision()
    return h


def test_bin_codes_bit_exact():
    """DHistogram.bin (T/DHistogram.java:272-306): bin codes must be bit-exact, incl. NA, +-Inf, -0.0, float-split check."""
    rng = np.random.default_rng(1)
    n = 100003
    x = (rng.normal(size=n) * 3).astype(np.float32).astype(np.float64)
    x[:7] = [NaN, np.inf, -np.inf, -0.0, 0.0, x.min(), x.max()]
    x64 = rng.normal(size=n) * 1e3  # not fp32-exact -> F64 storage
    xi = rng.integers(-100, 100, n).astype(np.float64)
    xi[5] = NaN
    X = np.stack([x, x64, xi], axis=1)
    P, e = _engine(X)
    assert e.column_info(0).storage == P.STORE_F32
    assert e.column_info(1).storage == P.STORE_F64
    assert e.column_info(2).storage == P.STORE_U8
    for col, isint in ((0, 0), (1, 0), (2, 1)):
        v = X[:, col]
        lo, hi = np.nanmin(v[np.isfinite(v)]), np.nanmax(v[np.isfinite(v)])
        for nbins in (1024, 256, 20, 13):
            for cf in (False, True):
                maxEx = O.lib().orc_find_maxEx(hi, isint)
                oh = O.Hist(nbins, 1024, isint, lo, maxEx, checkFloat=cf).init()
                d = P.make_desc(lo, maxEx, oh.nbin, step=oh.step, is_int=isint, check_float=cf)
                got = e.bin_codes(col, d).astype(np.int32)
                want = oh.bin_many(v)
                assert np.array_equal(got, want), (col, nbins, cf, int(np.sum(got != want)))
    e.close()


@pytest.mark.parametrize("weighted", [False, True])
def test_root_level_histogram_matches_update_histo(weighted):
    """updateHisto / updateHistoInt + reducePrecision (T/DHistogram.java:609-704, 734-740) at the root, all storages;
    shared-memory and global-atomics kernels must agree bit for bit (integer accumulation commutes)."""
    rng = np.random.default_rng(2)
    n = 150001
    X = np.stack([
        rng.normal(size=n).astype(np.float32).astype(np.float64),
        rng.lognormal(size=n),
        rng.integers(0, 7, n).astype(np.float64),
        rng.integers(-500, 3000, n).astype(np.float64),
        rng.choice([0.0, 1.09, 2.17], n).astype(np.float32).astype(np.float64),
    ], axis=1)
    X[rng.random(X.shape) < 0.03] = NaN
    work = (rng.random(n) - 0.5).astype(np.float32).astype(np.float64)
    w = rng.integers(0, 4, n).astype(np.float64) if weighted else None
    P, e = _engine(X, w=w)
    if weighted:  # w == 0 rows are DECIDED_ROW after init_model (MarkDecidedRows)
        assert np.array_equal(e.get_nids() == P.DECIDED_ROW, w == 0)
    e.set_work(work)
    e.begin_tree(0, 0, 1.0, float(np.max(np.abs(work))))
    args, descs = [], (P.HistDesc * X.shape[1])()
    for c in range(X.shape[1]):
        v = X[:, c]
        isint = 1 if c in (2, 3) else 0
        lo, hi = np.nanmin(v), np.nanmax(v)
        maxEx = O.lib().orc_find_maxEx(hi, isint)
        intopt = c == 2
        a = (1024, 1024, isint, lo, maxEx, intopt, True)
        oh = O.Hist(*a)
        args.append(a)
        descs[c] = P.make_desc(lo, maxEx, oh.nbin, step=oh.step, is_int=isint, int_opt=intopt, init_na=True, order=c)
    results = []
    for mode in (0, 1):
        e.set_hist_mode(mode)
        e.build_level(descs)
        r = e.finish_level()
        results.append(r)
    assert np.array_equal(results[0]["vals"], results[1]["vals"])
    assert np.array_equal(results[0]["min2"], results[1]["min2"]) and np.array_equal(results[0]["maxIn"], results[1]["maxIn"])
    r = results[0]
    rows = np.arange(n, dtype=np.int32) if w is None else np.nonzero(w != 0)[0].astype(np.int32)
    for c in range(X.shape[1]):
        oh = _oracle_hist(args[c], X[:, c], work[rows], w, rows, int_path=args[c][5])  # ys are gathered in leaf order
        ov = oh.vals()
        gv = r["vals"][r["offsets"][c]: r["offsets"][c] + ov.size]
        assert r["has_vals"][c] == 1
        assert np.array_equal(gv[0::3], ov[0::3]), c
        for j in (1, 2):
            scale = np.max(np.abs(ov[j::3]))
            assert np.max(np.abs(gv[j::3] - ov[j::3])) <= 1e-6 * scale, (c, j)
        assert r["min2"][c] == oh.min2 and r["maxIn"][c] == oh.maxIn, (c, r["min2"][c], oh.min2)
        assert r["wNA"][c] == ov[-3]
    e.close()


def test_score_decide_and_two_leaf_level():
    """score_decide (T/ScoreBuildHistogram2.java:139-168) + getChildNodeID (T/DTree.java:728-763): numeric split with
    NA direction, then a categorical bitset split with out-of-range levels; nids bit-exact vs a numpy restatement."""
    rng = np.random.default_rng(3)
    n = 70001
    x0 = rng.normal(size=n).astype(np.float32).astype(np.float64)
    x0[rng.random(n) < 0.05] = NaN
    x1 = rng.integers(0, 40, n).astype(np.float64)
    x1[rng.random(n) < 0.05] = NaN
    X = np.stack([x0, x1], axis=1)
    work = rng.normal(size=n).astype(np.float32).astype(np.float64)
    P, e = _engine(X)
    e.set_work(work)
    e.begin_tree(0, 0, 1.0, float(np.max(np.abs(work))))
    nids = np.zeros(n, dtype=np.int32)
    nids[rng.random(n) < 0.1] = P.OUT_OF_BAG  # sampled-out rows keep their OOB encoding below the root (q1)
    e.set_nids(nids)

    def root_descs(nleaves):
        d = (P.HistDesc * (nleaves * 2))()
        for l in range(nleaves):
            d[l * 2 + 0] = P.make_desc(-6.0, 6.0, 64, init_na=True, order=0)
            d[l * 2 + 1] = P.make_desc(0.0, 40.0, 40, step=1.0, is_int=1, init_na=True, order=1)
        return d

    e.build_level(root_descs(1))
    r0 = e.finish_level()
    # at the root OOB rows ARE histogrammed
    assert r0["vals"][r0["offsets"][0]: r0["offsets"][0] + 65 * 3][0::3].sum() == n
    assert np.array_equal(e.get_nids(), nids)
    # level 1: root (nid 0) splits on col 0 at 0.25f, NAs left -> children 1, 2
    dn = (P.DecidedNode * 1)()
    splat = np.float32(0.25)
    dn[0] = P.DecidedNode(nid=0, pid=-1, col=0, equal=0, nasplit=P.NA_LEFT, splat=float(splat), child=(C.c_int32 * 2)(1, 2))
    e.build_level(root_descs(2), leaf_offset=1, decided=dn)
    r1 = e.finish_level()
    go_right = np.where(np.isnan(x0), False, x0 >= float(splat))
    want = np.where(go_right, 2, 1).astype(np.int32)
    oob = nids == P.OUT_OF_BAG
    want_enc = np.where(oob, -want + P.OUT_OF_BAG, want).astype(np.int32)
    got = e.get_nids()
    assert np.array_equal(got, want_enc)
    inbag = ~oob
    for l in (0, 1):
        cnt = r1["vals"][r1["offsets"][l * 2]: r1["offsets"][l * 2] + 65 * 3][0::3].sum()
        assert cnt == np.sum(inbag & (want == l + 1))
    # level 2: node 1 splits on the categorical col 1 by bitset (levels 5..20, bitoff 5): in-set -> right,
    # out-of-range -> NA direction (right here); node 2 does not split (rows stay, code skip)
    nbits, bitoff = 16, 5
    bits = rng.integers(0, 2, nbits)
    pool = np.zeros(2, dtype=np.uint8)
    for i, b in enumerate(bits):
        if b:
            pool[i >> 3] |= 1 << (i & 7)
    dn2 = (P.DecidedNode * 2)()
    dn2[0] = P.DecidedNode(nid=1, pid=0, col=1, equal=3, nasplit=P.NA_RIGHT, splat=NaN, child=(C.c_int32 * 2)(3, 4),
                           bs_off=0, bs_nbits=nbits, bs_bitoff=bitoff)
    dn2[1] = P.DecidedNode(nid=2, pid=0, col=-1, equal=0, nasplit=P.NA_NONE, splat=NaN, child=(C.c_int32 * 2)(-1, -1))
    # the parent's child pointer is cleared by doNotSplit (T/DTree.java:504-511)
    dn0 = (P.DecidedNode * 3)()
    dn0[0] = P.DecidedNode(nid=0, pid=-1, col=0, equal=0, nasplit=P.NA_LEFT, splat=float(splat), child=(C.c_int32 * 2)(1, -1))
    dn0[1], dn0[2] = dn2[0], dn2[1]
    e.build_level(root_descs(2), leaf_offset=3, decided=dn0, bitset_pool=pool)
    e.finish_level()
    got2 = e.get_nids()
    lvl = np.where(np.isnan(x1), -1, x1).astype(np.int64)
    idx = lvl - bitoff
    in_range = (~np.isnan(x1)) & (idx >= 0) & (idx < nbits)
    contains = np.zeros(n, dtype=bool)
    contains[in_range] = bits[idx[in_range]] == 1
    right = np.where(in_range, contains, True)
    child = np.where(right, 4, 3)
    want2 = np.where(want == 1, child, 0).astype(np.int32)  # rows of node 2 go back to the parent (nid 0)
    want2_enc = np.where(oob, -want2 + P.OUT_OF_BAG, want2).astype(np.int32)
    assert np.array_equal(got2, want2_enc)
    e.close()


def _split_tuple(s):
    return (s.col, s.bin, s.nasplit, s.equal, s.se, s.se0, s.se1, s.n0, s.n1, s.p0, s.p1, s.tree_p0, s.tree_p1)


def test_find_splits_matches_find_best_split_point():
    """findBestSplitPoint + splatNumeric + fillBitSet (T/DTree.java:984-1487, 253-291, 1907-1941): one active column
    at a time, so the per-leaf result is that column's candidate. Histograms come from the device level pass and are
    fed to the oracle unchanged, so every field must match bit for bit."""
    rng = np.random.default_rng(4)
    n = 40000
    X = np.stack([
        rng.normal(size=n).astype(np.float32).astype(np.float64),
        rng.integers(0, 30, n).astype(np.float64),     # categorical, sorted-set path
        rng.integers(0, 9, n).astype(np.float64),      # small int
        rng.lognormal(size=n).astype(np.float32).astype(np.float64),
    ], axis=1)
    X[rng.random(X.shape) < 0.04] = NaN
    X[:, 3] = np.where(np.isnan(X[:, 3]), 1.0, X[:, 3])  # no NAs in col 3
    y = (X[:, 0] > 0.3).astype(np.float64) + 0.1 * (np.nan_to_num(X[:, 1]) % 4) + rng.normal(size=n) * 0.3
    work = y.astype(np.float32).astype(np.float64)
    P, e = _engine(X)
    e.set_work(work)
    e.begin_tree(0, 0, 1.0, float(np.max(np.abs(work))))
    specs = [(0, 0, 64), (1, 2, 1024), (2, 1, 20), (3, 0, 20), (0, 0, 1024), (3, 0, 256)]
    for (c, isint, nbins) in specs:
        for check_float in ((False, True) if isint == 0 else (False,)):
            v = X[:, c]
            lo, hi = np.nanmin(v), np.nanmax(v)
            maxEx = 30.0 if isint == 2 else O.lib().orc_find_maxEx(hi, isint)
            lo = 0.0 if isint == 2 else lo
            args = (nbins, 1024, isint, lo, maxEx, False, True, 1e-5, 0, check_float)
            oh = O.Hist(*args).init()
            descs = (P.HistDesc * X.shape[1])()
            for j in range(X.shape[1]):
                descs[j] = P.make_desc(0.0, 1.0, 1, active=False)
            descs[c] = P.make_desc(lo, maxEx, oh.nbin, step=oh.step, is_int=isint, init_na=True, check_float=check_float, order=0)
            e.build_level(descs)
            r = e.finish_level()
            gv = r["vals"][: (oh.nbin + 1) * 3]
            oh.set_vals(gv)
            for min_rows in (1.0, 10.0, 5000.0):
                got = e.find_splits(min_rows, 1e-5)[0]
                want = oh.find_split(c, min_rows)
                if want is None:
                    assert got.col == -1, (c, nbins, min_rows)
                    continue
                want_splat = oh.splat_numeric(want.bin, want.nasplit) if (want.nasplit != O.NAvsREST and want.equal in (0, 1)) else -1.0
                if check_float and np.isnan(want_splat):  # bestCol rejects it (T/DTree.java:644-651)
                    assert got.col == -1
                    continue
                assert _split_tuple(got) == _split_tuple(want), (c, nbins, min_rows, _split_tuple(got), _split_tuple(want))
                assert np.float32(got.splat).tobytes() == np.float32(want_splat).tobytes()
                if want.equal >= 2:
                    assert (got.bs_nbits, got.bs_bitoff, got.bs_nbytes) == (want.bs_nbits, want.bs_bitoff, want.bs_nbytes)
                    assert bytes(got.bs[:got.bs_nbytes]) == bytes(want.bs[:want.bs_nbytes])
    e.close()


def test_dtree_test_golden_vector_on_device():
    """TT/DTreeTest.java:12-43 (PUBDEV-6495) replayed on the device: weights {10,10,100,1}, two NA rows, 2 bins.
    min_rows = 20 splits, 21 does not (min_split_improvement 0.01), 21 with improvement -9 splits with _se == SE of
    the non-NA rows and below the first split's _se."""
    ws = np.array([10.0, 10.0, 100.0, 1.0])
    cs = np.array([NaN, NaN, 0.5, 1.5])
    ys = np.array([0.0, 1.0, 0.0, 1.0])
    P, e = _engine(cs.reshape(-1, 1), w=ws)
    e.set_work(ys)
    e.begin_tree(0, 0, 1.0, 1.0)
    descs = (P.HistDesc * 1)()
    descs[0] = P.make_desc(0.0, 2.0, 2, init_na=True)
    e.build_level(descs)
    r = e.finish_level()
    assert list(r["vals"]) == [100.0, 0.0, 0.0, 1.0, 1.0, 1.0, 20.0, 10.0, 10.0]
    s1 = e.find_splits(20.0, 0.01)[0]
    assert s1.col == 0
    s2 = e.find_splits(21.0, 0.01)[0]
    assert s2.col == -1
    s3 = e.find_splits(21.0, -9.0)[0]
    assert s3.col == 0
    m = ~np.isnan(cs)
    mean = np.sum(ws[m] * ys[m]) / np.sum(ws[m])
    se_non_na = float(np.sum(ws[m] * (ys[m] - mean) ** 2))
    assert abs(s3.se - se_non_na) <= 1e-12 and s3.se < s1.se
    # and field for field against the oracle on the same histogram
    for (mr, msi) in ((20.0, 0.01), (21.0, -9.0)):
        oh = O.Hist(2, 2, 0, 0, 2, False, True, msi, 123).init()
        oh.update(ws, cs, ys)
        assert _split_tuple(e.find_splits(mr, msi)[0]) == _split_tuple(oh.find_split(0, mr))
    e.close()


def test_sampler_bit_exact():
    """T/Sample.java:22-36 on PCG (W/util/RandomUtils.java:50-119): OOB marks identical to the oracle's."""
    n = 50000
    X = np.arange(n, dtype=np.float64).reshape(-1, 1)
    P = h2ot_lib.pkg()
    for row_offset in (0, 123457):
        e = P.Engine(nrows=n, ncols=1, row_offset=row_offset)
        e.load_column(0, X[:, 0])
        y = np.zeros(n)
        y[::97] = NaN
        e.load_response(y)
        e.finalize_frame()
        e.init_model([0.0])
        seed = 0x1234ABCD5678
        e.begin_tree(0, seed, 0.632, 1.0)
        got = e.get_nids()
        want = np.zeros(n, dtype=np.uint8)
        O.lib().orc_sample(seed, 0.632, row_offset, n, O.ptr(want))
        want_nids = np.where(np.isnan(y), P.DECIDED_ROW, 0).astype(np.int32)
        want_nids = np.where((want == 1) | np.isnan(y), P.OUT_OF_BAG, want_nids)
        assert np.array_equal(got, want_nids)
        e.close()


@pytest.mark.parametrize("weighted,check_float,wide", [(False, False, False), (False, True, False), (True, True, False), (False, False, True)])
def test_sparse_level_kernels_agree(weighted, check_float, wide):
    """A level where every leaf histograms its own few columns (UndecidedNode.scoreCols, T/DTree.java:443-499; q8:
    T/ScoreBuildHistogram2.java:421-422): the leaf-major kernel (mode 0 picks it), the column-major shared-memory
    kernels (mode 2) and the L2-atomics kernel (mode 1) must give identical bits, and match updateHisto
    (T/DHistogram.java:609-658) per (leaf, column)."""
    rng = np.random.default_rng(11)
    n, L, per_leaf = (60001, 150, 3) if wide else (260001, 150, 3)
    cols = [rng.normal(size=n).astype(np.float32).astype(np.float64) for _ in range(5)]
    cols += [rng.lognormal(size=n), rng.normal(size=n) * 1e-3 + 7.0]                       # f64 storage
    cols += [rng.integers(0, 9, n).astype(np.float64), rng.integers(-3, 200, n).astype(np.float64)]      # u8
    cols += [rng.integers(0, 5000, n).astype(np.float64), rng.integers(-900, 30000, n).astype(np.float64)]  # u16
    cols += [rng.choice([0.0, 1.09, 2.17], n).astype(np.float32).astype(np.float64)]
    is_int = [0] * 7 + [1] * 4 + [0]
    if wide:  # more than 512 columns: the kernel reads the column table from global memory instead of shared memory
        for j in range(600):
            cols.append(rng.integers(0, 50, n).astype(np.float64) if j % 2 else rng.normal(size=n).astype(np.float32).astype(np.float64))
            is_int.append(1 if j % 2 else 0)
    X = np.stack(cols, axis=1)
    X[rng.random(X.shape) < 0.03] = NaN
    C_ = X.shape[1]
    work = (rng.random(n) - 0.5).astype(np.float32).astype(np.float64)
    w = rng.integers(0, 3, n).astype(np.float64) if weighted else None
    P, e = _engine(X, w=w)
    e.set_work(work)
    e.begin_tree(0, 0, 1.0, float(np.max(np.abs(work)) * (2.0 if weighted else 1.0)))
    # rows sit directly in leaves 1..L: a root-like decided node without a split keeps its rows (nid - leaf_offset)
    leaf_of = rng.integers(0, L, n).astype(np.int32)
    nids = (leaf_of + 1).astype(np.int32)
    skipped = rng.random(n) < 0.2
    nids[skipped] = P.DECIDED_ROW
    if weighted:
        nids[w == 0] = P.DECIDED_ROW
    e.set_nids(nids)
    dn = (P.DecidedNode * (L + 1))()
    for i in range(L + 1):
        dn[i] = P.DecidedNode(nid=i, pid=-1, col=-1, equal=0, nasplit=P.NA_NONE, splat=NaN, child=(C.c_int32 * 2)(-1, -1))
    descs = (P.HistDesc * (L * C_))()
    spec = {}
    for l in range(L):
        for j, c in enumerate(rng.choice(C_, per_leaf, replace=False)):
            v = X[:, c]
            lo, hi = np.nanmin(v), np.nanmax(v)
            maxEx = O.lib().orc_find_maxEx(hi, is_int[c])
            nbins = int(rng.choice([20, 64, 256]))
            a = (nbins, 1024, is_int[c], lo, maxEx, False, True)
            oh = O.Hist(*a)
            spec[(l, int(c))] = a
            descs[l * C_ + int(c)] = P.make_desc(lo, maxEx, oh.nbin, step=oh.step, is_int=is_int[c], init_na=True, order=j,
                                                 check_float=check_float and is_int[c] == 0)
    results, launches = [], []
    for mode in (0, 2, 1):
        e.set_hist_mode(mode)
        e.stats_reset()
        e.build_level(descs, leaf_offset=1, decided=dn)
        results.append(e.finish_level())
        launches.append(e.stats().sparse_launches)
        assert np.array_equal(e.get_nids(), nids)
    assert launches == [1, 0, 0]
    for r in results[1:]:
        assert np.array_equal(results[0]["vals"], r["vals"])
        assert np.array_equal(results[0]["min2"], r["min2"]) and np.array_equal(results[0]["maxIn"], r["maxIn"])
        assert np.array_equal(results[0]["has_vals"], r["has_vals"])
    r = results[0]
    live = nids != P.DECIDED_ROW
    for (l, c) in list(spec)[::7]:
        rows = np.nonzero(live & (leaf_of == l))[0].astype(np.int32)
        if check_float and is_int[c] == 0:
            continue  # the float-split correction is checked against the oracle in test_bin_codes_bit_exact
        oh = _oracle_hist(spec[(l, c)], X[:, c], work[rows], w, rows)
        ov = oh.vals()
        off = r["offsets"][l * C_ + c]
        gv = r["vals"][off: off + ov.size]
        assert np.array_equal(gv[0::3], ov[0::3]), (l, c)
        for j in (1, 2):
            scale = max(np.max(np.abs(ov[j::3])), 1e-300)
            assert np.max(np.abs(gv[j::3] - ov[j::3])) <= 1e-6 * scale, (l, c, j)
        if rows.size and np.any(~np.isnan(X[rows, c])):
            assert r["min2"][l * C_ + c] == oh.min2 and r["maxIn"][l * C_ + c] == oh.maxIn, (l, c)
    e.close()


def _sp_hist(kind, nbins, isint, lo, maxEx, check_float, seed, pts=None):
    """Oracle DHistogram with split points: kind 'random' (HT_Random), 'global' (QuantilesGlobal), 'custom' (UniformRobust)."""
    oh = O.Hist(nbins, 1024, isint, lo, maxEx, False, True, 1e-5, seed, check_float)
    oh.set_type({"random": O.HT_Random, "global": O.HT_QuantilesGlobal, "custom": O.HT_UniformRobust}[kind], pts)
    return oh.init()


def test_split_point_binning_levels_and_splits():
    """histogram_type Random / QuantilesGlobal / UniformRobust: DHistogram.bin through Arrays.binarySearch over _splitPts
    (T/DHistogram.java:281-286), binAt through the points (:308-312). Bin codes bit-exact vs the oracle; a level of several
    leaves built by the leaf-major kernel == the L2-atomics kernel == updateHisto; split search (splat through binAt) ==
    findBestSplitPoint."""
    rng = np.random.default_rng(31)
    n = 120001
    X = np.stack([
        rng.normal(size=n).astype(np.float32).astype(np.float64),
        rng.lognormal(size=n),
        rng.integers(-40, 4000, n).astype(np.float64),
        np.round(rng.normal(size=n), 1),   # many values exactly on decimal grid points, zeros included
    ], axis=1)
    X[rng.random(X.shape) < 0.03] = NaN
    C_ = X.shape[1]
    is_int = [0, 0, 1, 0]
    y = (np.nan_to_num(X[:, 0]) > 0.2).astype(np.float64) + 0.2 * np.nan_to_num(X[:, 3]) + rng.normal(size=n) * 0.3
    work = y.astype(np.float32).astype(np.float64)
    P, e = _engine(X)
    e.set_work(work)
    e.begin_tree(0, 0, 1.0, float(np.max(np.abs(work))))

    def hist_for(c, kind, nbins, check_float, seed):
        v = X[:, c]
        lo, hi = np.nanmin(v), np.nanmax(v)
        maxEx = O.lib().orc_find_maxEx(hi, is_int[c])
        pts = None
        if kind == "global":
            pts = np.unique(np.nanquantile(v, np.linspace(0, 1, nbins + 5)))
        if kind == "custom":
            pts = np.unique(np.concatenate([[lo], np.nanquantile(v, np.linspace(0.05, 0.95, nbins)), [maxEx]]))
        return _sp_hist(kind, nbins, is_int[c], lo, maxEx, check_float, seed, pts), lo, maxEx

    # ---- 1. bin codes
    for c in range(C_):
        for kind in ("random", "global", "custom"):
            for nbins in (20, 257):
                for cf in ((False, True) if is_int[c] == 0 else (False,)):
                    oh, lo, maxEx = hist_for(c, kind, nbins, cf, seed=1234 + c)
                    sp = oh.split_points()
                    assert sp.size > 1 and oh.absolute == (kind != "random")
                    e.set_split_points(np.concatenate([[123.0, -5.0], sp]))  # the pool may hold other nodes' points too
                    d = P.make_desc(lo, maxEx, oh.nbin, step=oh.step, is_int=is_int[c], check_float=cf, splitpts=(2, sp.size),
                                    absolute=oh.absolute)
                    got = e.bin_codes(c, d).astype(np.int32)
                    want = oh.bin_many(X[:, c])
                    assert np.array_equal(got, want), (c, kind, nbins, cf, int(np.sum(got != want)))

    # ---- 2. a level of L leaves, every (leaf, column) with its own split points; sparse kernel == global kernel == oracle
    L = 6
    leaf_of = rng.integers(0, L, n).astype(np.int32)
    nids = (leaf_of + 1).astype(np.int32)
    nids[rng.random(n) < 0.1] = P.DECIDED_ROW
    e.set_nids(nids)
    dn = (P.DecidedNode * (L + 1))()
    for i in range(L + 1):
        dn[i] = P.DecidedNode(nid=i, pid=-1, col=-1, equal=0, nasplit=P.NA_NONE, splat=NaN, child=(C.c_int32 * 2)(-1, -1))
    descs = (P.HistDesc * (L * C_))()
    pool, hists = [], {}
    kinds = ["random", "global", "custom", "uniform"]
    for l in range(L):
        for c in range(C_):
            kind = kinds[(l + c) % 4]
            cf = is_int[c] == 0 and (l % 2 == 1)
            if kind == "uniform":
                v = X[:, c]
                lo, hi = np.nanmin(v), np.nanmax(v)
                maxEx = O.lib().orc_find_maxEx(hi, is_int[c])
                oh = O.Hist(64, 1024, is_int[c], lo, maxEx, False, True, 1e-5, 0, cf).init()
                descs[l * C_ + c] = P.make_desc(lo, maxEx, oh.nbin, step=oh.step, is_int=is_int[c], init_na=True, check_float=cf, order=c)
            else:
                oh, lo, maxEx = hist_for(c, kind, 50, cf, seed=77 * l + c)
                sp = oh.split_points()
                off = sum(len(p) for p in pool)
                pool.append(sp)
                descs[l * C_ + c] = P.make_desc(lo, maxEx, oh.nbin, step=oh.step, is_int=is_int[c], init_na=True, check_float=cf, order=c,
                                                splitpts=(off, sp.size), absolute=oh.absolute)
            hists[(l, c)] = oh
    e.set_split_points(np.concatenate(pool))
    results = []
    for mode in (0, 1):
        e.set_hist_mode(mode)
        e.stats_reset()
        e.build_level(descs, leaf_offset=1, decided=dn)
        results.append(e.finish_level())
        assert e.stats().sparse_launches == (1 if mode == 0 else 0)
    e.set_hist_mode(0)
    assert np.array_equal(results[0]["vals"], results[1]["vals"])
    assert np.array_equal(results[0]["min2"], results[1]["min2"]) and np.array_equal(results[0]["maxIn"], results[1]["maxIn"])
    r = results[0]
    live = nids != P.DECIDED_ROW
    for (l, c), oh in hists.items():
        rows = np.nonzero(live & (leaf_of == l))[0].astype(np.int32)
        oh.update(None, X[:, c], work[rows], rows)
        oh.reduce_precision()
        ov = oh.vals()
        off = r["offsets"][l * C_ + c]
        gv = r["vals"][off: off + ov.size]
        assert np.array_equal(gv[0::3], ov[0::3]), (l, c)
        for j in (1, 2):
            assert np.max(np.abs(gv[j::3] - ov[j::3])) <= 1e-6 * max(np.max(np.abs(ov[j::3])), 1e-300), (l, c, j)
        assert r["min2"][l * C_ + c] == oh.min2 and r["maxIn"][l * C_ + c] == oh.maxIn

    # ---- 3. split search on one column at a time (root level): splat goes through binAt of the split points
    e.set_nids(np.zeros(n, dtype=np.int32))
    for c in (0, 1, 2, 3):
        for kind in ("random", "global", "custom"):
            for cf in ((False, True) if is_int[c] == 0 else (False,)):
                oh, lo, maxEx = hist_for(c, kind, 40, cf, seed=5 + c)
                sp = oh.split_points()
                e.set_split_points(sp)
                d1 = (P.HistDesc * C_)()
                for j in range(C_):
                    d1[j] = P.make_desc(0.0, 1.0, 1, active=False)
                d1[c] = P.make_desc(lo, maxEx, oh.nbin, step=oh.step, is_int=is_int[c], init_na=True, check_float=cf, order=0,
                                    splitpts=(0, sp.size), absolute=oh.absolute)
                e.build_level(d1)
                rr = e.finish_level()
                oh.set_vals(rr["vals"][: (oh.nbin + 1) * 3])
                for min_rows in (1.0, 50.0):
                    got = e.find_splits(min_rows, 1e-5)[0]
                    want = oh.find_split(c, min_rows)
                    if want is None:
                        assert got.col == -1, (c, kind, min_rows)
                        continue
                    want_splat = oh.splat_numeric(want.bin, want.nasplit) if (want.nasplit != O.NAvsREST and want.equal in (0, 1)) else -1.0
                    if cf and np.isnan(want_splat):
                        assert got.col == -1
                        continue
                    assert _split_tuple(got) == _split_tuple(want), (c, kind, cf, min_rows, _split_tuple(got), _split_tuple(want))
                    assert np.float32(got.splat).tobytes() == np.float32(want_splat).tobytes(), (c, kind, cf, got.splat, want_splat)
    e.close()
