"""Edge cases of the level pass and the C-ABI error behaviour (GPU), mirroring what the reference's unit tests poke at:
all-NA and constant columns, tiny and ragged row counts, zero weights, +-Inf values, two-bin histograms, leaves that
receive no rows, min_rows larger than the data, and invalid arguments (status codes + messages, never a crash)."""
import ctypes as C

import numpy as np
import pytest

import h2ot_lib
import oracle_lib as O

pytestmark = pytest.mark.gpu
NaN = float("nan")


def _both(X, y, cats=None, w=None, **kw):
    P = h2ot_lib.pkg()
    dist = kw.pop("distribution", "gaussian")
    nclass = 2 if dist == "bernoulli" else 1
    fam = {"gaussian": O.GAUSSIAN, "bernoulli": O.BERNOULLI}[dist]
    om = O.Model(O.Frame(X, y, w=w, cat_cards=cats), O.default_params(family=fam, nclass=nclass, **kw), capture=True)
    gm = P.GBM(distribution=dist, nclass=nclass, **kw)
    gm.load([X[:, c] for c in range(X.shape[1])], y, weights=w, categorical=cats)
    gm.capture_levels(True)
    gm.train()
    st = h2ot_lib.compare_models(gm, om)
    gm.close()
    return st


def test_all_na_and_constant_columns_are_ignored():
    rng = np.random.default_rng(1)
    n = 5000
    X = np.stack([rng.normal(size=n).astype(np.float32).astype(np.float64), np.full(n, NaN), np.full(n, 3.25),
                  np.where(rng.random(n) < 0.5, 7.0, NaN)], axis=1)  # col 3: constant + NAs -> NAvsREST candidate only
    y = (X[:, 0] + np.where(np.isnan(X[:, 3]), 1.0, 0.0) + rng.normal(size=n) * 0.1).astype(np.float32).astype(np.float64)
    _both(X, y, ntrees=2, max_depth=3, nbins=20, seed=1)


@pytest.mark.parametrize("n", [1, 2, 3, 31, 4095, 4096, 4097])
def test_tiny_and_ragged_row_counts(n):
    rng = np.random.default_rng(n)
    X = rng.normal(size=(n, 3)).astype(np.float32).astype(np.float64)
    y = (X[:, 0] > 0).astype(np.float64) + 0.5 * X[:, 1]
    y = y.astype(np.float32).astype(np.float64)
    _both(X, y, ntrees=2, max_depth=3, nbins=8, min_rows=1.0, seed=3)


def test_zero_weights_na_responses_and_infinite_values():
    rng = np.random.default_rng(5)
    n = 9000
    X = rng.normal(size=(n, 4)).astype(np.float32).astype(np.float64)
    X[::97, 1] = np.inf
    X[::89, 1] = -np.inf
    X[::50, 2] = NaN
    w = rng.integers(0, 3, n).astype(np.float64)
    w[:100] = 0
    y = (X[:, 0] * 2 + np.nan_to_num(X[:, 2])).astype(np.float32).astype(np.float64)
    y[::41] = NaN
    _both(X, y, w=w, ntrees=2, max_depth=4, nbins=16, min_rows=2.0, seed=9)


def test_two_bins_and_min_rows_above_the_data():
    rng = np.random.default_rng(6)
    n = 3000
    X = rng.normal(size=(n, 3)).astype(np.float32).astype(np.float64)
    y = (X[:, 0] > 0.2).astype(np.float64)
    _both(X, y, distribution="bernoulli", ntrees=2, max_depth=3, nbins=2, nbins_top_level=2, seed=2)
    st = _both(X, y, distribution="bernoulli", ntrees=1, max_depth=3, nbins=20, min_rows=5000.0, seed=2)  # root cannot split
    assert st["levels"] == 1


def test_level_with_an_empty_leaf_and_single_valued_leaf():
    """A leaf that receives no rows keeps _vals == null (has_vals 0) and cannot split; a leaf whose rows all share one value
    reports min2 == maxIn."""
    P = h2ot_lib.pkg()
    n = 20000
    x = np.concatenate([np.zeros(n // 2), np.full(n // 2, 5.0)])
    X = np.stack([x, np.arange(n, dtype=np.float64)], axis=1)
    e = P.Engine(nrows=n, ncols=2)
    for j in range(2):
        e.load_column(j, X[:, j])
    e.load_response(np.zeros(n))
    e.finalize_frame()
    e.init_model([0.0])
    e.set_work(np.where(x > 1, 1.0, -1.0))
    e.begin_tree(0, 0, 1.0, 1.0)
    root = (P.HistDesc * 2)()
    root[0] = P.make_desc(0.0, 6.0, 6, step=1.0, is_int=1)
    root[1] = P.make_desc(0.0, float(n), 20, is_int=1)
    e.build_level(root)
    e.finish_level()
    # split on col 0 at 100f: every row goes left, the right child (nid 2) gets nothing
    dn = (P.DecidedNode * 1)()
    dn[0] = P.DecidedNode(nid=0, pid=-1, col=0, equal=0, nasplit=P.LEFT, splat=100.0, child=(C.c_int32 * 2)(1, 2))
    lv = (P.HistDesc * 4)()
    for l in range(2):
        lv[l * 2 + 0] = P.make_desc(0.0, 6.0, 6, step=1.0, is_int=1)
        lv[l * 2 + 1] = P.make_desc(0.0, float(n), 20, is_int=1, order=1)
    e.build_level(lv, leaf_offset=1, decided=dn)
    r = e.finish_level()
    assert list(r["has_vals"]) == [1, 1, 0, 0]
    assert r["min2"][0] == 0.0 and r["maxIn"][0] == 5.0
    assert r["min2"][2] == np.finfo(np.float64).max and r["maxIn"][2] == -np.finfo(np.float64).max  # untouched sentinels
    s = e.find_splits(10.0, 1e-5)
    assert s[0].col == 0 and s[1].col == -1
    # descend again: left child splits on col 0 between the two values -> two pure leaves, each single valued
    dn2 = (P.DecidedNode * 2)()
    dn2[0] = P.DecidedNode(nid=1, pid=0, col=0, equal=0, nasplit=P.LEFT, splat=float(s[0].splat), child=(C.c_int32 * 2)(3, 4))
    dn2[1] = P.DecidedNode(nid=2, pid=0, col=-1, equal=0, nasplit=P.NA_NONE, splat=NaN, child=(C.c_int32 * 2)(-1, -1))
    e.build_level(lv, leaf_offset=3, decided=dn2)
    r2 = e.finish_level()
    assert r2["min2"][0] == r2["maxIn"][0] == 0.0 and r2["min2"][2] == r2["maxIn"][2] == 5.0
    assert e.find_splits(10.0, 1e-5)[0].col in (-1, 1)  # col 0 is constant inside the leaf; only col 1 may still split
    e.close()


def test_error_codes_and_messages():
    P = h2ot_lib.pkg()
    L = P.load()
    with pytest.raises(P.H2OTError) as ei:
        P.Engine(nrows=0, ncols=1)
    assert ei.value.code == -1
    e = P.Engine(nrows=100, ncols=2)
    with pytest.raises(P.H2OTError) as ei:
        e.load_column(5, np.zeros(100))
    assert ei.value.code == -1 and "column" in str(ei.value)
    with pytest.raises(P.H2OTError):
        e.load_column(0, np.zeros(101))  # more rows than the engine holds
    e.load_column(0, np.zeros(100))
    with pytest.raises(P.H2OTError) as ei:
        e.finalize_frame()  # column 1 never loaded
    assert "never loaded" in str(ei.value)
    e.load_column(1, np.arange(100.0))
    e.load_response(np.zeros(100))
    descs = (P.HistDesc * 2)()
    descs[0] = P.make_desc(0.0, 1.0, 4)
    descs[1] = P.make_desc(0.0, 100.0, 4)
    with pytest.raises(P.H2OTError) as ei:
        e.build_level(descs)  # frame not finalized
    assert "finalized" in str(ei.value)
    e.finalize_frame()
    e.init_model([0.0])
    with pytest.raises(P.H2OTError):
        e.find_splits(1.0, 0.0)  # no level built
    bad = (P.HistDesc * 2)()
    bad[0] = P.make_desc(0.0, 1.0, 4, step=-1.0)
    bad[1] = P.make_desc(0.0, 100.0, 4)
    with pytest.raises(P.H2OTError) as ei:
        e.build_level(bad)
    assert ei.value.code == -1 and "step" in str(ei.value)
    dn = (P.DecidedNode * 1)()
    dn[0] = P.DecidedNode(nid=0, pid=-1, col=9, equal=0, nasplit=0, splat=0.0, child=(C.c_int32 * 2)(1, 2))
    with pytest.raises(P.H2OTError) as ei:
        e.build_level(descs, leaf_offset=1, decided=dn)
    assert "column" in str(ei.value)
    assert L.h2ot_destroy(None) == 0
    e.close()
    with pytest.raises(P.H2OTError):
        P.GBM(distribution="bernoulli", nbins_cats=100000).load([np.zeros(10)], np.zeros(10))


def test_two_half_level_step_protocol():
    """h2ot_level_step_begin / _end: one outstanding begin per class-tree, end only after begin; two class-trees side by side give
    what the one-call level step gives."""
    P = h2ot_lib.pkg()
    L = P.load()
    rng = np.random.default_rng(8)
    n = 5000
    X = rng.normal(size=(n, 2)).astype(np.float32).astype(np.float64)
    e = P.Engine(nrows=n, ncols=2, nclass_trees=2, nclass=3, family="multinomial")
    for c in range(2):
        e.load_column(c, X[:, c])
    e.load_response(rng.integers(0, 3, n).astype(np.float64))
    e.finalize_frame()
    e.init_model([0.0, 0.0])
    for k in range(2):
        e.set_work(rng.normal(size=n).astype(np.float32).astype(np.float64), k=k)
        e.begin_tree(k, 0, 1.0, 5.0)
    descs = (P.HistDesc * 2)()
    descs[0] = P.make_desc(-6.0, 6.0, 32, init_na=True, order=0)
    descs[1] = P.make_desc(-6.0, 6.0, 16, init_na=True, order=1)
    LC = 2

    def end(k):
        m2, mi, wn = np.zeros(LC), np.zeros(LC), np.zeros(LC)
        hv = np.zeros(LC, dtype=np.uint8)
        sp = (P.Split * 1)()
        rc = L.h2ot_level_step_end(e.h, k, m2.ctypes.data, mi.ctypes.data, wn.ctypes.data, hv.ctypes.data, C.cast(sp, C.c_void_p))
        return rc, (m2.tobytes(), mi.tobytes(), wn.tobytes(), hv.tobytes(), bytes(sp))

    def begin(k):
        return L.h2ot_level_step_begin(e.h, k, 0, 1, 0, None, None, 0, C.cast(descs, C.c_void_p), 1.0, 1e-5)

    assert end(0)[0] == -1                      # nothing begun
    assert begin(0) == 0 and begin(0) == -1     # second begin of the same class-tree
    assert begin(1) == 0
    r0, a0 = end(0)
    r1, a1 = end(1)
    assert r0 == 0 and r1 == 0 and end(1)[0] == -1
    assert a0 != a1                             # different work columns
    for k, want in ((0, a0), (1, a1)):          # again, one tree at a time: the same bytes
        assert begin(k) == 0
        rc, got = end(k)
        assert rc == 0 and got == want
    e.close()
