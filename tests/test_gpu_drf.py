"""GPU parity of DRF builds (T/drf/DRF.java on the shared level loop) against the CPU oracle: per-level histograms with the
float-split correction, per-node mtries draws, row sampling with out-of-bag votes, leaf values from the split's child means."""
import numpy as np
import pytest

import h2ot_lib
import oracle_lib as O

pytestmark = pytest.mark.gpu
NaN = float("nan")


def _run_both(X, y, cats=None, w=None, capture=True, mtries=-1, **kw):
    P = h2ot_lib.pkg()
    dist = kw.pop("distribution", "bernoulli")
    nclass = kw.pop("nclass", 2 if dist == "bernoulli" else 1)
    fam = {"gaussian": O.GAUSSIAN, "bernoulli": O.BERNOULLI, "multinomial": O.MULTINOMIAL}[dist]
    op = O.default_params(family=fam, nclass=nclass, **kw)
    om = O.Model(O.Frame(X, y, w=w, cat_cards=cats), op, capture=capture, drf=True, mtries=mtries)
    gm = P.DRF(distribution=dist, nclass=nclass, mtries=mtries, **kw)
    gm.load([X[:, c] for c in range(X.shape[1])], y, weights=w, categorical=cats)
    gm.capture_levels(capture)
    gm.train()
    return gm, om


def test_drf_binomial_deep_trees():
    X, y, cats = h2ot_lib.make_frame(n=20000, c=8, seed=31)
    y[::333] = NaN  # NA responses: always out of bag, histogrammed at the root with a zero work value (reference quirk)
    gm, om = _run_both(X, y, cats, ntrees=4, max_depth=10, nbins=20, min_rows=1.0, sample_rate=0.632, seed=4242)
    st = h2ot_lib.compare_models(gm, om)
    assert st["levels"] >= 30
    assert np.array_equal(gm.oobt(), om.oobt())
    # out-of-bag scoring as DRF.score1 does it: votes / number of trees that held the row out
    o = gm.oobt()
    ok = (o > 0) & ~np.isnan(y)
    p0 = gm.predictions()[ok] / o[ok]
    assert np.mean((p0 < 0.5) == (y[ok] == 1)) > 0.6
    gm.close()


def test_drf_regression_weights_mtries():
    X, _, cats = h2ot_lib.make_frame(n=15000, c=6, seed=32)
    rng = np.random.default_rng(1)
    y = (np.nan_to_num(X[:, 0]) * 3 + np.sin(np.nan_to_num(X[:, 1]) * 2) + rng.normal(size=len(X)) * 0.2).astype(np.float32).astype(np.float64)
    w = rng.integers(0, 3, len(X)).astype(np.float64)
    gm, om = _run_both(X, y, cats, w=w, distribution="gaussian", mtries=3, ntrees=3, max_depth=8, nbins=20, min_rows=5.0,
                       sample_rate=0.7, seed=99)
    h2ot_lib.compare_models(gm, om)
    assert np.array_equal(gm.oobt(), om.oobt())
    gm.close()


def test_drf_multinomial():
    rng = np.random.default_rng(5)
    n = 12000
    X = rng.normal(size=(n, 7)).astype(np.float32).astype(np.float64)
    X[rng.random(X.shape) < 0.02] = NaN
    y = np.argmax(np.nan_to_num(X[:, :4]) + 0.5 * rng.normal(size=(n, 4)), axis=1).astype(np.float64)
    gm, om = _run_both(X, y, distribution="multinomial", nclass=4, ntrees=2, max_depth=7, nbins=20, min_rows=1.0, sample_rate=0.632,
                       seed=7)
    h2ot_lib.compare_models(gm, om, K=4)
    assert np.array_equal(gm.oobt(), om.oobt())
    gm.close()


def test_drf_very_deep_tree_many_leaf_ranges():
    """Depth 16 with min_rows 1: hundreds of leaves per level, so the leaf-major kernel works through several leaf ranges per
    level; the tree must still be the oracle's, node for node."""
    rng = np.random.default_rng(77)
    n = 60000
    X = rng.normal(size=(n, 6)).astype(np.float32).astype(np.float64)
    X[:, 4] = rng.integers(0, 50, n)
    X[rng.random(X.shape) < 0.01] = NaN
    y = (np.nan_to_num(X[:, 0]) * np.nan_to_num(X[:, 1]) + 0.3 * rng.normal(size=n) > 0).astype(np.float64)
    gm, om = _run_both(X, y, capture=False, mtries=2, ntrees=1, max_depth=16, nbins=20, min_rows=1.0, sample_rate=0.8, seed=5)
    h2ot_lib.compare_models(gm, om, check_levels=False)
    assert len(gm.tree(0)[0]) > 2000
    assert np.array_equal(gm.oobt(), om.oobt())
    gm.close()
