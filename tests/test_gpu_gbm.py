"""GPU parity tests of whole GBM builds (host driver over the C ABI) against the CPU oracle, level by level:
the analogue of the reference's histogram-monitor test (TT/gbm/GBMTest.java:4722-4765)."""
import numpy as np
import pytest

import h2ot_lib
import oracle_lib as O

pytestmark = pytest.mark.gpu
NaN = float("nan")


def _run_both(X, y, cats=None, w=None, off=None, capture=True, **kw):
    P = h2ot_lib.pkg()
    dist = kw.pop("distribution", "bernoulli")
    nclass = kw.pop("nclass", 2 if dist == "bernoulli" else 1)
    fam = {"gaussian": O.GAUSSIAN, "bernoulli": O.BERNOULLI, "multinomial": O.MULTINOMIAL, "laplace": O.LAPLACE, "quantile": O.QUANTILE}[dist]
    ht = kw.pop("histogram_type", "AUTO")
    op = O.default_params(family=fam, nclass=nclass, histogram_type={"AUTO": 0, "Random": 1}[ht], **kw)
    kw["histogram_type"] = ht
    om = O.Model(O.Frame(X, y, w=w, off=off, cat_cards=cats), op, capture=capture)
    gm = P.GBM(distribution=dist, nclass=nclass, **kw)
    gm.load([X[:, c] for c in range(X.shape[1])], y, weights=w, offset=off, categorical=cats)
    gm.capture_levels(capture)
    gm.train()
    return gm, om


def test_bernoulli_levels_trees_and_metrics_match_oracle():
    X, y, cats = h2ot_lib.make_frame(n=30000, c=6, seed=11)
    gm, om = _run_both(X, y, cats, ntrees=6, max_depth=5, nbins=20, seed=1234)
    st = h2ot_lib.compare_models(gm, om)
    assert st["levels"] >= 6 * 4 and st["hists"] > 100
    # float-cast sums: the integer accumulation is exact, the oracle's sequential fp64 sum is not -> a handful of
    # 1-ulp(float) differences at most
    assert st["n_vals_diff"] <= 1e-3 * st["n_vals"], st
    F_g, F_o = gm.predictions(), om.F()
    assert abs(h2ot_lib.logloss(y, F_g) - h2ot_lib.logloss(y, F_o)) <= 1e-5
    assert abs(h2ot_lib.auc(y, F_g) - h2ot_lib.auc(y, F_o)) <= 1e-5
    assert np.allclose(gm.varimp(), om.varimp(), rtol=1e-5, atol=0)
    gm.close()


def test_config1_shape_100k_x_10():
    """BASELINE.json configs[0]: GBM bernoulli, 100k x 10 numeric, ntrees=10, max_depth=5, nbins=20."""
    rng = np.random.default_rng(42)
    n, c = 100000, 10
    X = rng.normal(size=(n, c)).astype(np.float32).astype(np.float64)
    beta = rng.normal(size=c)
    y = (rng.random(n) < 1 / (1 + np.exp(-(X @ beta + 0.5 * X[:, 0] * X[:, 1])))).astype(np.float64)
    gm, om = _run_both(X, y, ntrees=10, max_depth=5, nbins=20, seed=1234, capture=False)
    h2ot_lib.compare_models(gm, om, check_levels=False)
    F_g, F_o = gm.predictions(), om.F()
    assert abs(h2ot_lib.logloss(y, F_g) - h2ot_lib.logloss(y, F_o)) <= 1e-5
    assert abs(h2ot_lib.auc(y, F_g) - h2ot_lib.auc(y, F_o)) <= 1e-5
    gm.close()


def test_gaussian_weights_offset():
    X, _, cats = h2ot_lib.make_frame(n=20000, c=5, seed=5)
    rng = np.random.default_rng(6)
    y = np.nan_to_num(X[:, 0]) * 2 + np.nan_to_num(X[:, 2]) ** 2 + rng.normal(size=len(X)) * 0.1
    y = y.astype(np.float32).astype(np.float64)
    y[::501] = NaN  # NA responses are skipped everywhere
    w = rng.integers(0, 3, len(X)).astype(np.float64)
    off = (rng.random(len(X)) * 0.1).astype(np.float32).astype(np.float64)
    gm, om = _run_both(X, y, cats, w=w, off=off, distribution="gaussian", ntrees=4, max_depth=4, nbins=32, seed=99)
    h2ot_lib.compare_models(gm, om)
    gm.close()


def test_row_and_column_sampling():
    X, y, cats = h2ot_lib.make_frame(n=25000, c=8, seed=21)
    gm, om = _run_both(X, y, cats, ntrees=5, max_depth=5, nbins=20, seed=777, sample_rate=0.6, col_sample_rate=0.5,
                       col_sample_rate_per_tree=0.8)
    h2ot_lib.compare_models(gm, om)
    gm.close()


def test_multinomial_three_classes():
    rng = np.random.default_rng(3)
    n = 20000
    X = rng.normal(size=(n, 5)).astype(np.float32).astype(np.float64)
    X[rng.random(X.shape) < 0.01] = NaN
    y = np.argmax(np.nan_to_num(X[:, :3]) + 0.5 * rng.normal(size=(n, 3)), axis=1).astype(np.float64)
    gm, om = _run_both(X, y, distribution="multinomial", nclass=3, ntrees=4, max_depth=4, nbins=20, seed=31)
    h2ot_lib.compare_models(gm, om, K=3)
    gm.close()


def test_high_cardinality_categorical_and_min_rows():
    rng = np.random.default_rng(8)
    n = 40000
    card = 300
    eff = rng.normal(size=card)
    c0 = rng.integers(0, card, n).astype(np.float64)
    c1 = rng.normal(size=n).astype(np.float32).astype(np.float64)
    c2 = rng.integers(0, 2000, n).astype(np.float64)  # > nbins_cats levels: ordinal binned path
    X = np.stack([c0, c1, c2], axis=1)
    X[rng.random(X.shape) < 0.02] = NaN
    y = (eff[np.nan_to_num(c0).astype(int)] + 0.5 * c1 + rng.normal(size=n) * 0.2).astype(np.float32).astype(np.float64)
    gm, om = _run_both(X, y, {0: card, 2: 2000}, distribution="gaussian", ntrees=3, max_depth=4, nbins=20, nbins_cats=1024,
                       min_rows=50.0, seed=5)
    h2ot_lib.compare_models(gm, om)
    gm.close()


def test_pubdev6356_exact_predictions():
    """TT/gbm/GBMTest.java:4160-4232: 100 rows of (1,1), 100 of (2,2) and 800 rows with NA response; a depth-2
    learn_rate=1 gaussian tree must predict exactly 1.0 / 2.0."""
    P = h2ot_lib.pkg()
    N, cnt = 1000, 100
    x = np.zeros(N)
    y = np.full(N, NaN)
    x[:cnt], y[:cnt] = 1, 1
    x[cnt:2 * cnt], y[cnt:2 * cnt] = 2, 2
    gm = P.GBM(distribution="gaussian", nclass=1, seed=42, max_depth=2, ntrees=1, learn_rate=1.0)
    gm.load([x], y)
    gm.train()
    F = gm.predictions()
    assert np.all(F[:cnt] == 1.0) and np.all(F[cnt:2 * cnt] == 2.0)
    gm.close()


def test_na_predictor_root_split_is_na_vs_rest():
    """TT/SharedTreeTest.java:80-138 checkNAPredictor: root NAvsREST split, 3-node tree, zero training error."""
    P = h2ot_lib.pkg()
    x = np.array([NaN, 1, NaN, 1, NaN, 1])
    y = np.array([0, 1, 0, 1, 0, 1], dtype=np.float64)
    gm = P.GBM(distribution="bernoulli", seed=42, ntrees=1, learn_rate=1.0, min_rows=1.0)
    gm.load([x], y)
    gm.train()
    t, _ = gm.tree(0)
    assert len(t) == 3 and t[0].has_split and t[0].nasplit == P.NA_VS_REST
    F = gm.predictions()
    assert np.array_equal((1 / (1 + np.exp(-F)) > 0.5).astype(int), y.astype(int))
    gm.close()


def test_run_to_run_determinism():
    """TT/gbm/GBMTest.java:4801-4866 protocol (scaled down): identical models from identical seeds, bit for bit."""
    X, y, cats = h2ot_lib.make_frame(n=60000, c=6, seed=2)
    P = h2ot_lib.pkg()
    sigs = []
    for _ in range(3):
        gm = P.GBM(ntrees=3, max_depth=5, seed=1234, sample_rate=0.8)
        gm.load([X[:, c] for c in range(X.shape[1])], y, categorical=cats)
        gm.train()
        sig = [gm.predictions().tobytes()]
        for t in range(gm.ntrees):
            nodes, _ = gm.tree(t)
            sig.append(bytes(nodes))
        sigs.append(sig)
        gm.close()
    assert sigs[0] == sigs[1] == sigs[2]


def test_wide_frame_epsilon_shape():
    """BASELINE configs[4] (Epsilon-shape wide dense frame, nbins 64), scaled down in rows: 2000 columns force many column
    groups per level and a 2000 x 1025-bin root level."""
    rng = np.random.default_rng(12)
    n, c = 3000, 2000
    X = rng.normal(size=(n, c)).astype(np.float32)
    X /= np.linalg.norm(X, axis=1, keepdims=True)
    X = X.astype(np.float64)
    beta = np.zeros(c)
    beta[:20] = rng.normal(size=20) * 8
    y = (rng.random(n) < 1 / (1 + np.exp(-(X @ beta)))).astype(np.float64)
    gm, om = _run_both(X, y, ntrees=2, max_depth=5, nbins=64, seed=3, capture=False)
    h2ot_lib.compare_models(gm, om, check_levels=False)
    gm.close()


def test_covtype_shape_seven_classes():
    """BASELINE configs[4] (covtype-shape: 10 numeric + 44 binary columns, 7 imbalanced classes, nbins 64), scaled down in rows."""
    rng = np.random.default_rng(13)
    n = 30000
    num = rng.normal(size=(n, 10)).astype(np.float32).astype(np.float64) * [1, 2, 3, 1, 1, 5, 1, 1, 1, 4]
    binary = (rng.random((n, 44)) < rng.random(44) * 0.3).astype(np.float64)
    X = np.concatenate([num, binary], axis=1)
    score = np.stack([num[:, k % 10] * (1 + 0.2 * k) + binary[:, k] * 2 + rng.normal(size=n) for k in range(7)], axis=1) + \
        np.array([2.0, 2.5, 0.5, -1, 0, -0.5, -1.5])
    y = np.argmax(score, axis=1).astype(np.float64)
    gm, om = _run_both(X, y, distribution="multinomial", nclass=7, ntrees=2, max_depth=5, nbins=64, seed=17, capture=False)
    h2ot_lib.compare_models(gm, om, K=7, check_levels=False)
    gm.close()


def test_airlines_shape_mixed_with_high_cardinality_cats():
    """BASELINE configs[3] (airlines-shape: 8 integer columns, some with NAs, 4 categoricals two of which exceed nbins_cats),
    gaussian, scaled down in rows."""
    rng = np.random.default_rng(14)
    n = 60000
    cols = [rng.integers(1987, 2009, n), rng.integers(1, 13, n), rng.integers(1, 32, n), rng.integers(1, 8, n),
            rng.integers(0, 2400, n), rng.integers(0, 2400, n), rng.integers(30, 5000, n), rng.integers(0, 600, n)]
    cards = [300, 300, 5000, 20000]
    cats = {}
    for i, card in enumerate(cards):
        z = np.minimum(rng.zipf(1.3, n) - 1, card - 1)
        cols.append(z)
        cats[8 + i] = card
    X = np.stack(cols, axis=1).astype(np.float64)
    X[rng.random(n) < 0.02, 4] = NaN
    X[rng.random(n) < 0.02, 7] = NaN
    eff = rng.normal(size=300)
    y = (0.01 * (X[:, 6] - 800) / 100 + eff[X[:, 8].astype(int)] + 0.3 * (X[:, 3] == 5) + rng.normal(size=n)).astype(np.float32).astype(np.float64)
    gm, om = _run_both(X, y, cats, distribution="gaussian", ntrees=3, max_depth=5, nbins=20, nbins_cats=1024, seed=21, capture=True)
    h2ot_lib.compare_models(gm, om, ybound=2.0 * float(np.nanmax(np.abs(y))) + 1.0)
    gm.close()


def test_device_scoring_matches_metric_builders():
    """hex.tree.Score on the device: MSE / logloss from the resident Tree_k columns equal the
    MetricBuilder.perRow sums restated in numpy on the same predictions (H/ModelMetricsBinomial.java perRow, MathUtils.logloss)."""
    P = h2ot_lib.pkg()
    X, y, cats = h2ot_lib.make_frame(n=30000, c=6, seed=44)
    y[::211] = NaN
    w = np.random.default_rng(2).integers(0, 3, len(y)).astype(np.float64)
    ok = ~np.isnan(y) & (w != 0)
    gm = P.GBM(distribution="bernoulli", ntrees=5, max_depth=4, seed=3)
    gm.load([X[:, c] for c in range(X.shape[1])], y, weights=w, categorical=cats)
    gm.train()
    F = gm.predictions()
    p1 = 1 / (1 + np.minimum(1e19, np.exp(-F[ok])))
    pact = np.where(y[ok] == 1, p1, 1 - p1)
    want_ll = np.sum(w[ok] * np.minimum(-np.log(1e-15), -np.log(pact))) / np.sum(w[ok])
    want_mse = np.sum(w[ok] * (1 - pact) ** 2) / np.sum(w[ok])
    got = gm.score()
    assert abs(got["logloss"] - want_ll) <= 1e-12 * want_ll and abs(got["mse"] - want_mse) <= 1e-12 * want_mse
    # ... and the oracle's restatement of MetricBuilderBinomial.perRow (oracle/h2o_metrics_oracle.hpp) on the same columns
    om = O.score_metrics(O.BERNOULLI, [F], y, w)
    assert abs(got["logloss"] - om["logloss"]) <= 1e-12 * om["logloss"] and abs(got["mse"] - om["mse"]) <= 1e-12 * om["mse"]
    gm.close()
    # multinomial
    rng = np.random.default_rng(3)
    n = 20000
    X = rng.normal(size=(n, 4)).astype(np.float32).astype(np.float64)
    y = np.argmax(X[:, :3] + 0.5 * rng.normal(size=(n, 3)), axis=1).astype(np.float64)
    gm = P.GBM(distribution="multinomial", nclass=3, ntrees=3, max_depth=3, seed=3)
    gm.load([X[:, c] for c in range(4)], y)
    gm.train()
    Fk = np.stack([gm.predictions(k) for k in range(3)], axis=1)
    e = np.exp(Fk - Fk.max(axis=1, keepdims=True))
    p = e / e.sum(axis=1, keepdims=True)
    pact = p[np.arange(n), y.astype(int)]
    got = gm.score()
    assert abs(got["logloss"] - np.mean(-np.log(pact))) <= 1e-12 and abs(got["mse"] - np.mean((1 - pact) ** 2)) <= 1e-12
    om = O.score_metrics(O.MULTINOMIAL, [Fk[:, k] for k in range(3)], y)
    assert abs(got["logloss"] - om["logloss"]) <= 1e-12 and abs(got["mse"] - om["mse"]) <= 1e-12
    gm.close()
    # gaussian
    yg = (X[:, 0] * 2 + X[:, 1]).astype(np.float32).astype(np.float64)
    gm = P.GBM(distribution="gaussian", nclass=1, ntrees=3, max_depth=3, seed=3)
    gm.load([X[:, c] for c in range(4)], yg)
    gm.train()
    got = gm.score()
    assert abs(got["mse"] - np.mean((yg - gm.predictions()) ** 2)) <= 1e-12 * got["mse"] and got["logloss"] == 0.0
    om = O.score_metrics(O.GAUSSIAN, [gm.predictions()], yg)
    assert abs(got["mse"] - om["mse"]) <= 1e-12 * om["mse"]
    gm.close()


@pytest.mark.parametrize("dist", ["poisson", "gamma", "tweedie"])
def test_log_link_distributions(dist):
    """SURVEY 8f row 4 (first part): poisson / gamma / tweedie (H/DistributionFactory.java:274-379) — initFNum/initFDenom,
    negHalfGradient, gammaNum/gammaDenom with the log link applied to the leaf ratio (GammaPass.gamma, T/gbm/GBM.java:1315-1323)."""
    P = h2ot_lib.pkg()
    rng = np.random.default_rng({"poisson": 1, "gamma": 2, "tweedie": 3}[dist])
    n = 25000
    X, _, cats = h2ot_lib.make_frame(n=n, c=6, seed=77)
    mu = np.exp(0.4 * np.nan_to_num(X[:, 0]) - 0.3 * np.nan_to_num(X[:, 1]) + 0.5)
    if dist == "poisson":
        y = rng.poisson(mu).astype(np.float64)
    elif dist == "gamma":
        y = (rng.gamma(2.0, mu / 2.0) + 0.01).astype(np.float32).astype(np.float64)
    else:
        y = np.where(rng.random(n) < 0.4, 0.0, rng.gamma(2.0, mu / 2.0)).astype(np.float32).astype(np.float64)
    w = rng.integers(0, 3, n).astype(np.float64)
    off = (rng.random(n) * 0.2).astype(np.float32).astype(np.float64) if dist != "gamma" else None
    fam = {"poisson": O.POISSON, "gamma": O.GAMMA, "tweedie": O.TWEEDIE}[dist]
    kw = dict(ntrees=4, max_depth=4, nbins=20, seed=11)
    op = O.default_params(family=fam, nclass=1, tweedie_power=1.3, **kw)
    om = O.Model(O.Frame(X, y, w=w, off=off, cat_cards=cats), op, capture=True)
    gm = P.GBM(distribution=dist, nclass=1, tweedie_power=1.3, **kw)
    gm.load([X[:, c] for c in range(X.shape[1])], y, weights=w, offset=off, categorical=cats)
    gm.capture_levels(True)
    gm.train()
    h2ot_lib.compare_models(gm, om, ybound=2.0 * float(np.nanmax(np.abs(y))) + 1.0)
    gm.close()


def test_deep_levels_split_into_leaf_ranges(monkeypatch):
    """Levels whose leaves x bins exceed shared memory are cut into leaf-range groups (and a tiny budget forces the same
    path at shallow levels): results must not depend on the grouping."""
    X, y, cats = h2ot_lib.make_frame(n=40000, c=5, seed=61)
    kw = dict(ntrees=2, max_depth=8, nbins=256, min_rows=1.0, seed=5)
    gm, om = _run_both(X, y, cats, **kw)            # depth 8 x 257 bins: leaf ranges from level 6 on
    h2ot_lib.compare_models(gm, om)
    sig = [gm.predictions().tobytes()] + [bytes(gm.tree(t)[0]) for t in range(gm.ntrees)]
    gm.close()
    monkeypatch.setenv("H2OT_SMEM_BUDGET_KB", "24")  # one column per group, leaf ranges from level 3 on
    gm2, _ = _run_both(X, y, cats, **kw)
    h2ot_lib.compare_models(gm2, om)
    assert sig == [gm2.predictions().tobytes()] + [bytes(gm2.tree(t)[0]) for t in range(gm2.ntrees)]
    gm2.close()


@pytest.mark.parametrize("with_cat", [False, True])
def test_random_histogram_type_matches_oracle(with_cat):
    """histogram_type = Random (T/DHistogram.java:347-352, 742-749): every node draws its own sorted random split points from a
    PCG seeded by its range, bins by binary search and splits at binAt of those points. Levels, trees and predictions must
    match the oracle like the uniform build does (integer columns lose the integer fast path, :537-542)."""
    X, y, cats = h2ot_lib.make_frame(n=30000, c=6, seed=13, with_cat=with_cat)
    gm, om = _run_both(X, y, cats, ntrees=5, max_depth=5, nbins=20, seed=4321, histogram_type="Random")
    st = h2ot_lib.compare_models(gm, om)
    assert st["levels"] >= 5 * 4 and st["hists"] > 80
    # and it is not the uniform model
    gu, _ = _run_both(X, y, cats, ntrees=5, max_depth=5, nbins=20, seed=4321, capture=False)
    assert not np.array_equal(gm.predictions(), gu.predictions())
    gm.close()
    gu.close()


@pytest.mark.parametrize("dist", ["bernoulli", "multinomial"])
def test_predict_new_rows_matches_score_tree(dist):
    """Scoring NEW rows (SharedTreeMojoModel.scoreTree / scoreAllTrees + GbmMojoModel.score0's `+ init_f`): the device walks every
    finished tree root to leaf and sums the float leaf values in double, in build order -- bit for bit what the oracle's
    scoreTree gives row by row, NAs and never-seen categorical levels included."""
    X, y, cats = h2ot_lib.make_frame(n=30000, c=6, seed=91)
    K = 1
    if dist == "multinomial":
        y = np.argmax(np.nan_to_num(X[:, :3]), axis=1).astype(np.float64)
        K = 3
    gm, om = _run_both(X, y, cats, capture=False, distribution=dist, nclass=3 if K == 3 else 2, ntrees=6, max_depth=5, nbins=20, seed=77)
    rng = np.random.default_rng(92)
    m = 3000
    Xn = rng.normal(size=(m, 6)).astype(np.float32).astype(np.float64) * 1.5
    Xn[:, 3] = rng.integers(-10, 60, m)
    Xn[:, 4] = rng.integers(0, 20, m)          # levels 12..19 were never seen in training
    Xn[rng.random((m, 6)) < 0.05] = NaN
    got = gm.predict([Xn[:, c] for c in range(6)])
    assert got.shape == (K, m)
    want = np.zeros((K, m))
    for r in range(m):
        for k in range(K):
            s = 0.0
            for t in range(om.ntrees):
                s += float(om.score_tree(t, k, Xn[r]))
            want[k, r] = s + (om.init_f if K == 1 else 0.0)
    assert np.array_equal(got, want), int(np.sum(got != want))
    # on the training rows themselves the sum differs from Tree_k only by the per-tree float rounding of AddTreeContributions
    tr = gm.predict([X[:, c] for c in range(6)])
    assert np.max(np.abs(tr[0] - gm.predictions(0))) < 1e-5
    gm.close()


@pytest.mark.parametrize("case", ["numeric+small-cat", "wide-cat", "multinomial"])
def test_compressed_tree_bytes_match_dtree_compress(case):
    """DTree.compress (T/DTree.java:845-933, 972-982): the CompressedTree bytes the host driver emits for the trees grown on the
    device equal the oracle's restatement byte for byte (main and auxiliary stream), and scoring NEW rows by walking those bytes
    the way SharedTreeMojoModel.scoreTree does reproduces the device's own scoring of the node table (h2ot_score_tree)."""
    X, y, cats = h2ot_lib.make_frame(n=30000, c=6, seed=57)
    K, kw = 1, dict(ntrees=4, max_depth=5, nbins=20, seed=19)
    cats = dict(cats)
    if case == "wide-cat":
        rng = np.random.default_rng(3)
        X = np.concatenate([X, rng.integers(0, 90, (len(X), 1)).astype(np.float64)], axis=1)  # bitsets beyond 32 bits: compress3
        cats[6] = 90
        y = ((X[:, 6] % 7 < 3) ^ (y > 0)).astype(np.float64)
    elif case == "multinomial":
        y = np.argmax(np.nan_to_num(X[:, :3]), axis=1).astype(np.float64)
        K = 3
        kw.update(distribution="multinomial", nclass=3)
    else:
        kw.update(nbins_cats=8)  # 12 levels in 8 bins: group splits over binned levels
    gm, om = _run_both(X, y, cats, capture=False, **kw)
    rng = np.random.default_rng(58)
    Xn = X[rng.integers(0, len(X), 4000)].copy()
    Xn[rng.random(Xn.shape) < 0.05] = NaN
    Xn[:50, 4] = 15  # a level the training frame never had
    total = np.zeros((K, len(Xn)))
    kinds = set()
    for t in range(gm.ntrees):
        for k in range(K):
            bits, aux = gm.compressed_tree(t, k)
            obits, oaux = om.compressed_tree(t, k)
            assert bits == obits, (t, k, len(bits), len(obits))
            # auxiliary stream: node ids / subtree sizes exactly; the float-cast split statistics (n, p, se of the children) like
            # everywhere else in these tests, to 1e-6 (they come out of the fp64 split search, whose sums differ in the last bits)
            assert len(aux) == len(oaux) and len(aux) % 40 == 0
            ga = np.frombuffer(aux, dtype=np.uint8).reshape(-1, 40)
            oa = np.frombuffer(oaux, dtype=np.uint8).reshape(-1, 40)
            for lo, hi in ((0, 8), (32, 40)):
                assert np.array_equal(ga[:, lo:hi], oa[:, lo:hi]), (t, k)
            gf, of = ga[:, 8:32].copy().view(np.float32), oa[:, 8:32].copy().view(np.float32)
            # (se0 / se1 = wYY - wY^2 / w cancel to noise in a nearly pure child: absolute slack at the scale of the tree's errors)
            atol = np.zeros(6)
            atol[4:6] = 1e-6 * float(np.max(np.abs(of[:, 4:6]))) if len(of) else 0.0
            bad = np.argwhere(~(np.abs(gf - of) <= 3e-6 * np.abs(of) + atol))
            assert bad.size == 0, (t, k, bad[:4].tolist(), [(float(gf[i, j]), float(of[i, j])) for i, j in bad[:4]],
                                   ga[bad[0][0], :8].view(np.int32).tolist())
            total[k] += O.score_compressed(bits, Xn).astype(np.float64)
            for n in gm.tree(t, k)[0]:
                if n.has_split:
                    kinds.add(n.equal)
    got = gm.predict([Xn[:, c] for c in range(Xn.shape[1])])
    assert np.array_equal(got, total + (gm.init_f if K == 1 else 0.0))
    if case == "wide-cat":
        assert 3 in kinds
    if case == "numeric+small-cat":
        assert 0 in kinds
    gm.close()


def test_device_auc2_matches_the_streaming_builder():
    """hex.AUC2 on the device (h2ot_auc_builders + the host fold): one AUCBuilder per chunk of rows, folded in row order. Against
    the oracle's AUCBuilder run over the same chunking: identical bins and AUC when the device's p1 are fed to the oracle is not
    possible (p1 stays on the device), so the oracle gets p1 = 1 / (1 + exp(-F)) from the downloaded scores -- exp may differ from
    the device's by an ulp, which can move a bin centre in its last bits but not a count: counts must match exactly whenever no bin
    was merged, and the AUC within 1e-9 always; the exact (sorted) AUC bounds AUC2's own histogram error (< 1e-3)."""
    X, y, cats = h2ot_lib.make_frame(n=60000, c=6, seed=23)
    y[::97] = NaN  # rows without a response are not scored
    w = np.random.default_rng(4).integers(0, 3, len(y)).astype(np.float64)  # zero weights are skipped
    for weights, ntrees in ((None, 1), (w, 6)):
        gm, om = _run_both(X, y, cats, w=weights, capture=False, ntrees=ntrees, max_depth=3 if ntrees == 1 else 5, nbins=20, seed=5)
        r = gm.auc()
        F = gm.predictions()
        p1 = 1.0 / (1.0 + np.minimum(1e19, np.exp(-F)))
        a, ths, tps, fps = O.auc2(p1, y, w=weights, chunk_rows=r["chunk_rows"])
        assert abs(r["auc"] - a) <= 1e-9, (r["auc"], a)
        ok = ~np.isnan(y) & ((weights if weights is not None else np.ones(len(y))) > 0)
        ww = (weights if weights is not None else np.ones(len(y)))[ok]
        assert r["tps"].sum() == (ww * y[ok]).sum() and r["fps"].sum() == (ww * (1 - y[ok])).sum()
        assert len(r["ths"]) == len(ths)
        if ntrees == 1:  # a depth-3 tree: at most 8 distinct predictions, nothing is ever merged
            assert len(ths) <= 8 and np.array_equal(r["tps"], tps) and np.array_equal(r["fps"], fps)
            assert np.allclose(r["ths"], ths, rtol=1e-15, atol=0)
        if weights is None:
            assert abs(r["auc"] - h2ot_lib.auc(y[ok], F[ok])) < 1e-3
        gm.close()


@pytest.mark.parametrize("case", ["bernoulli+cats+sampling", "gaussian+weights+offset", "gaussian+weights", "multinomial", "deep+min_rows"])
def test_device_level_loop_grows_the_same_trees_as_the_host_loop(case, monkeypatch):
    """h2ot_tree_levels_begin / _end: all levels of a tree enqueued back to back, with the DecidedNode constructor and
    Split.nextLevelHistos (T/DTree.java:312-417, 703-726) on the device, against the level-by-level loop that does them on the host
    (H2OT_DEVICE_LOOP=0): identical trees (node numbering, splits, bitsets, predictions, varimp), and against the oracle."""
    P = h2ot_lib.pkg()
    X, y, cats = h2ot_lib.make_frame(n=40000, c=8, seed=71)
    K, w, off = 1, None, None
    kw = dict(ntrees=5, max_depth=6, nbins=20, seed=31)
    if case == "bernoulli+cats+sampling":
        kw.update(sample_rate=0.7, nbins_cats=8)
    elif case.startswith("gaussian"):
        # (without an offset gaussian trees are pipelined too: their fixed-point scale comes from the previous tree's residuals)
        rng = np.random.default_rng(2)
        y = (np.nan_to_num(X[:, 0]) * 2 + np.nan_to_num(X[:, 2]) ** 2 + rng.normal(size=len(X))).astype(np.float32).astype(np.float64)
        w = rng.integers(0, 3, len(X)).astype(np.float64)
        if case.endswith("offset"):
            off = (rng.random(len(X)) * 0.1).astype(np.float32).astype(np.float64)
        kw.update(distribution="gaussian", nclass=1, nbins=32, ntrees=8)
    elif case == "multinomial":
        # (noisy classes: with y a pure function of the features whole children become pure, their squared error is cancellation
        # noise around 0, and DTree's `se <= 1e-30` test -- T/DTree.java:330 -- falls either way: a near-tie of the engine's exact
        # sums against the oracle's sequential fp64 sums, not a property of the level loop under test)
        y = np.argmax(np.nan_to_num(X[:, :3]) + 0.7 * np.random.default_rng(8).normal(size=(len(X), 3)), axis=1).astype(np.float64)
        K = 3
        kw.update(distribution="multinomial", nclass=3, ntrees=3)
    else:
        kw.update(max_depth=8, min_rows=400.0, ntrees=3)  # many leaves stop early: empty slots, cut-off children
    sig = []
    # host loop; device loop with one synchronisation per tree; device loop with the trees in a pipeline (h2ot_tree_finish_enqueue:
    # leaf insertion, GammaPass, leaf values and AddTreeContributions without the host -- bernoulli / multinomial only, the others
    # fall back to the second mode by themselves)
    for loop, pipe in (("0", "0"), ("1", "0"), ("1", "1")):
        monkeypatch.setenv("H2OT_DEVICE_LOOP", loop)
        monkeypatch.setenv("H2OT_PIPELINE", pipe)
        gm, om = _run_both(X, y, cats, w=w, off=off, capture=False, **kw)
        nodes = [[bytes(gm.tree(t, k)[0]) for k in range(K)] for t in range(gm.ntrees)]
        sig.append((nodes, [gm.predictions(k).tobytes() for k in range(K)], gm.varimp().tobytes(), [gm.tree(t, 0)[1] for t in range(gm.ntrees)]))
        gm.close()
    for other in (1, 2):
        assert sig[0][3] == sig[other][3]
        assert sig[0][0] == sig[other][0], "trees differ between the host loop and the device loop (mode %d)" % other
        assert sig[0][1] == sig[other][1] and sig[0][2] == sig[other][2]
    monkeypatch.setenv("H2OT_DEVICE_LOOP", "1")
    monkeypatch.setenv("H2OT_PIPELINE", "1")
    gm, om = _run_both(X, y, cats, w=w, off=off, capture=False, **kw)
    h2ot_lib.compare_models(gm, om, K=K, check_levels=False)
    gm.close()


@pytest.mark.parametrize("capture", [True, False])
def test_sample_rate_per_class_and_pred_noise_bandwidth(capture):
    """SharedTreeParameters._sample_rate_per_class (T/Sample.java:14-30: the row's rate is picked by its response class, and rows
    are sampled even when sample_rate == 1) and GBMParameters._pred_noise_bandwidth (T/gbm/GBM.java:1450-1463: every leaf's value is
    scaled by 1 + nextGaussian * bandwidth, the generator re-seeded per tree, class and node) against the oracle -- level by
    level through the host loop (capture) and through the device loop."""
    X, y, cats = h2ot_lib.make_frame(n=30000, c=6, seed=41)
    gm, om = _run_both(X, y, cats, capture=capture, ntrees=5, max_depth=5, nbins=20, seed=77, sample_rate_per_class=[0.4, 0.9],
                       pred_noise_bandwidth=0.25)
    st = h2ot_lib.compare_models(gm, om, check_levels=capture)
    if capture:
        # out-of-bag rows follow the class rates: roughly 60 % of class 0 and 10 % of class 1 sit out of every tree
        nids = gm.capture_nids(1)
        oob = nids <= -2
        assert abs(oob[y == 0].mean() - 0.6) < 0.03 and abs(oob[y == 1].mean() - 0.1) < 0.03
    # the noise is real: the same build without it predicts differently
    g0, _ = _run_both(X, y, cats, capture=False, ntrees=5, max_depth=5, nbins=20, seed=77, sample_rate_per_class=[0.4, 0.9])
    assert not np.array_equal(g0.predictions(), gm.predictions())
    g0.close()
    gm.close()


@pytest.mark.parametrize("dist_kw", [dict(distribution="bernoulli"), dict(distribution="multinomial", nclass=3)])
def test_pipelined_build_one_iteration_at_a_time_with_accessors_in_between(dist_kw, monkeypatch):
    """h2ot_gbm_train may return with the last iteration still on the device (pipelined bernoulli / multinomial builds): every
    accessor has to wait for it and see a complete model. One iteration per call with reads in between must give the bytes of a
    build that synchronises after every tree (H2OT_PIPELINE=0)."""
    P = h2ot_lib.pkg()
    X, y, cats = h2ot_lib.make_frame(n=30000, c=6, seed=5)
    K = 1
    if dist_kw["distribution"] == "multinomial":
        y = np.argmax(np.nan_to_num(X[:, :3]) + 0.5 * np.random.default_rng(3).normal(size=(len(X), 3)), axis=1).astype(np.float64)
        K = 3
    kw = dict(ntrees=6, max_depth=5, nbins=20, seed=99, sample_rate=0.8, **dist_kw)
    out = []
    for pipe in ("0", "1"):
        monkeypatch.setenv("H2OT_PIPELINE", pipe)
        gm = P.GBM(**kw).load([X[:, c] for c in range(X.shape[1])], y, categorical=cats)
        seen = []
        for it in range(6):
            gm.train(1)
            if it % 2 == 0:
                assert gm.ntrees == it + 1                     # waits for the tree in flight
                seen.append(gm.predictions(0).copy())           # and sees its contribution
            if it == 3:
                assert len(gm.tree(3, K - 1)[0]) >= 1
        gm.flush()
        sc = gm.score()
        out.append(([bytes(gm.tree(t, k)[0]) for t in range(6) for k in range(K)], [gm.predictions(k).tobytes() for k in range(K)],
                    gm.varimp().tobytes(), [s.tobytes() for s in seen], sc))
        gm.close()
    assert out[0][0] == out[1][0], "pipelined trees differ from the tree-at-a-time build"
    assert out[0][1] == out[1][1] and out[0][2] == out[1][2] and out[0][3] == out[1][3]
    assert out[0][4] == out[1][4]


def test_timer_modes():
    """h2ot_set_timers: the model builder times the histogram launches only (mode 1); mode 2 times every stage; mode 0 none."""
    P = h2ot_lib.pkg()
    X, y, cats = h2ot_lib.make_frame(n=20000, c=5, seed=8)
    gm = P.GBM(ntrees=6, max_depth=4, nbins=20, seed=1).load([X[:, c] for c in range(X.shape[1])], y, categorical=cats)
    gm.train(2)
    e = gm.engine()
    s = e.stats()
    # (one histogram launch per level and storage type of the frame's columns)
    assert s.ms_hist > 0 and s.hist_launches >= 2 * 4 and s.ms_score == 0 and s.ms_split == 0 and s.ms_gamma == 0
    e.set_timers(2)
    gm.train(2)
    s2 = gm.engine().stats()
    assert s2.ms_score > 0 and s2.ms_split > 0 and s2.ms_gamma > 0 and s2.ms_hist > s.ms_hist
    e.set_timers(0)
    gm.train(2)
    s3 = gm.engine().stats()
    assert s3.ms_hist == s2.ms_hist and s3.hist_launches == 3 * s.hist_launches and s3.launches > s2.launches
    gm.close()


@pytest.mark.parametrize("dist,alpha", [("laplace", 0.5), ("quantile", 0.8), ("quantile", 0.25)])
def test_laplace_and_quantile_distributions(dist, alpha):
    """laplace / quantile GBM (H/DistributionFactory.java:406-434). The trees are grown on sign-like gradients,
    the initial prediction is the median / alpha-quantile of the response (GBM.getInitialValueQuantile, T/gbm/GBM.java:355-381) and
    every leaf is fitted with the quantile of its rows' residuals (fitBestConstantsQuantile :729-751 over
    Quantile.StratifiedQuantilesTask): h2ot_leaf_quantiles sorts (leaf, residual) on the device and interpolates like
    Quantile.linearInterpolate. Level by level against the oracle, with row sampling (out-of-bag rows take part in the leaf
    quantiles, as in the reference)."""
    X, y, cats = h2ot_lib.make_frame(n=30000, c=6, seed=17)
    rng = np.random.default_rng(4)
    y = (np.nan_to_num(X[:, 0]) * 1.5 - np.nan_to_num(X[:, 1]) + rng.standard_t(3, size=len(X))).astype(np.float32).astype(np.float64)
    y[::97] = NaN
    kw = dict(distribution=dist, nclass=1, ntrees=6, max_depth=4, nbins=20, seed=5, sample_rate=0.8)
    if dist == "quantile":
        kw["quantile_alpha"] = alpha
    gm, om = _run_both(X, y, cats, capture=True, **kw)
    h2ot_lib.compare_models(gm, om, ybound=1.0)
    ok = ~np.isnan(y)
    F = gm.predictions()
    # the fitted quantile: about alpha of the training rows lie below the prediction
    below = np.mean(y[ok] < F[ok])
    assert abs(below - alpha) < 0.08, below
    gm.close()
