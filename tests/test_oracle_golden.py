"""Pins the CPU oracle against the reference's own fixture-free unit tests.

Each test cites the reference test it replays (paths relative to /root/reference,
TT/ = h2o-algos/src/test/java/hex/tree/). No GPU, no reference files needed at run time.
"""
import ctypes as C
import math

import pytest

import numpy as np

import oracle_lib as O

NaN = float("nan")


# ---- TT/DTreeTest.java:12-43 testFindBestSplitPoint_pubdev6495
def _se_non_na(ws, cs, ys):
    ws, cs, ys = map(np.asarray, (ws, cs, ys))
    m = ~np.isnan(cs)
    sw, swy = 0.0, 0.0
    for w, y in zip(ws[m], ys[m]):
        sw += w
        swy += w * y
    mean = swy / sw
    se = 0.0
    for w, y in zip(ws[m], ys[m]):
        se += w * (y - mean) * (y - mean)
    return se


def test_find_best_split_point_pubdev6495():
    ws = [10, 10, 100, 1]
    cs = [NaN, NaN, 0.5, 1.5]
    ys = [0, 1, 0, 1]
    hs = O.Hist(2, 2, 0, 0, 2, False, True, 0.01, 123).init()
    hs.update(ws, cs, ys)
    s1 = hs.find_split(0, 20)
    assert s1 is not None
    s2 = hs.find_split(0, 21)
    assert s2 is None
    hsN = O.Hist(2, 2, 0, 0, 2, False, True, -9, 123).init()
    hsN.update(ws, cs, ys)
    s3 = hsN.find_split(0, 21)
    assert s3 is not None
    assert s3.se == _se_non_na(ws, cs, ys)
    assert s3.se < s1.se


# ---- TT/DTreeSplitTest.java:18-37 testSplatNumeric
def test_splat_numeric():
    mn, maxIn = 20.0000000708, 20.0000009324
    maxEx = O.lib().orc_find_maxEx(maxIn, 0)
    for check, expect_nan in ((False, False), (True, True)):
        h = O.Hist(4, 2, 0, mn, maxEx, False, False, 0, 42, check).init()
        h.update([1.0, 1.0], [mn, maxIn], [1.0, 1.0])
        splat = h.splat_numeric(1, O.NALeft)
        if expect_nan:
            assert math.isnan(splat)
        else:
            assert splat == 20.0
            assert mn > splat and maxIn > splat


# ---- TT/DHistogramTest.java:173-206 testUpdateHistoWithIntOpt (java.util.Random(42) inputs)
def test_update_histo_with_int_opt():
    N = 10000
    ys = np.empty(N)
    dataInt = np.empty(N, dtype=np.int32)
    O.lib().orc_java_random_intopt_inputs(42, N, O.ptr(ys), O.ptr(dataInt))
    # java.util.Random(42): first nextGaussian / nextInt(900) known answers
    assert abs(ys[0] - 1.1419053154730547) < 1e-15
    assert dataInt.min() >= 13 and dataInt.max() < 913
    weights = np.arange(N, dtype=np.float64)
    hopt = O.Hist(1000, 1024, 1, 0, 1000, True, False, -0.001, 42).init()
    hopt.update_int(weights, dataInt, ys)
    h = O.Hist(1000, 1024, 1, 0, 1000, False, False, -0.001, 42).init()
    h.update(weights, dataInt.astype(np.float64), ys)
    assert h.min2 == hopt.min2
    assert h.maxIn == hopt.maxIn
    assert np.array_equal(h.vals(), hopt.vals())


# ---- TT/DHistogramTest.java:307-319 testNonEmptyBins
def test_non_empty_bins():
    h = O.Hist(1000, 1024, 1, 0, 1000, False, False, -0.001, 42)
    assert h.non_empty_bins() == 0
    h.init()
    h.update(None, [10, NaN, 999], [0.1, 0.2, 0.3])
    assert h.non_empty_bins() == 2


# ---- TT/HistogramTest.java:158-172 testUniformAdaptiveRange
def test_uniform_adaptive_range():
    nbins, mn, maxEx = 13, 1.0, 6.900000000000001
    h = O.Hist(nbins, nbins, 0, mn, maxEx, False, False, 0, 1234).init()
    assert h.binAt(0) == mn
    assert h.binAt(nbins - 1) < maxEx
    assert h.bin(mn) == 0
    assert h.bin(maxEx - 1e-15) == nbins - 1


# ---- TT/HistogramTest.java:111-157 testSplits (UniformAdaptive arm): bin <-> binAt round trip <= 1e-6
def test_bin_binat_round_trip():
    nbins, mn, maxEx = 13, 1.0, 6.900000000000001
    h = O.Hist(nbins, nbins, 0, mn, maxEx, False, False, 0, 7).init()
    N = 10_000_000
    i = np.arange(N, dtype=np.float64)
    col = mn + i / N * (maxEx - mn)
    bins = h.bin_many(col)
    assert np.all(np.diff(bins) >= 0)
    first = np.searchsorted(bins, np.arange(nbins), side="left")
    l1 = col[first]
    l2 = np.array([h.binAt(b) for b in range(nbins)])
    assert np.allclose(l1, l2, rtol=0, atol=1e-6)


# ---- TT/HistogramTest.java:22-67: float-cast sums stay within 1e-6 of the fp64 sums
def test_reduce_precision_accuracy():
    rng = np.random.default_rng(0xDECAF)
    cs = rng.uniform(0, 100, 100_000)
    ys = rng.normal(size=100_000)
    h = O.Hist(100, 100, 0, 0, 100.0000001, False, False, 0, 1).init()
    h.update(None, cs, ys)
    d = h.vals().copy()
    h.reduce_precision()
    f = h.vals()
    assert np.array_equal(d[0::3], f[0::3])  # w untouched (T/DHistogram.java:734-740)
    nz = d != 0
    assert np.max(np.abs(d[nz] - f[nz]) / np.abs(d[nz])) < 1e-6
    assert np.array_equal(f[1::3], f[1::3].astype(np.float32).astype(np.float64))


# ---- TT/DHistogramTest.java:57-75 spirit: -0.0 bins like 0.0 on the UniformAdaptive path
def test_negative_zero_bin():
    h = O.Hist(20, 1024, 1, -1, 2, False, False, -0.001, 42).init()
    assert h.bin(-0.0) == h.bin(0.0) == 1


# ---- DHistogram ctor int/cat shrink rule, T/DHistogram.java:237-247
def test_ctor_shrink_rule():
    h = O.Hist(20, 1024, 1, 0, 2, False, False, 0, 1)  # boolean-like int column
    assert h.nbin == 2 and h.step == 1.0
    h = O.Hist(20, 1024, 2, 0, 300, False, False, 0, 1)  # categorical, 300 levels <= nbins_cats
    assert h.nbin == 300 and h.step == 1.0
    h = O.Hist(20, 1024, 2, 0, 5000, False, False, 0, 1)  # categorical above nbins_cats -> binned ordinal
    assert h.nbin == 1024 and h.step == 1024 / 5000.0
    h = O.Hist(1024, 1024, 0, -3.0, 5.0, False, False, 0, 1)
    assert h.nbin == 1024 and h.step == 1024 / 8.0


# ---- find_maxEx, T/DHistogram.java:454-459
def test_find_max_ex():
    L = O.lib()
    assert L.orc_find_maxEx(1.0, 1) == 2.0
    assert L.orc_find_maxEx(1.0, 0) == np.nextafter(1.0, 2.0)
    assert L.orc_find_maxEx(1e17, 1) == 1e17 + 16


# ---- PCG-XSH-RR known answers (pcg32 reference stream: seed 42, seq 54 is the canonical demo;
# here we pin H2O's seq=1 variant against an independent pure-Python restatement).
def _pcg_py(seed, n):
    M = (1 << 64) - 1
    inc = 3
    state = 0

    def nxt():
        nonlocal state
        old = state
        state = (old * 6364136223846793005 + inc) & M
        xs = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
        rot = old >> 59
        return ((xs >> rot) | (xs << ((-rot) & 31))) & 0xFFFFFFFF

    nxt()
    state = (state + seed) & M
    nxt()
    out = []
    for _ in range(n):
        hi, lo = nxt(), nxt()
        out.append(((hi << 32) | lo) & ((1 << 24) - 1))
    return np.array(out, dtype=np.float32) / np.float32(1 << 24)


def test_pcg_next_float():
    got = np.empty(64, dtype=np.float32)
    O.lib().orc_pcg_floats(1234, 64, O.ptr(got))
    assert np.array_equal(got, _pcg_py(1234, 64))
    got2 = np.empty(8, dtype=np.float32)
    O.lib().orc_pcg_floats(-77, 8, O.ptr(got2))
    assert np.array_equal(got2, _pcg_py((-77) & ((1 << 64) - 1), 8))


def test_sample_rate_statistics():
    out = np.empty(200_000, dtype=np.uint8)
    O.lib().orc_sample(0xC0FFEE, 0.63, 0, len(out), O.ptr(out))
    assert abs((1 - out.mean()) - 0.63) < 0.005
    out2 = np.empty(1000, dtype=np.uint8)
    O.lib().orc_sample(0xC0FFEE, 0.63, 5000, 1000, O.ptr(out2))  # chunk-independent seeding, T/Sample.java:31
    assert np.array_equal(out2, out[5000:6000])


# ---- TT/gbm/GBMTest.java:4160-4232 testPubDev6356 (exact predictions 1.0 / 2.0 / 1.0)
def _pubdev6356(with_na):
    N, cnt = 1000, 100
    x = np.zeros(N)
    y = np.full(N, NaN)
    x[:cnt], y[:cnt] = 1, 1
    x[cnt:2 * cnt], y[cnt:2 * cnt] = 2, 2
    if not with_na:
        x, y = x[:2 * cnt], y[:2 * cnt]
    fr = O.Frame(x.reshape(-1, 1), y)
    p = O.default_params(family=O.GAUSSIAN, nclass=1, seed=42, max_depth=2, ntrees=1, learn_rate=1.0)
    m = O.Model(fr, p)
    return [m.init_f + float(m.score_tree(0, 0, [v])) for v in (1.0, 2.0, 0.0)]


def test_pubdev6356_no_na():
    assert _pubdev6356(False) == [1.0, 2.0, 1.0]


def test_pubdev6356_na_response():
    assert _pubdev6356(True) == [1.0, 2.0, 1.0]


# ---- TT/SharedTreeTest.java:80-138 checkNAPredictor (GBM arm, numeric predictor): root NAvsREST, 3 nodes
def test_na_predictor_num():
    x = np.array([NaN, 1, NaN, 1, NaN, 1]).reshape(-1, 1)
    y = np.array([0, 1, 0, 1, 0, 1], dtype=np.float64)
    fr = O.Frame(x, y)
    p = O.default_params(family=O.BERNOULLI, nclass=2, seed=42, ntrees=1, learn_rate=1.0, min_rows=1.0)
    m = O.Model(fr, p)
    t = m.tree(0, 0)
    assert len(t) == 3
    assert t[0].kind == 1 and t[0].has_split and t[0].split.nasplit == O.NAvsREST
    F = m.F(0)
    pred = (1 / (1 + np.exp(-F)) > 0.5).astype(int)
    assert np.array_equal(pred, y.astype(int))


# ---- checkNAPredictor, categorical arm (single level "V" + NAs)
def test_na_predictor_cat():
    x = np.array([NaN, 0, NaN, 0, NaN, 0]).reshape(-1, 1)
    y = np.array([0, 1, 0, 1, 0, 1], dtype=np.float64)
    fr = O.Frame(x, y, cat_cards={0: 1})
    p = O.default_params(family=O.BERNOULLI, nclass=2, seed=42, ntrees=1, learn_rate=1.0, min_rows=1.0)
    m = O.Model(fr, p)
    t = m.tree(0, 0)
    assert len(t) == 3 and t[0].split.nasplit == O.NAvsREST


# ---- TT/gbm/GBMTest.java:4801-4866 determinism protocol (scaled down): identical models from identical seeds,
# and thread count / chunking must not change the float-cast histograms' chosen splits.
def _small_frame(n=20000, c=6, seed=7):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, c)).astype(np.float32).astype(np.float64)
    X[rng.random((n, c)) < 0.02] = NaN
    X[:, 3] = rng.integers(0, 5, n)
    logit = X[:, 0] - 0.5 * np.nan_to_num(X[:, 1]) + 0.3 * (X[:, 3] == 2)
    y = (rng.random(n) < 1 / (1 + np.exp(-np.nan_to_num(logit)))).astype(np.float64)
    return X, y


def _tree_sig(m):
    sig = []
    for t in range(m.ntrees):
        for n in m.tree(t, 0):
            sig.append((n.kind, n.pid, n.nid0, n.nid1, n.split.col if n.has_split else -1,
                        np.float32(n.splat).tobytes(), np.float32(n.pred).tobytes()))
    return sig


def test_reproducibility_and_thread_invariance():
    X, y = _small_frame()
    fr = O.Frame(X, y)
    p1 = O.default_params(ntrees=5, max_depth=4, nthreads=1, chunk_rows=4096)
    m1 = O.Model(fr, p1)
    m2 = O.Model(fr, p1)
    assert _tree_sig(m1) == _tree_sig(m2)
    assert np.array_equal(m1.F(0), m2.F(0))
    p3 = O.default_params(ntrees=5, max_depth=4, nthreads=4, chunk_rows=1024)
    m3 = O.Model(fr, p3)
    assert np.max(np.abs(m1.F(0) - m3.F(0))) < 1e-5
    ll = lambda F: float(np.mean(np.log1p(np.exp(-(2 * y - 1) * F))))
    assert ll(m1.F(0)) < ll(np.full(len(y), m1.init_f))


def test_multinomial_runs_and_improves():
    rng = np.random.default_rng(3)
    n = 5000
    X = rng.normal(size=(n, 4)).astype(np.float32).astype(np.float64)
    y = (np.argmax(X[:, :3] + 0.5 * rng.normal(size=(n, 3)), axis=1)).astype(np.float64)
    fr = O.Frame(X, y)
    p = O.default_params(family=O.MULTINOMIAL, nclass=3, ntrees=5, max_depth=3)
    m = O.Model(fr, p)
    F = np.stack([m.F(k) for k in range(3)], axis=1)
    acc = float(np.mean(np.argmax(F, axis=1) == y))
    assert acc > 0.6


# ---- histogram types with split points (T/DHistogram.java:283-300, 343-430; TT/HistogramTest.java:111-157, 210-269)
def _limit(pts, mn, mx):
    L = O.lib()
    L.orc_limit_to_range.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_void_p]
    p = np.ascontiguousarray(pts, dtype=np.float64)
    out = np.empty(len(p), dtype=np.float64)
    n = L.orc_limit_to_range(O.ptr(p), len(p), mn, mx, O.ptr(out))
    return list(out[:n])


def _pad(pts, n_new):
    L = O.lib()
    L.orc_pad_uniformly.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    p = np.ascontiguousarray(pts, dtype=np.float64)
    out = np.empty(max(len(p), n_new), dtype=np.float64)
    n = L.orc_pad_uniformly(O.ptr(p), len(p), n_new, O.ptr(out))
    return list(out[:n])


def _uniq(pts, mn, mx):
    L = O.lib()
    L.orc_make_unique_and_limit.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_void_p]
    p = np.ascontiguousarray(pts, dtype=np.float64)
    out = np.empty(len(p) + 2, dtype=np.float64)
    n = L.orc_make_unique_and_limit(O.ptr(p), len(p), mn, mx, O.ptr(out))
    return list(out[:n])


def test_make_unique_and_limit_to_range_golden():
    """HistogramTest.testShrinking .. testShrinking4 (:210-229): exact arrays."""
    assert _uniq([0.2, 0.28, 0.31, 0.32, 0.32, 0.4, 0.7, 0.81, 0.84], 0.3, 0.8) == [0.3, 0.31, 0.32, 0.4, 0.7]
    long_ = [-0.3, 0.2, 0.28, 0.28, 0.3, 0.3, 0.31, 0.32, 0.32, 0.4, 0.7, 0.7, 0.8, 0.8, 0.81, 0.84]
    assert _uniq(long_, 0.3, 0.8) == [0.3, 0.31, 0.32, 0.4, 0.7]
    assert _uniq(long_, 0.3, 0.9) == [0.3, 0.31, 0.32, 0.4, 0.7, 0.8, 0.81, 0.84]
    assert _uniq([0.31, 0.32, 0.32, 0.4, 0.7, 0.7], 0.3, 0.9) == [0.3, 0.31, 0.32, 0.4, 0.7]


def test_limit_to_range_and_pad_uniformly_golden():
    """HistogramTest.testShrinking5-9, testPadding, testPadding2, testPadding3: exact arrays."""
    b = [0.3, 0.31, 0.32, 0.4, 0.7]
    assert _limit(b, 0.31, 0.9) == [0.31, 0.32, 0.4, 0.7]
    assert _limit(b, 0.305, 0.9) == [0.3, 0.31, 0.32, 0.4, 0.7]
    assert _limit(b, 0.305, 0.699) == [0.3, 0.31, 0.32, 0.4]
    assert _limit(b, 0.7, 0.9) == [0.7]
    assert _limit(b, 0.8, 0.9) == [0.7]
    assert _pad(b, 9) == [0.3, 0.305, 0.31, 0.315, 0.32, 0.36, 0.4, 0.55, 0.7]
    assert _pad(b, 10) == [0.3, 0.3025, 0.3075, 0.31, 0.315, 0.32, 0.36, 0.4, 0.55, 0.7]
    assert _pad(b, 8) == [0.3, 0.305, 0.31, 0.315, 0.32, 0.36, 0.4, 0.7]


@pytest.mark.parametrize("histo_type", [O.HT_UniformAdaptive, O.HT_Random, O.HT_QuantilesGlobal])
def test_splits_bin_and_binAt_agree_for_every_histogram_type(histo_type):
    """HistogramTest.testSplits (:111-157): nbins 13 on [1, 6.900000000000001); scanning the range, the first value that lands in
    bin b must be binAt(b) to 1e-6 -- for UniformAdaptive, Random (any seed) and QuantilesGlobal with the test's 13 points."""
    pts = [1, 1.5, 2, 2.5, 3, 4, 5, 6.1, 6.2, 6.3, 6.7, 6.8, 6.85]
    for seed in ([0] if histo_type != O.HT_Random else [1, -77, 123456789012345]):
        h = O.Hist(13, 13, 0, 1.0, 6.900000000000001, seed=seed)
        h.set_type(histo_type, pts if histo_type == O.HT_QuantilesGlobal else None)
        h.init()
        assert h.nbin == 13
        sp = h.split_points()
        if histo_type == O.HT_QuantilesGlobal:
            assert h.absolute and list(sp) == [float(p) for p in pts]
        if histo_type == O.HT_Random:
            assert not h.absolute and len(sp) == 13 and sp[0] == 0 and np.all(np.diff(sp) >= 0) and sp[-1] < 13
        N = 2_000_000
        x = 1.0 + np.arange(N, dtype=np.float64) / N * (6.900000000000001 - 1.0)
        b = h.bin_many(x)
        assert b.min() == 0 and b.max() <= 12 and np.all(np.diff(b) >= 0)
        l2 = np.array([h.binAt(i) for i in range(13)])
        first = np.searchsorted(b, np.arange(13), side="left")
        for i in range(13):
            if first[i] < N and b[first[i]] == i:  # Random can leave a bin empty (two equal split points)
                assert abs(x[first[i]] - l2[i]) <= 1e-5, (i, x[first[i]], l2[i])
        assert h.bin(1.0) == 0 and h.bin(float("nan")) == 13 and h.bin(float("inf")) == 12


@pytest.mark.parametrize("fam,alpha", [(O.LAPLACE, 0.5), (O.QUANTILE, 0.8), (O.QUANTILE, 0.1)])
def test_quantile_families_initial_prediction_is_the_type7_quantile(fam, alpha):
    """GBM.getInitialValueQuantile (T/gbm/GBM.java:355-381) through Quantile's INTERPOLATE method
    (h2o-core/src/main/java/hex/quantile/Quantile.java:330-350): the order statistics around row prob (n - 1), interpolated linearly
    -- numpy's default ("linear", Hyndman-Fan type 7) is the same definition, evaluated with another formula (a few ulps apart)."""
    import h2ot_lib
    X, y, cats = h2ot_lib.make_frame(n=5001, c=4, seed=3)
    y = (np.nan_to_num(X[:, 0]) + np.random.default_rng(1).standard_t(3, size=len(X))).astype(np.float32).astype(np.float64)
    y[::13] = np.nan
    p = O.default_params(family=fam, nclass=1, ntrees=1, max_depth=2, nbins=20, seed=5, quantile_alpha=alpha)
    m = O.Model(O.Frame(X, y, cat_cards=cats), p, capture=False)
    want = np.quantile(y[~np.isnan(y)], alpha)
    assert abs(m.init_f - want) <= 1e-12 * max(1.0, abs(want))
    # and the leaves are fitted towards that quantile of the residuals: after one tree about alpha of the rows lie below the prediction
    F = m.F()
    ok = ~np.isnan(y)
    assert abs(np.mean(y[ok] < F[ok]) - alpha) < 0.1
