"""Test-side access to the product package (directory name is not an identifier) + comparison helpers."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pkg():
    build = importlib.import_module("h2o-3_b200.build")
    build.build()
    return importlib.import_module("h2o-3_b200")


def has_gpu():
    try:
        return pkg().device_count() > 0
    except Exception:
        return False


NaN = float("nan")


def make_frame(n=20000, c=6, seed=7, na_frac=0.02, with_cat=True, with_int=True):
    """fp32-exact normals, one small-int column, one categorical column, NAs everywhere; bernoulli response."""
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, c)).astype(np.float32).astype(np.float64)
    cats = {}
    if with_int and c > 3:
        X[:, 3] = rng.integers(-3, 40, n)
    if with_cat and c > 4:
        X[:, 4] = rng.integers(0, 12, n)
        cats[4] = 12
    logit = X[:, 0] - 0.5 * X[:, 1] + (0.3 * (X[:, 4] % 3 == 1) if (with_cat and c > 4) else 0)
    y = (rng.random(n) < 1 / (1 + np.exp(-logit))).astype(np.float64)
    if na_frac > 0:
        X[rng.random((n, c)) < na_frac] = NaN
    return X, y, cats


def compare_models(gm, om, K=1, check_levels=True, rel=1e-6, ybound=1.0):
    """gm: product GBM (trained, maybe with captures); om: oracle Model. Asserts parity:
    bit-exact nids / descriptors / tree structure, histogram sums within `rel` of the oracle's -- and per bin within one float
    ulp + the fixed-point quantum of the engine's accumulators (ybound: a bound on |work| of the build, 1 for classification)."""
    P = pkg()
    assert gm.ntrees == om.ntrees
    # sums of non-exact terms (offsets through exp, non-float responses) depend on the summation order, in the reference too
    assert gm.init_f == om.init_f or abs(gm.init_f - om.init_f) <= 1e-12 * max(1.0, abs(om.init_f))
    stats = dict(levels=0, hists=0, max_rel=0.0, n_vals=0, n_vals_diff=0)
    if check_levels:
        # The reference launches ScoreBuildHistogram2 even when a tree has no undecided leaf left (T/SharedTree.java:671-683):
        # that pass builds nothing and only walks rows of dead-end nodes back to their parents, which GammaPass / CollectPreds
        # undo anyway (they take "the parent's decision"). The engine's host driver skips it, so such captures are dropped here.
        ocaps = [om.capture(i) for i in range(om.ncaptures())]
        ocaps = [c for c in ocaps if c["nleaves"] > 0]
        assert gm.ncaptures() == len(ocaps)
        for i in range(gm.ncaptures()):
            gi = gm.capture_info(i)
            oc = ocaps[i]
            assert (gi["tree"], gi["k"], gi["leaf"], gi["nleaves"]) == (oc["tree"], oc["k"], oc["leaf"], oc["nleaves"]), (gi, i)
            assert np.array_equal(gm.capture_nids(i), oc["nids"]), "row->nid assignment differs at capture %d" % i
            ncols = gm.ncols
            for n in range(gi["nleaves"]):
                for c in range(ncols):
                    od = oc["descs"][n * ncols + c]
                    gd, gmin2, gmaxin, ghas, nv = gm.capture_desc(i, n, c)
                    active = bool(gd.flags & P.DESC_ACTIVE)
                    assert active == bool(od.present), (i, n, c)
                    if not active:
                        continue
                    assert (gd.nbin, gd.min, gd.maxEx, gd.step) == (od.nbin, od.min, od.maxEx, od.step), (i, n, c)
                    assert bool(ghas) == bool(od.has_vals), (i, n, c)
                    assert gmin2 == od.min2 and gmaxin == od.maxIn, (i, n, c, gmin2, od.min2, gmaxin, od.maxIn)
                    if not ghas:
                        continue
                    gv = gm.capture_vals(i, n, c, nv)
                    ov = oc["vals"][n * ncols + c]
                    assert gv.shape == ov.shape
                    # w is exact. wY / wYY are float-cast sums (reducePrecision): per bin they may differ from the oracle's by
                    # ONE float ulp (a cast landing on the other side of a rounding boundary) plus the fp64 summation noise of
                    # the oracle itself, which is relative to sum|y| <= sqrt(w * wYY) of the bin, not to a cancelled sum.
                    assert np.array_equal(gv[0::3], ov[0::3]), (i, n, c)
                    noise = 1e-13 * np.sqrt(np.abs(ov[0::3] * ov[2::3])) + 1e-300
                    # The engine adds round(v * 2^e) as int64 with 2^e ~ 2^61 / (rows * bound): a value is off by at most half
                    # that quantum (values within 2^-14 of the bound are exact for float inputs at 16M rows, more below), so a bin of
                    # w rows by at most w * rows * bound * 2^-61 -- far below a float ulp unless the bin's own values are tiny.
                    quantum = ov[0::3] * float(gm.nrows) * 2.0 ** -60
                    for j in (1, 2):
                        g_, o_ = gv[j::3], ov[j::3]
                        ulp = np.spacing(np.maximum(np.abs(g_), np.abs(o_)).astype(np.float32)).astype(np.float64)
                        bad = np.abs(g_ - o_) > ulp + (noise if j == 1 else noise * 0) + quantum * (ybound if j == 1 else ybound * ybound)
                        assert not bad.any(), (i, n, c, j, int(bad.sum()), g_[bad][:4], o_[bad][:4])
                        scale = max(1e-300, float(np.max(np.abs(o_))))
                        d = float(np.max(np.abs(g_ - o_))) / scale
                        stats["max_rel"] = max(stats["max_rel"], d)
                        assert d <= rel, (i, n, c, j, d)
                        stats["n_vals"] += g_.size
                        stats["n_vals_diff"] += int(np.sum(g_ != o_))
                    stats["hists"] += 1
            stats["levels"] += 1
    for t in range(gm.ntrees):
        for k in range(K):
            gt, gls = gm.tree(t, k)
            ot = om.tree(t, k)
            assert len(gt) == len(ot), "tree %d/%d: %d vs %d nodes" % (t, k, len(gt), len(ot))
            assert gls == om.leaves_start(t, k)
            for a, b in zip(gt, ot):
                assert (a.kind, a.pid, a.nid, a.nid0, a.nid1, a.has_split) == (b.kind, b.pid, b.nid, b.nid0, b.nid1, b.has_split), (t, k, a.nid)
                if a.has_split:
                    s = b.split
                    assert (a.col, a.bin, a.nasplit, a.equal) == (s.col, s.bin, s.nasplit, s.equal), (t, k, a.nid)
                    assert np.float32(a.splat).tobytes() == np.float32(b.splat).tobytes(), (t, k, a.nid)
                    for f in ("se", "se0", "se1", "n0", "n1", "p0", "p1", "tree_p0", "tree_p1"):
                        x, y = getattr(a, f), getattr(s, f)
                        # a child's squared error wYY - wY^2 / w cancels towards 0 in a nearly pure child: what is left is the float
                        # rounding of the bins' sums, i.e. noise at 1e-6 of the NODE's error, not of the child's
                        slack = 1e-6 * max(abs(a.se), abs(s.se)) if f in ("se0", "se1") else 0.0
                        assert x == y or abs(x - y) <= 1e-6 * max(abs(x), abs(y)) + slack, (t, k, a.nid, f, x, y)
                    if a.equal >= 2:
                        assert (a.bs_nbits, a.bs_bitoff, a.bs_nbytes) == (s.bs_nbits, s.bs_bitoff, s.bs_nbytes)
                        assert bytes(a.bs[:a.bs_nbytes]) == bytes(s.bs[:s.bs_nbytes])
                if a.kind == 2:
                    assert a.pred == b.pred or abs(a.pred - b.pred) <= 1e-6 * max(abs(a.pred), abs(b.pred)), (t, k, a.nid, a.pred, b.pred)
    for k in range(K):
        gF, oF = gm.predictions(k), om.F(k)
        assert np.max(np.abs(gF - oF)) <= 1e-5, float(np.max(np.abs(gF - oF)))
    print("compare_models: %d levels, %d histograms, %d of %d float-cast sums differ from the oracle's (each by <= 1 float ulp), "
          "largest difference %.3g of the column scale" % (stats["levels"], stats["hists"], stats["n_vals_diff"], stats["n_vals"], stats["max_rel"]))
    return stats


def logloss(y, F):
    p = 1 / (1 + np.exp(-F))
    p = np.clip(p, 1e-15, 1 - 1e-15)
    return float(-np.mean(y * np.log(p) + (1 - y) * np.log(1 - p)))


def auc(y, F):
    order = np.argsort(F, kind="stable")
    r = np.empty(len(F))
    r[order] = np.arange(1, len(F) + 1)
    # average ranks for ties
    Fs = F[order]
    i = 0
    while i < len(Fs):
        j = i
        while j + 1 < len(Fs) and Fs[j + 1] == Fs[i]:
            j += 1
        if j > i:
            r[order[i:j + 1]] = (i + 1 + j + 1) / 2
        i = j + 1
    npos = y.sum()
    nneg = len(y) - npos
    return float((r[y == 1].sum() - npos * (npos + 1) / 2) / (npos * nneg))
