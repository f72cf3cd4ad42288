"""Full-size checks of the BASELINE configurations.

Against the oracle, level by level (nids bit-exact, w exact, wY / wYY within one float ulp per bin, identical descriptors,
splits and trees) -- the protocol of GBMTest's histogram monitor (h2o-algos/src/test/java/hex/tree/gbm/GBMTest.java:4722-4765):
  * configs[1] (C2): one depth-6 GBM tree on the 11M x 28 Higgs-shape frame (+ a second tree without captures);
  * configs[2] (C3): two DRF trees, depth 12, mtries 5, sample_rate 0.63 on the same frame;
  * configs[3] (C4): airlines-shape 10M x 12 with 300 / 300 / 5000 / 20000-level categoricals, gaussian, depth 5.
The oracle runs on all host cores (a tree of C2 takes it about a second per million rows and core).

And through size-independent properties. Per level of a real tree build:
  * every active (leaf, col) histogram of a leaf holds the same total weight, sum(wY) and sum(wYY) (the rows of a leaf are
    the same whatever the column: a checksum of checksums across 28 columns);
  * the leaf totals add up to the number of in-bag rows and to the global sum of the work column;
  * the shared-memory and the global-atomics kernels agree bit for bit (integer accumulation);
  * a second run reproduces trees and predictions bit for bit.
"""
import numpy as np
import pytest

import h2ot_lib

pytestmark = pytest.mark.gpu


def test_higgs_shape_11m_x_28_level_invariants():
    import importlib
    bench = importlib.import_module("bench")
    P = h2ot_lib.pkg()
    N, C = 11_000_000, bench.NCOLS
    cols, y, _ = bench.make_frame(N, 0, N, pinned=False)
    sig = []
    for run in range(2):
        gm = P.GBM(distribution="bernoulli", ntrees=2, **bench.PARAMS)
        gm.load(cols, y)
        if run == 0:
            gm.capture_levels(True)
        gm.train()
        sig.append((gm.predictions().tobytes(), [bytes(gm.tree(t)[0]) for t in range(gm.ntrees)]))
        if run == 0:
            assert gm.ncaptures() == 12
            for i in range(gm.ncaptures()):
                info = gm.capture_info(i)
                tot_w = 0.0
                for n in range(info["nleaves"]):
                    sums = []
                    for c in range(C):
                        d, _, _, has, nv = gm.capture_desc(i, n, c)
                        if not (d.flags & P.DESC_ACTIVE) or not has:
                            continue
                        v = gm.capture_vals(i, n, c, nv)
                        sums.append((v[0::3].sum(), v[1::3].sum(), v[2::3].sum()))
                    assert len(sums) >= 2
                    s = np.array(sums)
                    assert np.all(s[:, 0] == s[0, 0]), (i, n)            # counts: exact
                    # wY / wYY: every bin was float-cast, so column totals agree to float precision of the bin sums
                    assert np.ptp(s[:, 1]) <= 1e-6 * max(1.0, np.abs(s[:, 1]).max() + s[0, 2]), (i, n, s[:, 1])
                    assert np.ptp(s[:, 2]) <= 1e-6 * s[0, 2], (i, n)
                    tot_w += s[0, 0]
                assert tot_w == N, (i, tot_w)
        gm.close()
    assert sig[0] == sig[1]


def test_full_size_root_level_kernel_variants_agree():
    import importlib
    bench = importlib.import_module("bench")
    P = h2ot_lib.pkg()
    N, C = 11_000_000, bench.NCOLS
    cols, y, _ = bench.make_frame(N, 0, N, pinned=False)
    e = P.Engine(nrows=N, ncols=C)
    for j in range(C):
        e.load_column(j, cols[j])
    e.load_response(y)
    e.finalize_frame()
    e.init_model([0.0])
    work = (y - 0.53).astype(np.float32).astype(np.float64)
    e.set_work(work)
    e.begin_tree(0, 0, 1.0, float(np.max(np.abs(work))))
    descs = (P.HistDesc * C)()
    for j in range(C):
        descs[j] = P.make_desc(float(np.min(cols[j])), float(np.nextafter(np.max(cols[j]), np.inf)), 1024, order=j)
    res = []
    for mode in (0, 1):
        e.set_hist_mode(mode)
        e.build_level(descs)
        res.append(e.finish_level())
    assert np.array_equal(res[0]["vals"], res[1]["vals"])
    assert np.array_equal(res[0]["min2"], res[1]["min2"]) and np.array_equal(res[0]["maxIn"], res[1]["maxIn"])
    v = res[0]["vals"]
    for j in range(C):
        blk = v[res[0]["offsets"][j]: res[0]["offsets"][j] + 1025 * 3]
        assert blk[0::3].sum() == N
        assert res[0]["min2"][j] == np.min(cols[j]) and res[0]["maxIn"][j] == np.max(cols[j])
    # exact integer accumulation: the total of wY over all bins of a column equals the fp64 sum of the work column
    # up to the float cast of each bin (<= 2^-24 relative per bin)
    tot = float(np.sum(work))
    for j in range(0, C, 9):
        blk = v[res[0]["offsets"][j]: res[0]["offsets"][j] + 1025 * 3]
        assert abs(blk[1::3].sum() - tot) <= 1e-6 * float(np.sum(np.abs(work)))
    e.close()


def test_full_size_drf_sparse_and_column_major_kernels_grow_the_same_trees():
    """BASELINE configs[2] shape (DRF, 11M x 28, mtries 5, sample 0.63; depth 9 here): the leaf-major kernel (default) and the
    column-major kernels (hist mode 2) must grow bit-identical trees and out-of-bag votes, and the default build must have
    used the leaf-major kernel on the levels below the root."""
    import importlib
    bench = importlib.import_module("bench")
    P = h2ot_lib.pkg()
    N = 11_000_000
    cols, y, _ = bench.make_frame(N, 0, N, pinned=False)
    sig, sparse = [], []
    for mode in (0, 2):
        gm = P.DRF(distribution="bernoulli", ntrees=2, max_depth=9, mtries=5, sample_rate=0.63, min_rows=1.0, nbins=20, seed=1234)
        gm.load(cols, y)
        eng = gm.engine()
        eng.set_hist_mode(mode)
        eng.stats_reset()
        gm.train()
        sparse.append(eng.stats().sparse_launches)
        nodes = [gm.tree(t)[0] for t in range(gm.ntrees)]
        sig.append((gm.predictions().tobytes(), gm.oobt().tobytes(), [len(n) for n in nodes],
                    [[(x.col, x.bin, x.nasplit, x.nid0, x.nid1, np.float32(x.splat).tobytes(), np.float32(x.pred).tobytes()) for x in n] for n in nodes]))
        gm.close()
    assert sparse[0] >= 2 * 7 and sparse[1] == 0, sparse
    assert sig[0] == sig[1]


def _oracle_threads():
    import os
    return max(1, os.cpu_count() or 1)


def test_c2_11m_x_28_one_tree_matches_oracle_level_by_level():
    """BASELINE configs[1] at full size against the oracle: every level of the first tree (nids of all 11M rows, every
    (leaf, column) histogram, min/max, descriptors, splits), then a second tree on the updated residuals (trees + predictions)."""
    import importlib
    bench = importlib.import_module("bench")
    import oracle_lib as O
    P = h2ot_lib.pkg()
    N = 11_000_000
    cols, y, _ = bench.make_frame(N, 0, N, pinned=False)
    kw = {k: v for k, v in bench.PARAMS.items() if k != "seed"}
    gm = P.GBM(distribution="bernoulli", ntrees=2, seed=1234, **kw)
    gm.load(cols, y)
    gm.capture_levels(True)
    gm.train(1)
    gm.capture_levels(False)
    gm.train(1)
    X = np.stack(cols, axis=1)
    fr = O.Frame(X, y)
    del X
    nt = _oracle_threads()
    for ntrees, capture in ((1, True), (2, False)):
        om = O.Model(fr, O.default_params(family=O.BERNOULLI, nclass=2, ntrees=ntrees, nthreads=nt, chunk_rows=max(4096, N // (4 * nt)),
                                          seed=1234, **kw), capture=capture)
        if capture:
            # the product captured the first tree only: compare its levels, its tree and nothing of the second tree
            class FirstTree:
                ntrees = 1
                nrows = gm.nrows
                init_f = gm.init_f
                ncols = gm.ncols
                ncaptures = gm.ncaptures
                capture_info = gm.capture_info
                capture_nids = gm.capture_nids
                capture_desc = gm.capture_desc
                capture_vals = gm.capture_vals
                tree = gm.tree

                @staticmethod
                def predictions(k=0):
                    return om.F(k)  # predictions are compared after the second tree
            st = h2ot_lib.compare_models(FirstTree, om)
            assert st["levels"] == 6 and st["hists"] == 63 * 28, st
            assert st["n_vals_diff"] <= 2e-2 * st["n_vals"], st
        else:
            h2ot_lib.compare_models(gm, om, check_levels=False)
        del om
    gm.close()


def test_c3_drf_11m_x_28_depth12_mtries5_matches_oracle():
    """BASELINE configs[2] at full size: DRF, depth 12, mtries 5, sample_rate 0.63 -- two trees, every level against the oracle."""
    import importlib
    bench = importlib.import_module("bench")
    import oracle_lib as O
    P = h2ot_lib.pkg()
    N = 11_000_000
    cols, y, _ = bench.make_frame(N, 0, N, pinned=False)
    kw = dict(ntrees=2, max_depth=12, nbins=20, min_rows=1.0, sample_rate=0.63, seed=1234)
    gm = P.DRF(distribution="bernoulli", mtries=5, **kw)
    gm.load(cols, y)
    gm.capture_levels(True)
    gm.train()
    X = np.stack(cols, axis=1)
    fr = O.Frame(X, y)
    del X
    nt = _oracle_threads()
    om = O.Model(fr, O.default_params(family=O.BERNOULLI, nclass=2, nthreads=nt, chunk_rows=max(4096, N // (4 * nt)), **kw),
                 capture=True, drf=True, mtries=5)
    st = h2ot_lib.compare_models(gm, om)
    assert st["levels"] >= 2 * 12, st
    assert np.array_equal(gm.oobt(), om.oobt())
    gm.close()


def test_c4_airlines_shape_10m_x_12_matches_oracle():
    """BASELINE configs[3] shape at 10M rows (one GPU's share of the 100M x 12 frame on 8 GPUs is 12.5M): 8 integer columns
    (two with NAs) + 4 Zipf categoricals with 300 / 300 / 5000 / 20000 levels, gaussian, depth 5, nbins 20, nbins_cats 1024."""
    import oracle_lib as O
    P = h2ot_lib.pkg()
    N = 10_000_000
    rng = np.random.default_rng(42)
    ranges = [(1987, 2008), (1, 12), (1, 31), (1, 7), (0, 2400), (0, 2400), (-60, 600), (11, 4983)]
    cards = [300, 300, 5000, 20000]
    cols, cats = [], {}
    y = np.zeros(N)
    for c in range(12):
        if c < 8:
            lo, hi = ranges[c]
            v = rng.integers(lo, hi + 1, N).astype(np.float64)
            if c in (4, 6):
                v[rng.random(N) < 0.02] = np.nan
            if c in (5, 6):
                y += np.nan_to_num(v) * (0.01 if c == 5 else 0.5)
        else:
            card = cards[c - 8]
            v = np.minimum(rng.zipf(1.3, N) - 1, card - 1).astype(np.float64)
            y += (v % 7) * 2.0
            cats[c] = card
        cols.append(v)
    y += rng.normal(size=N) * 10
    y = y.astype(np.float32).astype(np.float64)
    kw = dict(ntrees=2, max_depth=5, nbins=20, nbins_cats=1024, seed=1234)
    gm = P.GBM(distribution="gaussian", nclass=1, **kw)
    gm.load(cols, y, categorical=cats)
    gm.capture_levels(True)
    gm.train()
    X = np.stack(cols, axis=1)
    fr = O.Frame(X, y, cat_cards=cats)
    del X
    nt = _oracle_threads()
    om = O.Model(fr, O.default_params(family=O.GAUSSIAN, nclass=1, nthreads=nt, chunk_rows=max(4096, N // (4 * nt)), **kw), capture=True)
    st = h2ot_lib.compare_models(gm, om, ybound=2.0 * float(np.max(np.abs(y - y.mean()))))
    assert st["levels"] == 10, st
    gm.close()
