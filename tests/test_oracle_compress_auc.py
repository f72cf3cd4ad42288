"""Oracle restatements of DTree.compress / SharedTreeMojoModel.scoreTree (oracle/h2o_compress_oracle.hpp) and of AUC2
(oracle/h2o_auc_oracle.hpp), pinned on the CPU by the reference's own invariants (no byte vectors exist in the reference:
CompressedTreeTest trains on smalldata/logreg/prostate.csv, which is not shipped):
  * DecidedNode.size() == bytes written (the assert at DTree.java:929) -- checked inside the oracle, -1 on violation;
  * scoreTree over the bytes == the walk over the node table for every row (CompressedTreeTest.testNodeIdAssignment's check),
    including bitset splits in both formats (compress2: <= 32 levels, compress3: offset + length), NA-vs-rest splits and stumps;
  * with at most 400 distinct predictions AUCBuilder never merges bins and AUC2's trapezoid sum is the exact AUC (AUC2.java:36, 262-277);
  * a single builder and per-chunk builders folded by reduce() agree when nothing has to be merged.
"""
import ctypes as C
import os

import numpy as np
import pytest

import h2ot_lib
import oracle_lib as O

NaN = float("nan")


def _model(X, y, cats, **kw):
    p = O.default_params(family=O.BERNOULLI, nclass=2, **kw)
    return O.Model(O.Frame(X, y, cat_cards=cats), p)


def test_compressed_bytes_score_like_the_node_table():
    X, y, cats = h2ot_lib.make_frame(n=6000, c=6, seed=3)
    X[:, 5] = np.where(np.arange(len(X)) % 3 == 0, NaN, X[:, 5])  # a column with many NAs: NA-vs-rest candidates
    cats = dict(cats)
    rng = np.random.default_rng(0)
    Xw = np.concatenate([X, rng.integers(0, 90, (len(X), 1)).astype(np.float64)], axis=1)  # 90 levels: bitsets beyond 32 bits
    cats[6] = 90
    yw = ((Xw[:, 6] % 7 < 3) ^ (y > 0)).astype(np.float64)
    seen = set()
    for XX, yy, cc, kw in ((X, y, cats, dict(ntrees=4, max_depth=5, nbins=20, nbins_cats=8, seed=7)),
                           (Xw, yw, cats, dict(ntrees=4, max_depth=6, nbins=20, nbins_cats=1024, seed=11)),
                           (X, y, cats, dict(ntrees=1, max_depth=3, nbins=20, min_rows=1e9, seed=1))):  # stump
        om = _model(XX, yy, {k: v for k, v in cc.items() if k < XX.shape[1]}, **kw)
        for t in range(om.ntrees):
            bits, aux = om.compressed_tree(t)
            nodes = om.tree(t)
            n_dec = sum(1 for n in nodes if n.kind == 1 and n.has_split and _reachable(nodes, n.nid))
            assert len(aux) == 40 * n_dec
            got = O.score_compressed(bits, XX)
            want = np.array([om.score_tree(t, 0, XX[r]) for r in range(len(XX))], dtype=np.float32)
            assert np.array_equal(got, want)
            for n in nodes:
                if n.kind == 1 and n.has_split:
                    seen.add((n.split.equal, n.split.nasplit == O.NAvsREST))
            if len(nodes) == 1:
                seen.add("stump")
                assert bits[:3] == b"\x00\xff\xff" and len(bits) == 7
    assert (0, False) in seen and (2, False) in seen and (3, False) in seen and "stump" in seen, seen


def _reachable(nodes, nid):
    while nodes[nid].pid >= 0:
        p = nodes[nodes[nid].pid]
        if nid not in (p.nid0, p.nid1):
            return False
        nid = p.nid
    return nid == 0


def test_auc2_is_exact_without_merges_and_chunking_agrees():
    rng = np.random.default_rng(5)
    n = 20000
    act = (rng.random(n) < 0.3).astype(np.float64)
    pred = np.round(np.clip(0.3 + 0.25 * (act - 0.3) + 0.2 * rng.standard_normal(n), 0, 1), 2)  # <= 101 distinct values
    a1, ths, tps, fps = O.auc2(pred, act)
    assert len(ths) == len(np.unique(pred)) and np.array_equal(ths, np.unique(pred))
    assert tps.sum() == act.sum() and fps.sum() == n - act.sum()
    assert abs(a1 - h2ot_lib.auc(act, pred)) < 1e-12
    a2, ths2, tps2, fps2 = O.auc2(pred, act, chunk_rows=1500)
    # reduce() merges the chunks' equal thresholds through combine_centers, which may move a centre by an ulp; counts are exact
    assert np.allclose(ths, ths2, rtol=1e-14, atol=0) and np.array_equal(tps, tps2) and np.array_equal(fps, fps2) and a1 == a2
    w = rng.integers(0, 3, n).astype(np.float64)
    a3, _, tps3, fps3 = O.auc2(pred, act, w=w, chunk_rows=777)
    assert tps3.sum() == (w * act).sum() and fps3.sum() == (w * (1 - act)).sum()
    # continuous predictions: 400 bins, within the streaming histogram's own error of the exact value, and always in (0, 1)
    predc = np.clip(0.3 + 0.25 * (act - 0.3) + 0.2 * rng.standard_normal(n), 0, 1)
    a4, ths4, _, _ = O.auc2(predc, act, chunk_rows=4096)
    assert len(ths4) == 400 and np.all(np.diff(ths4) > 0)
    assert abs(a4 - h2ot_lib.auc(act, predc)) < 1e-3
    # edge cases: one class only (AUC2.java:263-264)
    assert O.auc2(pred, np.ones(n))[0] == 1.0 and O.auc2(pred, np.zeros(n))[0] == 0.0


def test_auc_builder_known_answer_of_the_reference_test():
    """h2o-core/src/test/java/hex/AUCBuilderTest.java:16-26 (testPerRow): a 10-bin builder fed i / 100 for i < 100, all actuals 1,
    ends with these thresholds (the test's own tolerance: 1e-5)."""
    ths, tps, fps, _ = O.auc_builder(10, np.arange(100) / 100.0, np.ones(100, dtype=np.int32))
    assert len(ths) == 10
    assert np.allclose(ths, [0.05, 0.16, 0.25, 0.335, 0.445, 0.555, 0.655, 0.76, 0.875, 0.965], rtol=0, atol=1e-5)
    assert tps.sum() == 100 and fps.sum() == 0


def test_auc2_equals_the_exact_auc_like_the_reference_test():
    """AUC2Test.perfectAUC (h2o-core/src/test/java/hex/AUC2Test.java:66-74) checks AUC2.perfectAUC against a sort-based exact AUC on
    these inputs; with fewer than 400 distinct predictions the streaming builder merges nothing, so the oracle's AUC2 must give the
    same exact values: (0,1 | 0,1) = 1, (0,1,0 | 1,0,1) = 0, (.1,.2,.2,.9 | 0,1,1,1) = 1, (.1,.2,.2,.9 | 0,1,0,1) = 0.875 -- ties
    count half (the diagonal across a tied group)."""
    for p, a, want in (([0, 1], [0, 1], 1.0), ([0, 1, 0], [1, 0, 1], 0.0), ([0.1, 0.2, 0.2, 0.9], [0, 1, 1, 1], 1.0),
                       ([0.1, 0.2, 0.2, 0.9], [0, 1, 0, 1], 0.875)):
        got = O.auc2(np.array(p, dtype=np.float64), np.array(a, dtype=np.float64))[0]
        assert abs(got - want) < 1e-15, (p, a, got, want)


REF_AUC_ROWS = "/root/reference/h2o-core/src/test/resources/hex/aucbuilder.csv.gz"


@pytest.mark.skipif(not os.path.exists(REF_AUC_ROWS), reason="the reference tree (and its test resources) exists only in the build container")
def test_auc_builder_fast_path_equals_merge_path_on_the_reference_rows():
    """AUCBuilderTest.testPerRow_compat (h2o-core/src/test/java/hex/AUCBuilderTest.java:28-60): on the reference's own 240 618
    (prediction, actual) rows the fast path of perRow must leave the builder exactly as insert + mergeOneBin does, after every
    row. Read from the reference tree where it is (nothing is copied into this repository)."""
    import gzip
    rows = np.loadtxt(gzip.open(REF_AUC_ROWS, "rt"), delimiter=",")
    pred = np.ascontiguousarray(rows[:, 0])
    act = np.ascontiguousarray(rows[:, 1].astype(np.int32))
    L = O.lib()
    L.orc_auc_builder_compat.restype = C.c_longlong
    L.orc_auc_builder_compat.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_longlong]
    bad = L.orc_auc_builder_compat(400, O.ptr(pred), O.ptr(act), len(pred))
    assert bad == -1, "fast path and merge path differ after row %d" % bad


REF_MOJOS = "/root/reference/h2o-genmodel/src/test/resources/hex/genmodel"


def _reference_trees():
    """(label, mojo version, bits, aux) of every compressed tree the reference ships as a test fixture: the isolation-forest trees
    lying open in the resources, and the GBM / DRF models inside the MOJO zips (stacked-ensemble zips hold whole sub-models)."""
    import glob
    import zipfile
    out = []
    iso = os.path.join(REF_MOJOS, "algos/isofor")
    for f in sorted(glob.glob(os.path.join(iso, "trees", "t??_???.bin"))):
        out.append(("isofor/" + os.path.basename(f), 1.30, open(f, "rb").read(), open(f[:-4] + "_aux.bin", "rb").read()))
    for z in ("mojo.zip", "algos/ensemble/binomial_titanic.zip", "algos/ensemble/multinomial.zip", "algos/ensemble/regression.zip",
              "algos/ensemble/binomial.zip"):
        zf = zipfile.ZipFile(os.path.join(REF_MOJOS, z))
        names = set(zf.namelist())
        versions = {}
        for n in names:
            if n.endswith("model.ini"):
                for line in zf.read(n).decode().splitlines():
                    if line.startswith("mojo_version"):
                        versions[n[:-len("model.ini")]] = float(line.split("=")[1])
        for n in sorted(names):
            if n.endswith(".bin") and not n.endswith("_aux.bin") and "/trees/t" in "/" + n and n[:-4] + "_aux.bin" in names:
                prefix = n[:n.index("trees/")]
                out.append((z + ":" + n, versions.get(prefix, 0.0), zf.read(n), zf.read(n[:-4] + "_aux.bin")))
    return out


@pytest.mark.skipif(not os.path.isdir(REF_MOJOS), reason="the reference tree (and its test resources) exists only in the build container")
def test_trees_serialised_by_the_reference_survive_decode_and_recompress():
    """DTree.compress against bytes the reference itself wrote: every GBM / DRF / isolation-forest tree among the reference's MOJO
    test fixtures is decoded into a node table (oracle decompressTree) and serialised again by the oracle's restatement of
    DTree.compress / DecidedNode.size / compress / IcedBitSet.compress2/3 -- the bytes, and the auxiliary records, must come out
    identical. (The second word of an auxiliary record is the node's parent id in old MOJOs and the size of its left subtree in
    new ones -- SharedTreeMojoModel.AuxInfo.reserved -- so it is compared where it holds the count.)"""
    L = O.lib()
    L.orc_recompress_tree.restype = C.c_int
    trees = _reference_trees()
    assert len(trees) > 500
    n_counts = 0
    kinds_total = np.zeros(4, dtype=np.int64)
    for label, version, bits, aux in trees:
        ob, oa = C.create_string_buffer(len(bits) + 64), C.create_string_buffer(len(aux) + 64)
        nb, na = C.c_longlong(len(ob)), C.c_longlong(len(oa))
        res = (C.c_int * (len(aux) // 40 + 1))()
        err = C.create_string_buffer(256)
        kinds = (C.c_int * 4)()
        rc = L.orc_recompress_tree(bits, C.c_longlong(len(bits)), aux, C.c_longlong(len(aux)), ob, C.byref(nb), oa, C.byref(na), res, kinds, err, 256)
        assert rc == 0, (label, err.value)
        assert ob.raw[:nb.value] == bits, label
        got, want = np.frombuffer(oa.raw[:na.value], dtype=np.int32).reshape(-1, 10), np.frombuffer(aux, dtype=np.int32).reshape(-1, 10)
        assert got.shape == want.shape, label
        keep = [0, 2, 3, 4, 5, 6, 7, 8, 9]
        assert np.array_equal(got[:, keep], want[:, keep]), label
        if len(want) and np.array_equal(got[:, 1], want[:, 1]):
            n_counts += 1
        else:
            assert version < 1.30 or len(want) == 0, (label, version)  # new MOJOs carry the left-subtree node count, as the oracle writes it
        kinds_total += np.array(list(kinds))
    assert n_counts > 50
    # the fixtures exercise every split encoding: numeric thresholds, 32-bit and long category bitsets, NA-vs-rest
    print("decided nodes by kind (numeric, bitset32, bitset, NAvsREST):", kinds_total)
    assert kinds_total[0] > 1000 and kinds_total[1] + kinds_total[2] > 0


@pytest.mark.skipif(not os.path.isdir(REF_MOJOS), reason="the reference tree (and its test resources) exists only in the build container")
def test_score_tree_reader_known_answer_of_the_reference_mojo_test():
    """GbmMojoModelTest.testScore0 (h2o-genmodel/src/test/java/hex/genmodel/algos/gbm/GbmMojoModelTest.java:32-37): the 10-tree GBM
    MOJO under algos/gbm/calibrated scores this row as {0.5416688, 0.4583312} (1e-5). The oracle's restatement of
    SharedTreeMojoModel.scoreTree reads the reference's tree bytes; the rest is GbmMojoModel.unifyPreds for a 2-class multinomial
    model with one tree per iteration (f = sum of the trees + init_f, classes scored f and -f, GBM_rescale = softmax)."""
    d = os.path.join(REF_MOJOS, "algos/gbm/calibrated")
    init_f = None
    for line in open(os.path.join(d, "model.ini")):
        if line.startswith("init_f"):
            init_f = float(line.split("=")[1])
    row = np.array([[18.7, 1.51, 1.003, 132.53, 1.15, 0.2, 1.153, 8.3, 0.34, 0.0, 0.0]])
    f = init_f
    for t in range(10):
        bits = open(os.path.join(d, "trees", "t00_%03d.bin" % t), "rb").read()
        f += float(O.score_compressed(bits, row)[0])
    p = np.exp([f, -f])
    p /= p.sum()
    assert np.allclose(p, [0.5416688, 0.4583312], rtol=0, atol=1e-5), p


def test_metric_sums_restatement_on_known_values():
    """MetricBuilderBinomial / Multinomial / Regression.perRow + MathUtils.logloss restated (oracle/h2o_metrics_oracle.hpp), on
    values whose sums are known in closed form: a bernoulli model at f = 0 predicts 1/2 for every row (err = 1/2, logloss = ln 2),
    rows with a missing actual or zero weight do not count, a certain wrong prediction is capped at -ln(1e-15), a 3-class model
    with equal scores predicts 1/3 (logloss = ln 3, err^2 = 4/9), regression sums (y - f)^2."""
    n = 1000
    y = (np.arange(n) % 2).astype(np.float64)
    y[::50] = np.nan
    w = np.ones(n)
    w[1::50] = 0
    live = (~np.isnan(y)) & (w != 0)
    m = O.score_metrics(O.BERNOULLI, [np.zeros(n)], y, w)
    assert m["count"] == live.sum() and m["wcount"] == live.sum()
    assert abs(m["logloss"] - np.log(2)) < 1e-12 and abs(m["mse"] - 0.25) < 1e-12  # (row-by-row sums: a few ulps per add)
    m = O.score_metrics(O.BERNOULLI, [np.full(4, 800.0)], np.array([0.0, 0.0, 1.0, 1.0]))   # p1 == 1: two certain mistakes
    assert abs(m["logloss"] - 0.5 * -np.log(1e-15)) < 1e-12 and abs(m["mse"] - 0.5) < 1e-15
    m = O.score_metrics(O.MULTINOMIAL, [np.full(9, 0.3)] * 3, np.arange(9) % 3)
    assert abs(m["logloss"] - np.log(3)) < 1e-12 and abs(m["mse"] - 4.0 / 9.0) < 1e-12
    yg = np.array([1.0, 2.0, 4.0])
    m = O.score_metrics(O.GAUSSIAN, [np.array([0.5, 2.5, 1.0])], yg, np.array([1.0, 2.0, 1.0]))
    assert abs(m["mse"] - (0.25 + 2 * 0.25 + 9.0) / 4.0) < 1e-15 and m["logloss"] == 0
