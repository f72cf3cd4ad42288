"""tests/golden/gbm_small.json: (CPU) the oracle still reproduces the committed fixture bit for bit; (GPU) the engine,
driven through the C ABI, builds the trees the fixture holds."""
import hashlib
import json
import os
import sys

import numpy as np
import pytest

import h2ot_lib

sys.path.insert(0, os.path.join(h2ot_lib.ROOT, "tests", "golden"))
import make_golden as G  # noqa: E402

FIXTURE = json.load(open(os.path.join(h2ot_lib.ROOT, "tests", "golden", "gbm_small.json")))


@pytest.mark.parametrize("name", sorted(G.JOBS))
def test_oracle_reproduces_fixture(name):
    got = json.loads(json.dumps(G.run_oracle(name), sort_keys=True))
    assert got == FIXTURE[name]


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(G.JOBS))
def test_engine_matches_fixture(name):
    P = h2ot_lib.pkg()
    j = G.JOBS[name]
    want = FIXTURE[name]
    X, y, cats = G.job_data(name)
    K = 3 if j["family"] == "multinomial" else 1
    gm = P.GBM(distribution=j["family"], nclass=3 if K == 3 else (2 if j["family"] == "bernoulli" else 1), **j["params"])
    gm.load([X[:, c] for c in range(X.shape[1])], y, categorical=cats)
    gm.capture_levels(True)
    gm.train()
    assert gm.init_f == want["init_f"]
    assert gm.ntrees == len(want["trees"])
    for t in range(gm.ntrees):
        for k in range(K):
            nodes, _ = gm.tree(t, k)
            wt = want["trees"][t][k]
            assert len(nodes) == len(wt)
            for n, w in zip(nodes, wt):
                assert (n.kind, n.pid, n.nid, n.nid0, n.nid1) == (w["kind"], w["pid"], w["nid"], w["nid0"], w["nid1"])
                if n.has_split:
                    assert (n.col, n.bin, n.nasplit, n.equal, G.f32hex(n.splat), n.n0, n.n1) == \
                        (w["col"], w["bin"], w["nasplit"], w["equal"], w["splat"], w["n0"], w["n1"]), (t, k, n.nid)
                    if n.equal >= 2:
                        assert bytes(n.bs[:n.bs_nbytes]).hex() == w["bs"]
                else:
                    assert "col" not in w
                if n.kind == 2:
                    wp = np.frombuffer(bytes.fromhex(w["pred"]), dtype="<f4")[0]
                    assert n.pred == wp or abs(n.pred - wp) <= 1e-6 * max(abs(n.pred), abs(wp)), (t, k, n.nid, n.pred, wp)
    # row->node assignments and bin counts of every level are exact: same digest as the fixture
    h = hashlib.sha256()
    assert gm.ncaptures() == want["ncaptures"]
    for i in range(gm.ncaptures()):
        info = gm.capture_info(i)
        h.update(gm.capture_nids(i).tobytes())
        for n in range(info["nleaves"]):
            for c in range(gm.ncols):
                d, _, _, has, nv = gm.capture_desc(i, n, c)
                if (d.flags & P.DESC_ACTIVE) and has:
                    h.update(np.ascontiguousarray(gm.capture_vals(i, n, c, nv)[0::3]).tobytes())
    assert h.hexdigest() == want["levels_digest"]
    for k in range(K):
        F = gm.predictions(k)
        assert np.allclose(F[:32], want["F_head"][k], rtol=1e-6, atol=1e-7)
        assert abs(float(np.sum(F)) - want["F_sum"][k]) <= 1e-6 * max(1.0, abs(want["F_sum"][k]))
    gm.close()
