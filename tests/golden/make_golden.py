#!/usr/bin/env python
"""Regenerates tests/golden/gbm_small.json from the CPU oracle (the reference itself is Java and cannot run in this image,
see DESIGN.md section 5). The fixture pins the oracle against drift and gives the GPU tests a committed target:
tree structure, split columns/bins/NA directions, float split values and leaf predictions (as hex), per-level histogram
digests and final predictions digests for three small seeded jobs.   python tests/golden/make_golden.py"""
import hashlib
import json
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import h2ot_lib  # noqa: E402
import oracle_lib as O  # noqa: E402

JOBS = {
    "bernoulli": dict(frame=dict(n=4000, c=5, seed=101), params=dict(ntrees=3, max_depth=4, nbins=20, seed=1234), family="bernoulli"),
    "gaussian_sampled": dict(frame=dict(n=3000, c=5, seed=102), params=dict(ntrees=2, max_depth=3, nbins=16, seed=77, sample_rate=0.7),
                             family="gaussian"),
    "multinomial": dict(frame=dict(n=3000, c=4, seed=103, with_cat=False), params=dict(ntrees=2, max_depth=3, nbins=20, seed=5), family="multinomial"),
}


def job_data(name):
    j = JOBS[name]
    X, y, cats = h2ot_lib.make_frame(**j["frame"])
    if j["family"] == "gaussian":
        y = (np.nan_to_num(X[:, 0]) * 2 - np.nan_to_num(X[:, 1])).astype(np.float32).astype(np.float64)
    if j["family"] == "multinomial":
        y = np.argmax(np.nan_to_num(X[:, :3]), axis=1).astype(np.float64)
    return X, y, cats


def f32hex(v):
    return struct.pack("<f", float(v)).hex()


def tree_record(nodes):
    out = []
    for n in nodes:
        rec = dict(kind=n.kind, pid=n.pid, nid=n.nid, nid0=n.nid0, nid1=n.nid1)
        if n.has_split:
            s = n.split if hasattr(n, "split") else n
            rec.update(col=s.col, bin=s.bin, nasplit=s.nasplit, equal=s.equal, splat=f32hex(n.splat), n0=s.n0, n1=s.n1)
            if s.equal >= 2:
                rec["bs"] = bytes(s.bs[:s.bs_nbytes]).hex()
        if n.kind == 2:
            rec["pred"] = f32hex(n.pred)
        out.append(rec)
    return out


def run_oracle(name):
    j = JOBS[name]
    X, y, cats = job_data(name)
    fam = {"gaussian": O.GAUSSIAN, "bernoulli": O.BERNOULLI, "multinomial": O.MULTINOMIAL}[j["family"]]
    nclass = 3 if j["family"] == "multinomial" else (2 if j["family"] == "bernoulli" else 1)
    m = O.Model(O.Frame(X, y, cat_cards=cats), O.default_params(family=fam, nclass=nclass, **j["params"]), capture=True)
    K = 3 if j["family"] == "multinomial" else 1
    rec = dict(init_f=m.init_f, trees=[[tree_record(m.tree(t, k)) for k in range(K)] for t in range(m.ntrees)])
    # histogram counts are exact in every implementation: digest of the w column of every captured histogram
    h = hashlib.sha256()
    ncap = 0
    for i in range(m.ncaptures()):
        c = m.capture(i)
        if c["nleaves"] == 0:  # the reference's pass over a tree with no undecided leaf left: builds nothing (see tests/h2ot_lib.py)
            continue
        ncap += 1
        h.update(c["nids"].tobytes())
        for v in c["vals"]:
            if v.size:
                h.update(np.ascontiguousarray(v[0::3]).tobytes())
    rec["levels_digest"] = h.hexdigest()
    rec["ncaptures"] = ncap
    rec["F_f32_digest"] = [hashlib.sha256(m.F(k).astype(np.float32).tobytes()).hexdigest() for k in range(K)]
    rec["F_head"] = [[float(v) for v in m.F(k)[:32]] for k in range(K)]
    rec["F_sum"] = [float(np.sum(m.F(k))) for k in range(K)]
    return rec


def main():
    out = {name: run_oracle(name) for name in JOBS}
    with open(os.path.join(HERE, "gbm_small.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote", os.path.join(HERE, "gbm_small.json"))


if __name__ == "__main__":
    main()
