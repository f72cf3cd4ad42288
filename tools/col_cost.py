#!/usr/bin/env python
"""Histogram time of the root level by column kind on the Higgs-shape frame: which columns cost what (ms per level per column)."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

P = importlib.import_module("h2o-3_b200")
N = 11_000_000
cols, y, _ = bench.make_frame(N, 0, N, pinned=False)
C = len(cols)
e = P.Engine(nrows=N, ncols=C, family="bernoulli", nclass=2)
for c in range(C):
    e.load_column(c, cols[c])
e.load_response(y)
e.finalize_frame()
e.init_model([0.0])
e.set_work((y - 0.5).astype(np.float32).astype(np.float64))
e.begin_tree(0, 0, 1.0, 0.5)
nuniq = [len(np.unique(cols[c][:200000])) for c in range(C)]
kinds = {"few-valued (<= 8 values)": [c for c in range(C) if nuniq[c] <= 8], "continuous": [c for c in range(C) if nuniq[c] > 8]}
for name, sel in list(kinds.items()) + [("all", list(range(C)))]:
    if not sel:
        continue
    descs = (P.HistDesc * C)()
    for c in range(C):
        v = cols[c]
        lo, hi = float(np.nanmin(v)), float(np.nanmax(v))
        descs[c] = P.make_desc(lo, np.nextafter(hi, np.inf), 1024, init_na=False, order=c, active=c in sel)
    for rep in range(3):
        e.stats_reset()
        e.build_level(descs)
        s = e.stats()
        if rep == 0:
            e.finish_level(want_vals=False)  # feeds the planner's per-column bin concentration (replicas / warp aggregation)
    print("%-26s %2d columns: %.3f ms per level, %.4f ms per column" % (name, len(sel), s.ms_hist, s.ms_hist / len(sel)))
print("per column (alone at the root, 1024 bins):")
for c in range(C):
    descs = (P.HistDesc * C)()
    for j in range(C):
        v = cols[j]
        lo, hi = float(np.nanmin(v)), float(np.nanmax(v))
        descs[j] = P.make_desc(lo, np.nextafter(hi, np.inf), 1024, init_na=False, order=j, active=j == c)
    for rep in range(3):
        e.stats_reset()
        e.build_level(descs)
        s = e.stats()
        if rep == 0:
            e.finish_level(want_vals=False)
    v = cols[c][:400000]
    b = np.minimum(((v - v.min()) * (1024 / (np.nextafter(v.max(), np.inf) - v.min()))).astype(np.int64), 1023)
    p = np.bincount(b, minlength=1024) / b.size
    print("  col %2d  %6d distinct  share-a-bin prob %.4f : %.4f ms" % (c, nuniq[c], float((p * p).sum()), s.ms_hist))
e.close()
