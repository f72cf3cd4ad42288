#!/usr/bin/env python
"""BASELINE configs[2]: DRF on the Higgs-shape frame (depth 12, mtries 5, sample_rate 0.63, nbins 20). Per-tree time."""
import importlib
import os
os.environ.setdefault("H2OT_STAGE_TIMERS", "1")  # these harnesses print every stage's time
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

P = importlib.import_module("h2o-3_b200")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 11_000_000
cols, y, _ = bench.make_frame(N, 0, N, pinned=False)
gm = P.DRF(distribution="bernoulli", ntrees=5, max_depth=12, mtries=5, sample_rate=0.63, min_rows=1.0, nbins=20, seed=1234)
gm.load(cols, y)
eng = gm.engine()
gm.train(1)
eng.stats_reset()
t0 = time.perf_counter()
gm.train(3)
dt = (time.perf_counter() - t0) / 3
s = eng.stats()
print("DRF %d x 28 depth 12 mtries 5: %.1f ms/tree  (hist %.1f ms in %d launches, score %.2f, split %.2f, finalize %.2f, collect %.2f) nodes=%d" % (
    N, dt * 1e3, s.ms_hist / 3, s.hist_launches // 3, s.ms_score / 3, s.ms_split / 3, s.ms_finalize / 3, s.ms_addtree / 3, len(gm.tree(1)[0])))
