#!/usr/bin/env python
"""BASELINE configs[3] shape: airlines-like N x 12 (8 integer columns, some with NAs, + 4 categoricals with 300 / 300 / 5000 /
20000 levels, Zipf), gaussian response, depth 5, nbins 20, nbins_cats 1024. Columns are generated and uploaded one at a time, so the
host never holds the frame. Per-tree time and stage breakdown."""
import ctypes as C
import importlib
import os
os.environ.setdefault("H2OT_STAGE_TIMERS", "1")  # these harnesses print every stage's time
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = importlib.import_module("h2o-3_b200")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
try:
    import psutil
    N = min(N, int(psutil.virtual_memory().available * 0.25 / 8 / 4))  # a few N-sized float64 temporaries at most
except Exception:
    pass
rng = np.random.default_rng(42)
ranges = [(1987, 2008), (1, 12), (1, 31), (1, 7), (0, 2400), (0, 2400), (-60, 600), (11, 4983)]
cards = [300, 300, 5000, 20000]
gm = P.GBM(distribution="gaussian", nclass=1, ntrees=8, max_depth=5, nbins=20, nbins_cats=1024, seed=1234)
L = gm.L
gm.nrows, gm.ncols = N, 12
gm.g = C.c_void_p()
P._check(L.h2ot_gbm_create(C.byref(gm.params), 0, N, 12, 0, 0, 0, 0, C.byref(gm.g)))
y = np.zeros(N)
t0 = time.perf_counter()
for c in range(12):
    if c < 8:
        lo, hi = ranges[c]
        v = rng.integers(lo, hi + 1, N).astype(np.float64)
        if c in (4, 6):
            v[rng.random(N) < 0.02] = np.nan
        if c in (5, 6):
            y += np.nan_to_num(v) * (0.01 if c == 5 else 0.5)
        P._check(L.h2ot_gbm_set_column(gm.g, c, v.ctypes.data_as(C.c_void_p), 0, 0))
    else:
        card = cards[c - 8]
        v = np.minimum(rng.zipf(1.3, N) - 1, card - 1).astype(np.float64)
        y += (v % 7) * 2.0
        P._check(L.h2ot_gbm_set_column(gm.g, c, v.ctypes.data_as(C.c_void_p), 1, card))
    del v
y += rng.normal(size=N) * 10
y = y.astype(np.float32).astype(np.float64)
P._check(L.h2ot_gbm_set_response(gm.g, y.ctypes.data_as(C.c_void_p)))
P._check(L.h2ot_gbm_init(gm.g))
t_load = time.perf_counter() - t0
eng = gm.engine()
gm.train(2)
eng.stats_reset()
t0 = time.perf_counter()
gm.train(4)
dt = (time.perf_counter() - t0) / 4
s = eng.stats()
print("airlines shape %d x 12 (4 categoricals up to 20000 levels), depth 5: %.1f ms/tree  kernels: hist %.1f score %.2f split %.2f finalize %.2f "
      "residuals %.2f gamma %.2f addtree %.2f ; hist %.0f GB/s algorithmic ; generate+load %.0f s" % (
          N, dt * 1e3, s.ms_hist / 4, s.ms_score / 4, s.ms_split / 4, s.ms_finalize / 4, s.ms_residuals / 4, s.ms_gamma / 4, s.ms_addtree / 4,
          s.hist_algo_bytes / (s.ms_hist * 1e-3) / 1e9, t_load))
