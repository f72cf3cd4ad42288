#!/usr/bin/env python
"""BASELINE configs[4a] shape: covtype-like 581k x 54 (10 numeric + 44 binary), 7 classes, nbins 64, depth 5. Per-iteration time."""
import importlib
import os
os.environ.setdefault("H2OT_STAGE_TIMERS", "1")  # these harnesses print every stage's time
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = importlib.import_module("h2o-3_b200")
rng = np.random.default_rng(42)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 581012
X = [rng.normal(size=N).astype(np.float32).astype(np.float64) for _ in range(10)] + [(rng.random(N) < 0.1).astype(np.float64) for _ in range(44)]
z = np.stack([X[k] + 0.5 * X[(k + 3) % 10] for k in range(7)], axis=1)
y = np.argmax(z + rng.gumbel(size=z.shape), axis=1).astype(np.float64)
gm = P.GBM(distribution="multinomial", nclass=7, ntrees=8, max_depth=5, nbins=64, seed=1234)
gm.load(X, y)
eng = gm.engine()
gm.train(2)
eng.stats_reset()
t0 = time.perf_counter()
gm.train(5)
dt = (time.perf_counter() - t0) / 5
s = eng.stats()
print("multinomial %d x 54, 7 classes, depth 5: %.2f ms/iteration (7 trees)  kernels: hist %.2f score %.2f split %.2f finalize %.2f residuals %.2f gamma %.2f addtree %.2f ; launches/iter %d" % (
    N, dt * 1e3, s.ms_hist / 5, s.ms_score / 5, s.ms_split / 5, s.ms_finalize / 5, s.ms_residuals / 5, s.ms_gamma / 5, s.ms_addtree / 5, s.launches // 5))
