#!/usr/bin/env python
"""BASELINE configs[4b] shape: Epsilon-like 400k x 2000 dense fp32 columns, nbins 64, depth 5; 7-class multinomial (the family the
config states) and bernoulli. Per-iteration time."""
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
N, Cc = 400_000, 2000
X = rng.normal(size=(Cc, N)).astype(np.float32)
X /= np.sqrt((X.astype(np.float64) ** 2).sum(axis=0)).astype(np.float32)  # unit-norm rows
cols = [X[c].astype(np.float64) for c in range(Cc)]
z = np.stack([X[k * 3] * 40 + X[k * 3 + 1] * 20 for k in range(7)], axis=1)
y7 = np.argmax(z + rng.gumbel(size=z.shape), axis=1).astype(np.float64)
for name, kw, y in (("multinomial 7-class", dict(distribution="multinomial", nclass=7), y7),
                    ("bernoulli", dict(distribution="bernoulli"), (y7 < 3).astype(np.float64))):
    gm = P.GBM(ntrees=4, max_depth=5, nbins=64, seed=1234, **kw)
    gm.load(cols, y)
    eng = gm.engine()
    gm.train(1)
    eng.stats_reset()
    t0 = time.perf_counter()
    gm.train(2)
    dt = (time.perf_counter() - t0) / 2
    s = eng.stats()
    print("epsilon shape %d x %d, %s, depth 5: %.1f ms/iteration  kernels: hist %.1f score %.2f split %.2f finalize %.2f ; hist %.0f GB/s algorithmic" % (
        N, Cc, name, dt * 1e3, s.ms_hist / 2, s.ms_score / 2, s.ms_split / 2, s.ms_finalize / 2, s.hist_algo_bytes / (s.ms_hist * 1e-3) / 1e9))
    gm.close()
