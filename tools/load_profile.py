#!/usr/bin/env python
"""Where does GBM.load() spend its time? (frame upload from pinned host memory + ingest + init)"""
import ctypes as C
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import torch  # noqa: E402

P = importlib.import_module("h2o-3_b200")
N = 11_000_000
cols, y, hold = bench.make_frame(N, 0, N, pinned=True)
L = P.load()
for rep in range(2):
    torch.cuda.synchronize()
    t = [time.perf_counter()]
    g = C.c_void_p()
    prm = P.GBM(distribution="bernoulli", ntrees=1, **bench.PARAMS).params
    P._check(L.h2ot_gbm_create(C.byref(prm), 0, N, bench.NCOLS, 0, 0, 0, 0, C.byref(g)))
    t.append(time.perf_counter())
    for c in range(bench.NCOLS):
        P._check(L.h2ot_gbm_set_column(g, c, cols[c].ctypes.data_as(C.c_void_p), 0, 0))
    t.append(time.perf_counter())
    P._check(L.h2ot_gbm_set_response(g, y.ctypes.data_as(C.c_void_p)))
    t.append(time.perf_counter())
    P._check(L.h2ot_gbm_init(g))
    t.append(time.perf_counter())
    P._check(L.h2ot_gbm_train(g, 1))
    t.append(time.perf_counter())
    L.h2ot_gbm_destroy(g)
    t.append(time.perf_counter())
    names = ["create", "set_column x28 (%.2f GB)" % (N * 28 * 8 / 1e9), "set_response", "init (finalize frame + init_f + init model)", "first tree", "destroy"]
    print("rep", rep, " | ".join("%s %.1f ms" % (n, 1e3 * (b - a)) for n, a, b in zip(names, t[:-1], t[1:])))
    gb = N * 28 * 8 / 1e9
    print("   column upload rate: %.1f GB/s" % (gb / (t[2] - t[1])))
