#!/usr/bin/env python
"""Per-level timing of the histogram pass on a Higgs-shape frame (engine-level harness, not the contract bench).
Rows are dealt to L leaves at random, each (leaf, col) gets an nbins histogram over the column's range.
  python tools/hist_bench.py [--rows 11000000] [--reps 5]
Prints per configuration: ms per launch (engine CUDA events), rows x cols / s, algorithmic GB/s."""
import argparse
import ctypes as C
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (data generator)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=11_000_000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--configs", default="1:1024,2:512,4:256,8:256,16:256,32:256")
    ap.add_argument("--mode", type=int, default=0, help="1: force the global-atomics kernel")
    args = ap.parse_args()
    importlib.import_module("h2o-3_b200.build").build()
    P = importlib.import_module("h2o-3_b200")
    n, c = args.rows, bench.NCOLS
    cols, y, _ = bench.make_frame(n, 0, n, pinned=False)
    e = P.Engine(nrows=n, ncols=c)
    for j in range(c):
        e.load_column(j, cols[j])
    e.load_response(y)
    e.finalize_frame()
    e.init_model([0.0])
    e.set_hist_mode(args.mode)
    rng = np.random.default_rng(0)
    work = (y - 0.53).astype(np.float32).astype(np.float64) + (rng.random(n) * 1e-3).astype(np.float32)
    e.set_work(work)
    e.begin_tree(0, 0, 1.0, float(np.max(np.abs(work))))
    lo = [float(np.min(v)) for v in cols]
    hi = [float(np.nextafter(np.max(v), np.inf)) for v in cols]
    for cfg in args.configs.split(","):
        L, nb = (int(t) for t in cfg.split(":"))
        leaf0 = 0 if L == 1 else 1000
        if L > 1:
            nids = (leaf0 + rng.integers(0, L, n)).astype(np.int32)
            e.set_nids(nids)
            dn = (P.DecidedNode * L)()
            for l in range(L):
                dn[l] = P.DecidedNode(nid=leaf0 + l, pid=-1, col=-1, equal=0, nasplit=0, splat=float("nan"), child=(C.c_int32 * 2)(-1, -1))
        else:
            e.set_nids(np.zeros(n, dtype=np.int32))
            dn = None
        descs = (P.HistDesc * (L * c))()
        for l in range(L):
            for j in range(c):
                descs[l * c + j] = P.make_desc(lo[j], hi[j], nb, order=j)
        e.build_level(descs, leaf_offset=leaf0, decided=dn)  # warm-up
        if L == 1:
            e.finish_level(want_vals=False)  # the root level feeds the planner's per-column bin concentration
        e.stats_reset()
        for _ in range(args.reps):
            e.build_level(descs, leaf_offset=leaf0, decided=dn)
        s = e.stats()
        ms = s.ms_hist / s.hist_launches
        print("L=%3d nbins=%5d : hist %.3f ms/launch (%d launches/level)  score %.3f ms  %.1f G rows*cols/s  %.0f GB/s algorithmic (%.1f%% of 6533.5)" % (
            L, nb, s.ms_hist / args.reps, s.hist_launches // args.reps, s.ms_score / args.reps, n * c / (s.ms_hist / args.reps) / 1e6,
            s.hist_algo_bytes / args.reps / (s.ms_hist / args.reps) / 1e6, 100 * s.hist_algo_bytes / args.reps / (s.ms_hist / args.reps) / 1e6 / 6533.5), flush=True)
    e.close()


if __name__ == "__main__":
    main()
