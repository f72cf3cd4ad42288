#!/usr/bin/env python
"""bench.py — the hex.tree histogram hot path on B200: GBM bernoulli on a Higgs-shape synthetic frame
(BASELINE.json configs[1]: 11M x 28 numeric, max_depth 6, nbins 256, nbins_top_level 1024).

One "step" = one boosting iteration (GBM.buildNextKTrees: residuals, <=6 level passes of score_decide + histogram
build + split search, GammaPass, AddTreeContributions) on the frame resident in HBM.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this engine
  python bench.py --impl reference [--steps K] [--warmup W]      # CPU restatement of the reference path (oracle port)

Prints ONE JSON line (rank 0). `value` = rows x active-cols histogrammed per second over the K timed steps (all ranks);
`e2e` = the same metric for a whole job from pinned HOST buffers (frame upload + encode + init + K trees + read-back
of predictions) through the public API; `roofline` = the histogram kernel's algorithmic bytes
over its CUDA-event time inside the timed region; `cpu_baseline` = the oracle's restatement timed on host cores.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "hist_build_rows_x_cols_per_s"
UNIT = "rows*cols/s"
PARAMS = dict(max_depth=6, nbins=256, nbins_top_level=1024, nbins_cats=1024, min_rows=10.0, learn_rate=0.1, seed=1234)
NCOLS = 28


def higgs_like_column(col, nrows, seed=42):
    """Seeded per column so that every rank can cut the same global frame. All values are fp32-exact."""
    rng = np.random.default_rng([seed, col])
    if col >= 21:  # 7 "high-level" positive skewed features
        v = rng.lognormal(mean=0.0, sigma=0.5, size=nrows)
    elif col % 7 == 3:  # b-tag like: three values
        v = rng.choice(np.array([0.0, 1.09, 2.17]), size=nrows, p=[0.5, 0.3, 0.2])
    elif col % 7 == 5:  # azimuth
        v = rng.uniform(-np.pi, np.pi, size=nrows)
    elif col % 7 == 1:  # momentum magnitudes
        v = np.abs(rng.standard_normal(nrows)) + 0.25
    else:
        v = rng.standard_normal(nrows)
    return v.astype(np.float32).astype(np.float64)


def make_frame(nrows, r0, r1, pinned):
    """Returns (columns[NCOLS] of float64 arrays for global rows [r0,r1), y). pinned: allocate in pinned host memory."""
    n = r1 - r0
    cols = []
    if pinned:
        import torch
        hold = [torch.empty(n, dtype=torch.float64, pin_memory=True) for _ in range(NCOLS + 1)]
        bufs = [h.numpy() for h in hold]
    else:
        hold = None
        bufs = [np.empty(n, dtype=np.float64) for _ in range(NCOLS + 1)]
    logit = np.zeros(nrows, dtype=np.float64)
    for c in range(NCOLS):
        v = higgs_like_column(c, nrows)
        if c in (0, 2, 4):
            logit += 0.6 * v * (1 if c != 2 else -1)
        elif c == 1:
            logit += 0.4 * (v - 1.0)
        elif c in (21, 22, 25):
            logit += 0.8 * (np.log(v) if c != 22 else -np.log(v))
        elif c == 3:
            logit += 0.3 * (v > 1.0)
        bufs[c][:] = v[r0:r1]
        cols.append(bufs[c])
    logit += 0.5 * higgs_like_column(0, nrows) * higgs_like_column(4, nrows) + 0.12
    u = np.random.default_rng([42, 1000]).random(nrows)
    bufs[NCOLS][:] = (u < 1.0 / (1.0 + np.exp(-logit)))[r0:r1]
    return cols, bufs[NCOLS], hold


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "10"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """dram bytes per histogram-kernel launch from the committed ncu capture (profiles/hist_traffic.json), else None."""
    p = os.path.join(ROOT, "profiles", "hist_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def cpu_baseline(nrows_sample, budget_s, threads):
    """Oracle port (C++ restatement of the reference CPU path, NOT the JVM) on the first nrows_sample rows, all host
    threads; runs whole trees until ~budget_s of work. Returns (rows*cols/s, sec/tree, trees, wall)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    cols, y, _ = make_frame(nrows_sample, 0, nrows_sample, pinned=False)
    X = np.stack(cols, axis=1)
    fr = O.Frame(X, y)
    del X, cols
    trees, wall, rowcols = 0, 0.0, 0.0
    nt = 1
    while True:
        p = O.default_params(family=O.BERNOULLI, nclass=2, ntrees=nt, nthreads=threads,
                             chunk_rows=max(4096, nrows_sample // (4 * max(1, threads))), **PARAMS)
        m = O.Model(fr, p)
        t_hist, t_total, rc = m.timing()
        trees, wall, rowcols = nt, t_total, rc
        del m
        if wall >= 0.5 * budget_s or nt >= 64:
            break
        nt = min(64, max(nt + 1, int(nt * budget_s / max(wall, 1e-3) * 0.8)))
    return rowcols / wall, wall / trees, trees, wall


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = min(args.rows, args.ref_rows)
    steps, warm = max(1, args.steps), max(0, args.warmup)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    cols, y, _ = make_frame(sample, 0, sample, pinned=False)
    X = np.stack(cols, axis=1)
    fr = O.Frame(X, y)
    del X, cols
    chunk = max(4096, sample // (4 * threads))
    # each step = one boosting iteration on the sample; the oracle trains from scratch, so run warm+steps trees once
    # and time only the last `steps` of them by differencing two runs (trees are identical up to that point)
    t_w = 0.0
    rc_w = 0.0
    if warm > 0:
        mw = O.Model(fr, O.default_params(family=O.BERNOULLI, nclass=2, ntrees=warm, nthreads=threads, chunk_rows=chunk, **PARAMS))
        _, t_w, rc_w = mw.timing()
        del mw
    m = O.Model(fr, O.default_params(family=O.BERNOULLI, nclass=2, ntrees=warm + steps, nthreads=threads, chunk_rows=chunk, **PARAMS))
    _, t_all, rc_all = m.timing()
    dt = max(1e-9, t_all - t_w)
    rc = rc_all - rc_w
    val = rc / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": 1e3 * dt / steps, "sec_per_tree": dt / steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "GBM bernoulli Higgs-shape %dx%d (bounded sample of 11Mx28) depth 6 nbins 256" % (sample, NCOLS),
                   "rows": sample, "cols": NCOLS, **{k: PARAMS[k] for k in ("max_depth", "nbins", "nbins_top_level")}},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "first %d rows of the 11Mx28 frame, %d trees timed after %d warm-up trees; C++ restatement of "
                                   "the reference CPU path (no JVM on this box)" % (sample, steps, warm)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_b200(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    importlib.import_module("h2o-3_b200.build").build()
    P = importlib.import_module("h2o-3_b200")

    N = args.rows
    r0, r1 = (N * rank) // world, (N * (rank + 1)) // world
    cols, y, hold = make_frame(N, r0, r1, pinned=True)
    comm = None
    if world > 1:
        uid = torch.zeros(P.COMM_ID_BYTES, dtype=torch.uint8)
        if rank == 0:
            uid = torch.from_numpy(P.comm_unique_id().copy())
        uid = uid.cuda()
        dist.broadcast(uid, 0)
        comm = (world, rank, uid.cpu().numpy())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def new_gbm():
        g = P.GBM(distribution="bernoulli", ntrees=args.steps + args.warmup, device=local, **PARAMS)
        return g

    # ---------------- device-resident arm: W warm-up + K timed boosting iterations
    gm = new_gbm()
    gm.load(cols, y, row_offset=r0, nrows_global=N, comm=comm)
    eng = gm.engine()
    for _ in range(args.warmup):
        gm.train(1)
    barrier()
    s0 = eng.stats()
    t0s = gm.timing()
    t0s = type('T', (), dict(levels=t0s.levels, level_s=t0s.level_s, total_s=t0s.total_s, hist_rows_cols=t0s.hist_rows_cols))
    snap0 = {f: getattr(s0, f) for f, _ in s0._fields_}
    rc0 = t0s.hist_rows_cols
    clk = ClockSampler(local)
    if rank == 0:
        clk.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        gm.train(1)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = clk.stop() if rank == 0 else None
    s1 = eng.stats()
    d = {f: getattr(s1, f) - snap0[f] for f, _ in s1._fields_}
    rowcols = gm.timing().hist_rows_cols - rc0  # global rows x active cols over the timed level passes
    t = torch.tensor([ms, float(d["launches"])], dtype=torch.float64, device="cuda")
    if world > 1:
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ts = t.clone()
        dist.all_reduce(ts, op=dist.ReduceOp.SUM)
        ms, launches = float(tm[0]), int(ts[1])
    else:
        launches = int(d["launches"])
    value = rowcols / (ms * 1e-3)
    peak, peak_src = measured_peak()
    hist_ms = d["ms_hist"]
    n_hl = max(1, d["hist_launches"])
    ach = d["hist_algo_bytes"] / (hist_ms * 1e-3) / 1e9 if hist_ms > 0 else 0.0
    traffic = ncu_traffic()
    roofline = {"bound": "hbm", "kernel": "hist_smem_kernel", "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": d["hist_algo_bytes"] / n_hl, "ms_per_launch": hist_ms / n_hl,
                "launches": int(d["hist_launches"]),
                "kernel_rows_x_cols_per_s": d["hist_rows_cols"] / (hist_ms * 1e-3) if hist_ms > 0 else 0.0,
                "stage_ms_per_step": {k: d[k] / args.steps for k in ("ms_score", "ms_hist", "ms_finalize", "ms_split", "ms_residuals",
                                                                        "ms_gamma", "ms_addtree")}}
    # secondary ceiling (SURVEY 8d asks for it): the kernel's time goes into shared-memory atomics, 5 per (row, column) item with
    # unit weights; random-bin ATOMS.ADD runs at 12.4 lanes/clk/SM = 3490 G/s on this part (conflict free: 31 lanes/clk/SM;
    # tools/ubench_atomics2.cu "clean" lines, profiles/r01_ubench_atomics2_b200.txt)
    atoms = 5.0 * d["hist_rows_cols"] / (hist_ms * 1e-3) / 1e9 if hist_ms > 0 else 0.0
    roofline["smem_atomics"] = {"achieved": atoms, "peak": 3490.0, "unit": "Gatom/s", "frac": atoms / 3490.0,
                                "peak_source": "measured, tools/ubench_atomics2.cu (random-bin pattern) on B200"}
    t1s = gm.timing()
    levels_per_step = (t1s.levels - t0s.levels) / args.steps
    roofline["host_ms_per_step"] = {"inside_level_calls": 1e3 * (t1s.level_s - t0s.level_s) / args.steps,
                                    "whole_step_wall": 1e3 * (t1s.total_s - t0s.total_s) / args.steps}
    gm.close()

    # ---------------- end-to-end arm: whole job from pinned host buffers through the public API
    barrier()
    g2 = new_gbm()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pred_hold = torch.empty(len(y), dtype=torch.float64).pin_memory()  # results land in pinned memory, like the inputs leave it
    f0.record()
    g2.load(cols, y, row_offset=r0, nrows_global=N, comm=comm)
    g2.train(args.steps)
    preds = g2.predictions(out=pred_hold.numpy())
    ntree_nodes = sum(len(g2.tree(t)[0]) for t in range(g2.ntrees))
    f1.record()
    barrier()
    ms2 = f0.elapsed_time(f1)
    st2 = g2.engine().stats()
    rc2 = g2.timing().hist_rows_cols
    t2 = torch.tensor([ms2], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        ms2 = float(t2[0])
    e2e = {"value": rc2 / (ms2 * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(st2.h2d_bytes / args.steps),
           "d2h_bytes_per_step": int(st2.d2h_bytes / args.steps), "ms_total": ms2, "sec_per_tree": ms2 * 1e-3 / args.steps,
           "what": "GBM.load (frame upload from pinned host + rollups + encode) + init + %d trees + read-back of the predictions (into pinned memory) and "
                   "the trees (%d nodes)" % (args.steps, ntree_nodes)}
    assert np.isfinite(preds).all()
    g2.close()

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "sec_per_tree": ms * 1e-3 / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": "GBM bernoulli Higgs-shape %dx%d, max_depth 6, nbins 256, nbins_top_level 1024 (BASELINE configs[1])" % (N, NCOLS),
                       "rows": N, "cols": NCOLS, "levels_per_step": levels_per_step, "parallelism": "rows sharded over %d GPU(s)" % world,
                       "l2": "inputs larger than L2 (frame %.2f GB resident)" % (N * NCOLS * 4 / 1e9),
                       "arithmetic": "bins in f64 (reference formula), sums of w, w*y, w*y*y as int64 fixed point, float-cast like reducePrecision",
                       **{k: PARAMS[k] for k in ("max_depth", "nbins", "nbins_top_level")}},
            "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, spt, trees, wall = cpu_baseline(min(N, args.ref_rows), args.cpu_budget_s, threads)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sec_per_tree_on_sample": spt,
                                    "sample": "first %d rows of the frame, %d trees, %.1f s wall; C++ restatement of the reference CPU "
                                              "path (oracle), not the JVM (no JDK on this box)" % (min(N, args.ref_rows), trees, wall)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=11_000_000)
    ap.add_argument("--ref-rows", type=int, default=1_100_000, help="rows of the bounded CPU sample")
    ap.add_argument("--cpu-budget-s", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
