#!/usr/bin/env python
"""bench.py — the hex.tree histogram hot path on B200. Default workload: GBM bernoulli on a Higgs-shape synthetic frame
(BASELINE.json configs[1]: 11M x 28 numeric, max_depth 6, nbins 256, nbins_top_level 1024); `--config` selects the other
BASELINE configurations (c1 100k x 10, c3 DRF depth 12 mtries 5, c4 airlines shape 100M x 12 with categoricals, c5a covtype shape
7 classes, c5b Epsilon shape 400k x 2000 7 classes), each with the same roofline / e2e / cpu_baseline treatment.

One "step" = one boosting iteration (GBM.buildNextKTrees: residuals, <=max_depth level passes of score_decide + histogram
build + split search, GammaPass, AddTreeContributions; K class-trees for multinomial; DRF: one tree) on the frame resident in HBM.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config c2]   # this engine
  python bench.py --impl reference [--steps K] [--warmup W]           # CPU restatement of the reference path (oracle port), same config

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
ATOMS_PER_ITEM = 5.0  # shared-memory atomics per (row, column) item in hist_smem_kernel with unit weights
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


def _buffers(n, ncols, pinned):
    if pinned:
        import torch
        hold = [torch.empty(n, dtype=torch.float64, pin_memory=True) for _ in range(ncols + 1)]
        return [h.numpy() for h in hold], hold
    return [np.empty(n, dtype=np.float64) for _ in range(ncols + 1)], None


def make_frame(nrows, r0, r1, pinned):
    """Higgs shape. Returns (columns[NCOLS] of float64 arrays for global rows [r0,r1), y, keep-alive). pinned: pinned host memory."""
    n = r1 - r0
    cols = []
    bufs, hold = _buffers(n, NCOLS, pinned)
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


def _frame_c1(nrows, r0, r1, pinned):
    """BASELINE configs[0]: 100k x 10 numeric, bernoulli."""
    bufs, hold = _buffers(r1 - r0, 10, pinned)
    rng = np.random.default_rng(7)
    X = rng.normal(size=(nrows, 10)).astype(np.float32).astype(np.float64)
    y = (rng.random(nrows) < 1 / (1 + np.exp(-(X[:, 0] - 0.5 * X[:, 1] + 0.3 * X[:, 2] * X[:, 3])))).astype(np.float64)
    for c in range(10):
        bufs[c][:] = X[r0:r1, c]
    bufs[10][:] = y[r0:r1]
    return bufs[:10], bufs[10], hold


def _frame_c4(nrows, r0, r1, pinned):
    """BASELINE configs[3] shape: airlines-like, 8 integer columns (two with NAs) + 4 Zipf categoricals (300 / 300 / 5000 / 20000 levels)."""
    bufs, hold = _buffers(r1 - r0, 12, pinned)
    ranges = [(1987, 2008), (1, 12), (1, 31), (1, 7), (0, 2400), (0, 2400), (-60, 600), (11, 4983)]
    cards = [300, 300, 5000, 20000]
    y = np.zeros(nrows)
    for c in range(12):
        rng = np.random.default_rng([43, c])
        if c < 8:
            lo, hi = ranges[c]
            v = rng.integers(lo, hi + 1, nrows).astype(np.float64)
            if c in (4, 6):
                v[rng.random(nrows) < 0.02] = np.nan
            if c in (5, 6):
                y += np.nan_to_num(v) * (0.01 if c == 5 else 0.5)
        else:
            v = np.minimum(rng.zipf(1.3, nrows) - 1, cards[c - 8] - 1).astype(np.float64)
            y += (v % 7) * 2.0
        bufs[c][:] = v[r0:r1]
        del v
    y += np.random.default_rng([43, 100]).normal(size=nrows) * 10
    bufs[12][:] = y.astype(np.float32)[r0:r1]
    return bufs[:12], bufs[12], hold


def _frame_c5a(nrows, r0, r1, pinned):
    """BASELINE configs[4] covtype shape: 10 numeric + 44 binary columns, 7 classes."""
    bufs, hold = _buffers(r1 - r0, 54, pinned)
    rng = np.random.default_rng(44)
    X = [rng.normal(size=nrows).astype(np.float32).astype(np.float64) for _ in range(10)] + [(rng.random(nrows) < 0.1).astype(np.float64) for _ in range(44)]
    z = np.stack([X[k] + 0.5 * X[(k + 3) % 10] for k in range(7)], axis=1)
    y = np.argmax(z + rng.gumbel(size=z.shape), axis=1).astype(np.float64)
    for c in range(54):
        bufs[c][:] = X[c][r0:r1]
    bufs[54][:] = y[r0:r1]
    return bufs[:54], bufs[54], hold


def _frame_c5b(nrows, r0, r1, pinned):
    """BASELINE configs[4] Epsilon shape: 2000 dense fp32 columns, unit-norm rows, 7 classes."""
    Cc = 2000
    bufs, hold = _buffers(r1 - r0, Cc, pinned)
    rng = np.random.default_rng(45)
    X = rng.normal(size=(Cc, nrows)).astype(np.float32)
    X /= np.sqrt((X.astype(np.float64) ** 2).sum(axis=0)).astype(np.float32)
    z = np.stack([X[k * 3] * 40 + X[k * 3 + 1] * 20 for k in range(7)], axis=1)
    y = np.argmax(z + rng.gumbel(size=z.shape), axis=1).astype(np.float64)
    for c in range(Cc):
        bufs[c][:] = X[c, r0:r1]
    bufs[Cc][:] = y[r0:r1]
    return bufs[:Cc], bufs[Cc], hold


# name -> workload. "model": keyword arguments of the product's GBM / DRF; "oracle": the matching oracle parameters.
WORKLOADS = {
    "c1": dict(rows=100_000, ncols=10, frame=_frame_c1, cls="GBM", cats={}, K=1,
               text="GBM bernoulli synthetic %dx%d numeric, max_depth 5, nbins 20 (BASELINE configs[0])",
               model=dict(distribution="bernoulli", max_depth=5, nbins=20, min_rows=10.0, learn_rate=0.1, seed=1234)),
    "c2": dict(rows=11_000_000, ncols=NCOLS, frame=make_frame, cls="GBM", cats={}, K=1,
               text="GBM bernoulli Higgs-shape %dx%d, max_depth 6, nbins 256, nbins_top_level 1024 (BASELINE configs[1])",
               model=dict(distribution="bernoulli", **PARAMS)),
    "c3": dict(rows=11_000_000, ncols=NCOLS, frame=make_frame, cls="DRF", cats={}, K=1,
               text="DRF Higgs-shape %dx%d, max_depth 12, mtries 5, sample_rate 0.63, nbins 20 (BASELINE configs[2])",
               model=dict(distribution="bernoulli", max_depth=12, mtries=5, sample_rate=0.63, min_rows=1.0, nbins=20, seed=1234)),
    "c4": dict(rows=100_000_000, ncols=12, frame=_frame_c4, cls="GBM", cats={8: 300, 9: 300, 10: 5000, 11: 20000}, K=1,
               text="GBM gaussian airlines-shape %dx%d mixed, 4 categoricals up to 20000 levels, max_depth 5, nbins 20, nbins_cats 1024 (BASELINE configs[3])",
               model=dict(distribution="gaussian", nclass=1, max_depth=5, nbins=20, nbins_cats=1024, seed=1234)),
    "c5a": dict(rows=581_012, ncols=54, frame=_frame_c5a, cls="GBM", cats={}, K=7,
                text="GBM multinomial 7-class covtype-shape %dx%d, max_depth 5, nbins 64 (BASELINE configs[4])",
                model=dict(distribution="multinomial", nclass=7, max_depth=5, nbins=64, seed=1234)),
    "c5b": dict(rows=400_000, ncols=2000, frame=_frame_c5b, cls="GBM", cats={}, K=7,
                text="GBM multinomial 7-class Epsilon-shape %dx%d dense, max_depth 5, nbins 64 (BASELINE configs[4])",
                model=dict(distribution="multinomial", nclass=7, max_depth=5, nbins=64, seed=1234)),
}


def oracle_model(O, fr, wl, ntrees, threads, chunk):
    """The oracle's build of workload wl (C++ restatement of the reference CPU path): returns an oracle Model."""
    m = dict(wl["model"])
    fam = {"bernoulli": O.BERNOULLI, "gaussian": O.GAUSSIAN, "multinomial": O.MULTINOMIAL}[m.pop("distribution")]
    nclass = m.pop("nclass", 2)
    mtries = m.pop("mtries", -1)
    p = O.default_params(family=fam, nclass=nclass, ntrees=ntrees, nthreads=threads, chunk_rows=chunk, **m)
    return O.Model(fr, p, drf=wl["cls"] == "DRF", mtries=mtries)


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
                                       "-lms", os.environ.get("H2OT_BENCH_CLOCK_MS", "10")], stdout=self.f, stderr=subprocess.DEVNULL)
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


def ncu_traffic(config, rows_per_gpu):
    """dram bytes per histogram-kernel launch from the committed ncu capture (profiles/hist_traffic.json) -- only when that
    capture was taken on this config and this many rows per GPU; otherwise None (the bytes scale with the shard)."""
    p = os.path.join(ROOT, "profiles", "hist_traffic.json")
    if os.path.exists(p):
        try:
            j = json.load(open(p))
            if j.get("config", "c2") == config and int(j.get("rows_per_gpu", 11_000_000)) == int(rows_per_gpu):
                return j.get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def cpu_baseline(wl, nrows_sample, budget_s, threads):
    """Oracle port (C++ restatement of the reference CPU path, NOT the JVM) on the first nrows_sample rows, all host
    threads; runs whole iterations until ~budget_s of work. Returns (rows*cols/s, sec/iteration, iterations, wall)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    cols, y, _ = wl["frame"](nrows_sample, 0, nrows_sample, False)
    X = np.stack(cols, axis=1)
    fr = O.Frame(X, y, cat_cards=wl["cats"])
    del X, cols
    trees, wall, rowcols = 0, 0.0, 0.0
    nt = 1
    chunk = max(4096, nrows_sample // (4 * max(1, threads)))
    while True:
        m = oracle_model(O, fr, wl, nt, threads, chunk)
        t_hist, t_total, rc = m.timing()
        trees, wall, rowcols = nt, t_total, rc
        del m
        if wall >= 0.5 * budget_s or nt >= 64:
            break
        nt = min(64, max(nt + 1, int(nt * budget_s / max(wall, 1e-3) * 0.8)))
    return rowcols / wall, wall / trees, trees, wall


def run_reference(args):
    """Reference arm: the oracle's multithreaded restatement of the reference CPU path (no JVM here) on the SAME config
    and the same number of rows as the engine's arm, all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.config]
    rows = args.rows or wl["rows"]
    threads = os.cpu_count() or 1
    sample = min(rows, args.ref_rows) if args.ref_rows else rows
    steps, warm = max(1, args.steps), max(0, args.warmup)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    cols, y, _ = wl["frame"](sample, 0, sample, False)
    X = np.stack(cols, axis=1)
    fr = O.Frame(X, y, cat_cards=wl["cats"])
    del X, cols
    chunk = max(4096, sample // (4 * threads))
    # each step = one boosting iteration; the oracle trains from scratch, so run warm and warm+steps iterations and time the
    # last `steps` of them by differencing the two runs (the trees are identical up to that point)
    t_w = 0.0
    rc_w = 0.0
    if warm > 0:
        mw = oracle_model(O, fr, wl, warm, threads, chunk)
        _, t_w, rc_w = mw.timing()
        del mw
    m = oracle_model(O, fr, wl, warm + steps, threads, chunk)
    _, t_all, rc_all = m.timing()
    dt = max(1e-9, t_all - t_w)
    rc = rc_all - rc_w
    val = rc / dt
    what = "all %d rows" % sample if sample == rows else "first %d of %d rows" % (sample, rows)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": 1e3 * dt / steps, "sec_per_tree": dt / steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["text"] % (sample, wl["ncols"]), "name": args.config, "rows": sample, "cols": wl["ncols"],
                   **{k: v for k, v in wl["model"].items() if k in ("max_depth", "nbins", "nbins_top_level", "mtries", "sample_rate")}},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "%s of the frame, %d iterations timed after %d warm-up iterations; C++ restatement of "
                                   "the reference CPU path (no JVM on this box)" % (what, steps, warm)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def tree_bytes(gm, K):
    out = []
    for t in range(gm.ntrees):
        for k in range(K):
            nodes, ls = gm.tree(t, k)
            out.append(bytes(nodes))
            out.append(int(ls).to_bytes(4, "little", signed=True))
    return b"".join(out)


def run_b200(args):
    import hashlib

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

    wl = WORKLOADS[args.config]
    N = args.rows or wl["rows"]
    K = wl["K"]
    r0, r1 = (N * rank) // world, (N * (rank + 1)) // world
    cols, y, hold = wl["frame"](N, r0, r1, True)
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

    PROF_STEPS = 3  # untimed steps after the timed region, with every stage's CUDA-event timer on (stage_ms_per_step)

    def new_gbm(ntrees=None, device=local):
        return getattr(P, wl["cls"])(ntrees=ntrees or (args.steps + args.warmup + PROF_STEPS), device=device, **wl["model"])

    # ---------------- device-resident arm: W warm-up + K timed boosting iterations
    gm = new_gbm()
    gm.load(cols, y, categorical=wl["cats"], row_offset=r0, nrows_global=N, comm=comm)
    eng = gm.engine()
    for _ in range(args.warmup):
        gm.train(1)
    gm.flush()  # (pipelined builds return with the last tree still on the device)
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
    gm.flush()  # the host waits for the last tree and mirrors it: the events below bracket the wall time of the K steps
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = clk.stop() if rank == 0 else None
    s1 = eng.stats()
    t1s = gm.timing()
    t1s = type('T', (), dict(levels=t1s.levels, level_s=t1s.level_s, total_s=t1s.total_s, hist_rows_cols=t1s.hist_rows_cols))
    d = {f: getattr(s1, f) - snap0[f] for f, _ in s1._fields_}
    rowcols = t1s.hist_rows_cols - rc0  # global rows x active cols over the timed level passes
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
    traffic = ncu_traffic(args.config, r1 - r0)
    roofline = {"bound": "hbm", "kernel": "hist_smem_kernel", "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": d["hist_algo_bytes"] / n_hl, "ms_per_launch": hist_ms / n_hl,
                "launches": int(d["hist_launches"]),
                "kernel_rows_x_cols_per_s": d["hist_rows_cols"] / (hist_ms * 1e-3) if hist_ms > 0 else 0.0,
                "stage_ms_per_step": None}
    # stage breakdown: a separate, untimed pass. A timed section costs two event records in the stream (~3 us each way), so the
    # timed region above times the histogram launches only (h2ot_set_timers mode 1) and the other stages are timed here.
    eng.set_timers(2)
    sp0 = eng.stats()
    snap_p = {f: getattr(sp0, f) for f, _ in sp0._fields_}
    for _ in range(PROF_STEPS):
        gm.train(1)
    gm.flush()
    sp1 = eng.stats()
    roofline["stage_ms_per_step"] = {k: (getattr(sp1, k) - snap_p[k]) / PROF_STEPS for k in (
        "ms_score", "ms_hist", "ms_finalize", "ms_split", "ms_residuals", "ms_gamma", "ms_addtree", "ms_exchange")}
    roofline["stage_ms_per_step"]["note"] = "%d extra steps after the timed region with every stage timed; ms_finalize is part of ms_split when the levels run back to back" % PROF_STEPS
    eng.set_timers(1)
    # secondary ceiling (SURVEY 8d asks for it): the kernel's time goes into shared-memory atomics, ATOMS_PER_ITEM per (row, column)
    # item with unit weights; random-bin ATOMS.ADD runs at 12.4 lanes/clk/SM = 3490 G/s on this part (conflict free: 31 lanes/clk/SM;
    # tools/ubench_atomics2.cu "clean" lines, profiles/r01_ubench_atomics2_b200.txt)
    atoms = ATOMS_PER_ITEM * d["hist_rows_cols"] / (hist_ms * 1e-3) / 1e9 if hist_ms > 0 else 0.0
    roofline["smem_atomics"] = {"achieved": atoms, "peak": 3490.0, "unit": "Gatom/s", "frac": atoms / 3490.0, "atomics_per_item": ATOMS_PER_ITEM,
                                "peak_source": "measured, tools/ubench_atomics2.cu (random-bin pattern) on B200"}
    levels_per_step = (t1s.levels - t0s.levels) / args.steps
    roofline["host_ms_per_step"] = {"inside_level_calls": 1e3 * (t1s.level_s - t0s.level_s) / args.steps,
                                    "whole_step_wall": 1e3 * (t1s.total_s - t0s.total_s) / args.steps}
    # ---------------- parity of the sharded build: every rank's trees must be the same bytes, and the same bytes as a
    # single-GPU build of the whole frame (rank 0 builds it after the timed region; GBMTest.java:1085-1143 is the protocol)
    parity = None
    mine = tree_bytes(gm, K)
    my_preds = [gm.predictions(k) for k in range(K)] if wl["cls"] == "GBM" else None
    gm.close()
    if world > 1:
        digests = [None] * world
        dist.all_gather_object(digests, hashlib.sha256(mine).hexdigest())
        same = all(x == digests[0] for x in digests)
        single = None
        if rank == 0:
            fcols, fy, _ = wl["frame"](N, 0, N, False)
            g1 = new_gbm()
            g1.load(fcols, fy, categorical=wl["cats"])
            g1.train()
            single = tree_bytes(g1, K) == mine
            if my_preds is not None:
                for k in range(K):
                    single = single and bool(np.array_equal(g1.predictions(k)[r0:r1], my_preds[k]))
            g1.close()
            del fcols, fy
            parity = {"trees_identical_across_ranks": bool(same), "trees_identical_to_1gpu": bool(single), "trees": args.steps + args.warmup + PROF_STEPS,
                      "sha256": digests[0][:16]}
        barrier()
    else:
        parity = {"trees_identical_across_ranks": True, "trees_identical_to_1gpu": True, "trees": args.steps + args.warmup + PROF_STEPS,
                  "sha256": hashlib.sha256(mine).hexdigest()[:16]}

    # ---------------- end-to-end arm: whole job from pinned host buffers through the public API
    barrier()
    g2 = new_gbm(ntrees=args.steps)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pred_hold = torch.empty(len(y), dtype=torch.float64).pin_memory()  # results land in pinned memory, like the inputs leave it
    f0.record()
    g2.load(cols, y, categorical=wl["cats"], row_offset=r0, nrows_global=N, comm=comm)
    g2.train(args.steps)
    preds = g2.predictions(out=pred_hold.numpy())
    ntree_nodes = sum(len(g2.tree(t, k)[0]) for t in range(g2.ntrees) for k in range(K))
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
           "what": "load (frame upload from pinned host + rollups + encode) + init + %d iterations + read-back of the predictions (into pinned memory) and "
                   "the trees (%d nodes)" % (args.steps, ntree_nodes)}
    assert np.isfinite(preds).all()
    g2.close()

    line = None
    if rank == 0:
        bpe = sum(1 for _ in cols) and 4
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "sec_per_tree": ms * 1e-3 / args.steps / K, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": wl["text"] % (N, wl["ncols"]), "name": args.config,
                       "rows": N, "cols": wl["ncols"], "levels_per_step": levels_per_step, "parallelism": "rows sharded over %d GPU(s)" % world,
                       "l2": "inputs larger than L2 (frame %.2f GB resident per GPU)" % ((r1 - r0) * wl["ncols"] * bpe / 1e9) if (r1 - r0) * wl["ncols"] * bpe > 126e6
                             else "frame %.0f MB per GPU fits L2: a step still streams it once per level and rewrites nids, residuals and predictions in between" % ((r1 - r0) * wl["ncols"] * bpe / 1e6),
                       "arithmetic": "bins in f64 (reference formula), sums of w, w*y, w*y*y as int64 fixed point, float-cast like reducePrecision",
                       **{k: v for k, v in wl["model"].items() if k in ("max_depth", "nbins", "nbins_top_level", "mtries", "sample_rate")}},
            "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "clocks": clocks, "parity": parity,
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            sample = min(N, args.ref_rows) if args.ref_rows else N
            v, spt, trees, wall = cpu_baseline(wl, sample, args.cpu_budget_s, threads)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sec_per_tree_on_sample": spt / K,
                                    "sample": "%s rows of the frame, %d iterations, %.1f s wall; C++ restatement of the reference CPU "
                                              "path (oracle), not the JVM (no JDK on this box)" % ("all %d" % sample if sample == N else "first %d" % sample, trees, wall)}
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
    ap.add_argument("--config", default="c2", choices=sorted(WORKLOADS), help="BASELINE configuration (default c2 = configs[1])")
    ap.add_argument("--rows", type=int, default=0, help="global rows (default: the configuration's own)")
    ap.add_argument("--ref-rows", type=int, default=0, help="rows of the CPU arm / cpu_baseline (default 0: all rows, the same config)")
    ap.add_argument("--cpu-budget-s", type=float, default=20.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
