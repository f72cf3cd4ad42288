"""ctypes binding of the CPU oracle (oracle/liboracle.so). TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_SO = os.path.join(ORACLE_DIR, "liboracle.so")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class OrcSplit(C.Structure):
    _fields_ = [("col", C.c_int), ("bin", C.c_int), ("nasplit", C.c_int), ("equal", C.c_int),
                ("se", C.c_double), ("se0", C.c_double), ("se1", C.c_double), ("n0", C.c_double),
                ("n1", C.c_double), ("p0", C.c_double), ("p1", C.c_double), ("tree_p0", C.c_double),
                ("tree_p1", C.c_double), ("bs_nbits", C.c_int), ("bs_bitoff", C.c_int),
                ("bs_nbytes", C.c_int), ("bs", C.c_ubyte * 132)]


class OrcNode(C.Structure):
    _fields_ = [("kind", C.c_int), ("pid", C.c_int), ("nid", C.c_int), ("nid0", C.c_int), ("nid1", C.c_int),
                ("splat", C.c_float), ("pred", C.c_float), ("has_split", C.c_int), ("split", OrcSplit)]


class OrcParams(C.Structure):
    _fields_ = [("ntrees", C.c_int), ("max_depth", C.c_int), ("nbins", C.c_int), ("nbins_cats", C.c_int),
                ("nbins_top_level", C.c_int), ("family", C.c_int), ("nclass", C.c_int), ("nthreads", C.c_int),
                ("min_rows", C.c_double), ("min_split_improvement", C.c_double), ("sample_rate", C.c_double),
                ("col_sample_rate", C.c_double), ("col_sample_rate_per_tree", C.c_double),
                ("col_sample_rate_change_per_level", C.c_double), ("learn_rate", C.c_double),
                ("learn_rate_annealing", C.c_double), ("max_abs_leafnode_pred", C.c_double),
                ("seed", C.c_longlong), ("chunk_rows", C.c_longlong), ("tweedie_power", C.c_double),
                ("histogram_type", C.c_longlong), ("pred_noise_bandwidth", C.c_double), ("n_rate_per_class", C.c_longlong),
                ("sample_rate_per_class", C.c_double * 16)]


class OrcHistDesc(C.Structure):
    _fields_ = [("present", C.c_int), ("has_vals", C.c_int), ("isInt", C.c_int), ("intOpt", C.c_int),
                ("nbin", C.c_int), ("min", C.c_double), ("maxEx", C.c_double), ("step", C.c_double),
                ("min2", C.c_double), ("maxIn", C.c_double)]


GAUSSIAN, BERNOULLI, MULTINOMIAL, POISSON, GAMMA, TWEEDIE, LAPLACE, QUANTILE = 0, 1, 2, 3, 4, 5, 6, 7
HT_UniformAdaptive, HT_Random, HT_QuantilesGlobal, HT_UniformRobust = 0, 1, 2, 3
NA_None, NAvsREST, NALeft, NARight, Left, Right = 0, 1, 2, 3, 4, 5


def default_params(**kw):
    """GBM defaults: T/SharedTreeModel.java:51-98, T/gbm/GBMModel.java:24-41."""
    p = OrcParams(ntrees=50, max_depth=5, nbins=20, nbins_cats=1024, nbins_top_level=1024, family=BERNOULLI,
                  nclass=2, nthreads=1, min_rows=10.0, min_split_improvement=1e-5, sample_rate=1.0,
                  col_sample_rate=1.0, col_sample_rate_per_tree=1.0, col_sample_rate_change_per_level=1.0,
                  learn_rate=0.1, learn_rate_annealing=1.0, max_abs_leafnode_pred=float(np.finfo(np.float64).max),
                  seed=1234, chunk_rows=0, tweedie_power=1.5, histogram_type=0)
    for k, v in kw.items():
        if k == "sample_rate_per_class":
            if v is not None:
                p.n_rate_per_class = len(v)
                for i, r in enumerate(v):
                    p.sample_rate_per_class[i] = r
            continue
        if k == "quantile_alpha":  # (one slot for the family's parameter)
            p.tweedie_power = v
            continue
        setattr(p, k, v)
    return p


def build():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    src_m = max(os.path.getmtime(os.path.join(ORACLE_DIR, f)) for f in ("h2o_tree_oracle.hpp", "h2o_chunk_oracle.hpp", "h2o_compress_oracle.hpp",
                                                                         "h2o_auc_oracle.hpp", "h2o_metrics_oracle.hpp", "oracle_capi.cpp"))
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < src_m:
        build()
    L = C.CDLL(_SO)
    L.orc_hist_new.restype = C.c_void_p
    L.orc_hist_new.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int, C.c_double,
                               C.c_longlong, C.c_int]
    for name in ("orc_hist_free", "orc_hist_init", "orc_hist_reduce_precision"):
        getattr(L, name).argtypes = [C.c_void_p]
        getattr(L, name).restype = None
    L.orc_hist_nbin.argtypes = [C.c_void_p]
    L.orc_hist_nbin.restype = C.c_int
    for name in ("orc_hist_step", "orc_hist_min2", "orc_hist_maxIn"):
        getattr(L, name).argtypes = [C.c_void_p]
        getattr(L, name).restype = C.c_double
    L.orc_hist_bin.argtypes = [C.c_void_p, C.c_double]
    L.orc_hist_bin.restype = C.c_int
    L.orc_hist_bin_many.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p]
    L.orc_hist_bin_many.restype = None
    L.orc_hist_binAt.argtypes = [C.c_void_p, C.c_int]
    L.orc_hist_binAt.restype = C.c_double
    L.orc_hist_nonEmptyBins.argtypes = [C.c_void_p]
    L.orc_hist_nonEmptyBins.restype = C.c_int
    L.orc_hist_update.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    L.orc_hist_update.restype = None
    L.orc_hist_update_int.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    L.orc_hist_update_int.restype = None
    L.orc_hist_vals.argtypes = [C.c_void_p, C.c_void_p]
    L.orc_hist_vals.restype = C.c_int
    L.orc_hist_set_vals.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.orc_hist_set_vals.restype = None
    L.orc_find_maxEx.argtypes = [C.c_double, C.c_int]
    L.orc_find_maxEx.restype = C.c_double
    L.orc_find_split.argtypes = [C.c_void_p, C.c_int, C.c_double, C.POINTER(OrcSplit)]
    L.orc_find_split.restype = C.c_int
    L.orc_splat_numeric.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.orc_splat_numeric.restype = C.c_float
    L.orc_pcg_floats.argtypes = [C.c_longlong, C.c_int, C.c_void_p]
    L.orc_pcg_ints.argtypes = [C.c_longlong, C.c_int, C.c_int, C.c_void_p]
    L.orc_pcg_next_long.argtypes = [C.c_longlong, C.c_int]
    L.orc_pcg_next_long.restype = C.c_longlong
    L.orc_tree_score_cols.argtypes = [C.c_int, C.c_int, C.c_int, C.c_longlong, C.c_double, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                      C.c_void_p]
    L.orc_tree_score_cols.restype = None
    L.orc_sample.argtypes = [C.c_longlong, C.c_double, C.c_longlong, C.c_longlong, C.c_void_p]
    L.orc_java_random_intopt_inputs.argtypes = [C.c_longlong, C.c_int, C.c_void_p, C.c_void_p]
    L.orc_equals_within_one_small_ulp_f.argtypes = [C.c_float, C.c_float]
    L.orc_equals_within_one_small_ulp_f.restype = C.c_int
    L.orc_frame_new.argtypes = [C.c_longlong, C.c_int]
    L.orc_frame_new.restype = C.c_void_p
    L.orc_frame_free.argtypes = [C.c_void_p]
    L.orc_frame_set_col.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int]
    for name in ("orc_frame_set_y", "orc_frame_set_w", "orc_frame_set_off"):
        getattr(L, name).argtypes = [C.c_void_p, C.c_void_p]
    L.orc_gbm_train.argtypes = [C.c_void_p, C.POINTER(OrcParams), C.c_int]
    L.orc_gbm_train.restype = C.c_void_p
    L.orc_drf_train.argtypes = [C.c_void_p, C.POINTER(OrcParams), C.c_int, C.c_int]
    L.orc_drf_train.restype = C.c_void_p
    L.orc_model_oobt.argtypes = [C.c_void_p, C.c_void_p]
    L.orc_model_free.argtypes = [C.c_void_p]
    L.orc_model_init_f.argtypes = [C.c_void_p]
    L.orc_model_init_f.restype = C.c_double
    L.orc_model_ntrees.argtypes = [C.c_void_p]
    L.orc_model_K.argtypes = [C.c_void_p]
    L.orc_model_F.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.orc_model_timing.argtypes = [C.c_void_p, c_double_p, c_double_p, c_double_p]
    L.orc_model_varimp.argtypes = [C.c_void_p, C.c_void_p]
    L.orc_model_tree_len.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.orc_model_tree_leaves_start.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.orc_model_tree_nodes.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    L.orc_model_score_tree.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    L.orc_model_score_tree.restype = C.c_float
    L.orc_model_tree_compress.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    L.orc_model_tree_compress.restype = C.c_longlong
    L.orc_score_compressed.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p]
    L.orc_score_compressed.restype = None
    L.orc_auc2.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_longlong, c_int_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_auc2.restype = C.c_double
    L.orc_model_ncaptures.argtypes = [C.c_void_p]
    L.orc_capture_info.argtypes = [C.c_void_p, C.c_int, c_int_p, c_int_p, c_int_p, c_int_p]
    L.orc_capture_desc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(OrcHistDesc)]
    L.orc_capture_vals.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    L.orc_capture_nids.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    _lib = L
    return L


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Hist:
    """Thin wrapper over oracle DHistogram (T/DHistogram.java)."""

    def __init__(self, nbins, nbins_cats, isInt, mn, maxEx, intOpt=False, initNA=False, msi=0.0, seed=0,
                 checkFloat=False):
        self.L = lib()
        self.h = self.L.orc_hist_new(nbins, nbins_cats, isInt, mn, maxEx, int(intOpt), int(initNA), msi, seed,
                                     int(checkFloat))
        assert self.h, "DHistogram ctor failed"

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_hist_free(self.h)
            self.h = None

    def init(self):
        self.L.orc_hist_init(self.h)
        return self

    def set_type(self, histo_type, pts=None, can_refine=True):
        """HistogramType (HT_*) and, for QuantilesGlobal / UniformRobust, the global / custom split points; before init()."""
        self.L.orc_hist_set_type.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int]
        self.L.orc_hist_set_type.restype = None
        v = None if pts is None else np.ascontiguousarray(pts, dtype=np.float64)
        self.L.orc_hist_set_type(self.h, histo_type, ptr(v), 0 if v is None else len(v), int(can_refine))
        return self

    def split_points(self):
        self.L.orc_hist_split_points.argtypes = [C.c_void_p, C.c_void_p]
        n = self.L.orc_hist_split_points(self.h, None)
        out = np.empty(n, dtype=np.float64)
        self.L.orc_hist_split_points.restype = C.c_int
        self.L.orc_hist_split_points(self.h, ptr(out))
        return out

    @property
    def absolute(self):
        self.L.orc_hist_absolute.argtypes = [C.c_void_p]
        return bool(self.L.orc_hist_absolute(self.h))

    nbin = property(lambda s: s.L.orc_hist_nbin(s.h))
    step = property(lambda s: s.L.orc_hist_step(s.h))
    min2 = property(lambda s: s.L.orc_hist_min2(s.h))
    maxIn = property(lambda s: s.L.orc_hist_maxIn(s.h))

    def bin(self, x):
        return self.L.orc_hist_bin(self.h, float(x))

    def bin_many(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        out = np.empty(len(x), dtype=np.int32)
        self.L.orc_hist_bin_many(self.h, ptr(x), len(x), ptr(out))
        return out

    def binAt(self, b):
        return self.L.orc_hist_binAt(self.h, b)

    def update(self, ws, cs, ys, rows=None):
        cs = np.ascontiguousarray(cs, dtype=np.float64)
        ys = np.ascontiguousarray(ys, dtype=np.float64)
        ws = None if ws is None else np.ascontiguousarray(ws, dtype=np.float64)
        rows = np.arange(len(cs), dtype=np.int32) if rows is None else np.ascontiguousarray(rows, dtype=np.int32)
        self.L.orc_hist_update(self.h, ptr(ws), ptr(cs), ptr(ys), ptr(rows), len(rows), 0)

    def update_int(self, ws, cs, ys, rows=None):
        cs = np.ascontiguousarray(cs, dtype=np.int32)
        ys = np.ascontiguousarray(ys, dtype=np.float64)
        ws = None if ws is None else np.ascontiguousarray(ws, dtype=np.float64)
        rows = np.arange(len(cs), dtype=np.int32) if rows is None else np.ascontiguousarray(rows, dtype=np.int32)
        self.L.orc_hist_update_int(self.h, ptr(ws), ptr(cs), ptr(ys), ptr(rows), len(rows), 0)

    def vals(self):
        n = self.L.orc_hist_vals(self.h, None)
        out = np.empty(n, dtype=np.float64)
        self.L.orc_hist_vals(self.h, ptr(out))
        return out

    def set_vals(self, v):
        v = np.ascontiguousarray(v, dtype=np.float64)
        self.L.orc_hist_set_vals(self.h, ptr(v), len(v))

    def reduce_precision(self):
        self.L.orc_hist_reduce_precision(self.h)

    def non_empty_bins(self):
        return self.L.orc_hist_nonEmptyBins(self.h)

    def find_split(self, col, min_rows):
        s = OrcSplit()
        ok = self.L.orc_find_split(self.h, col, float(min_rows), C.byref(s))
        return s if ok else None

    def splat_numeric(self, bin_, nasplit):
        return self.L.orc_splat_numeric(self.h, bin_, nasplit)


class Frame:
    def __init__(self, X, y, w=None, off=None, cat_cards=None):
        """X: [nrows, ncols] float64 (NaN = NA); cat_cards: {col: cardinality} for categoricals."""
        self.L = lib()
        X = np.asarray(X, dtype=np.float64)
        self.nrows, self.ncols = X.shape
        self.f = self.L.orc_frame_new(self.nrows, self.ncols)
        cat_cards = cat_cards or {}
        for c in range(self.ncols):
            col = np.ascontiguousarray(X[:, c])
            self.L.orc_frame_set_col(self.f, c, ptr(col), int(c in cat_cards), int(cat_cards.get(c, 0)))
        y = np.ascontiguousarray(y, dtype=np.float64)
        self.L.orc_frame_set_y(self.f, ptr(y))
        if w is not None:
            w = np.ascontiguousarray(w, dtype=np.float64)
            self.L.orc_frame_set_w(self.f, ptr(w))
        if off is not None:
            off = np.ascontiguousarray(off, dtype=np.float64)
            self.L.orc_frame_set_off(self.f, ptr(off))

    def __del__(self):
        if getattr(self, "f", None):
            self.L.orc_frame_free(self.f)
            self.f = None


class Model:
    def __init__(self, frame, params, capture=False, drf=False, mtries=-1):
        self.L = lib()
        self.frame = frame
        if drf:
            self.m = self.L.orc_drf_train(frame.f, C.byref(params), int(mtries), int(capture))
        else:
            self.m = self.L.orc_gbm_train(frame.f, C.byref(params), int(capture))
        assert self.m, "oracle training failed"

    def __del__(self):
        if getattr(self, "m", None):
            self.L.orc_model_free(self.m)
            self.m = None

    init_f = property(lambda s: s.L.orc_model_init_f(s.m))
    ntrees = property(lambda s: s.L.orc_model_ntrees(s.m))
    K = property(lambda s: s.L.orc_model_K(s.m))

    def F(self, k=0):
        out = np.empty(self.frame.nrows, dtype=np.float64)
        self.L.orc_model_F(self.m, k, ptr(out))
        return out

    def oobt(self):
        out = np.empty(self.frame.nrows, dtype=np.float64)
        self.L.orc_model_oobt(self.m, ptr(out))
        return out

    def timing(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        self.L.orc_model_timing(self.m, C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

    def varimp(self):
        out = np.empty(self.frame.ncols, dtype=np.float32)
        self.L.orc_model_varimp(self.m, ptr(out))
        return out

    def tree(self, t, k=0):
        n = self.L.orc_model_tree_len(self.m, t, k)
        arr = (OrcNode * n)()
        self.L.orc_model_tree_nodes(self.m, t, k, arr)
        return list(arr)

    def leaves_start(self, t, k=0):
        return self.L.orc_model_tree_leaves_start(self.m, t, k)

    def score_tree(self, t, k, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        return self.L.orc_model_score_tree(self.m, t, k, ptr(x))

    def compressed_tree(self, t, k=0):
        """(bits, aux) of DTree.compress for tree (t, k)."""
        out = []
        for aux in (0, 1):
            n = self.L.orc_model_tree_compress(self.m, t, k, aux, None)
            assert n >= 0, "DecidedNode.compress: size() differs from the bytes written"
            b = np.zeros(max(1, n), dtype=np.uint8)
            self.L.orc_model_tree_compress(self.m, t, k, aux, ptr(b))
            out.append(b[:n].tobytes())
        return tuple(out)

    def ncaptures(self):
        return self.L.orc_model_ncaptures(self.m)

    def capture(self, i):
        t, k, leaf, nl = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self.L.orc_capture_info(self.m, i, C.byref(t), C.byref(k), C.byref(leaf), C.byref(nl))
        ncols = self.frame.ncols
        descs, vals = [], []
        for n in range(nl.value):
            for c in range(ncols):
                d = OrcHistDesc()
                self.L.orc_capture_desc(self.m, i, n, c, ncols, C.byref(d))
                descs.append(d)
                nv = self.L.orc_capture_vals(self.m, i, n, c, ncols, None)
                v = np.empty(nv, dtype=np.float64)
                if nv:
                    self.L.orc_capture_vals(self.m, i, n, c, ncols, ptr(v))
                vals.append(v)
        nids = np.empty(self.frame.nrows, dtype=np.int32)
        self.L.orc_capture_nids(self.m, i, ptr(nids))
        return dict(tree=t.value, k=k.value, leaf=leaf.value, nleaves=nl.value, descs=descs, vals=vals, nids=nids)


def score_compressed(bits, X):
    """SharedTreeMojoModel.scoreTree over CompressedTree bytes for every row of X ([n, ncols] float64, NaN == NA)."""
    X = np.ascontiguousarray(X, dtype=np.float64)
    b = np.frombuffer(bits, dtype=np.uint8).copy()
    out = np.empty(X.shape[0], dtype=np.float32)
    lib().orc_score_compressed(ptr(b), len(bits), ptr(X), X.shape[0], X.shape[1], ptr(out))
    return out


def auc2(pred, act, w=None, chunk_rows=1 << 62):
    """AUC2 with one AUCBuilder per chunk_rows consecutive rows, folded left to right. Returns (auc, ths, tps, fps)."""
    pred = np.ascontiguousarray(pred, dtype=np.float64)
    act = np.ascontiguousarray(act, dtype=np.float64)
    w = None if w is None else np.ascontiguousarray(w, dtype=np.float64)
    n = C.c_int(0)
    ths, tps, fps = (np.zeros(800) for _ in range(3))
    a = lib().orc_auc2(ptr(pred), ptr(act), ptr(w), len(pred), min(int(chunk_rows), 1 << 62), C.byref(n), ptr(ths), ptr(tps), ptr(fps))
    return a, ths[:n.value], tps[:n.value], fps[:n.value]


def auc_builder(nbins, pred, act, w=None):
    """One AUC2.AUCBuilder of nbins bins fed row by row. Returns (ths, tps, fps, sqe) of its bins."""
    pred = np.ascontiguousarray(pred, dtype=np.float64)
    act = np.ascontiguousarray(act, dtype=np.int32)
    w = None if w is None else np.ascontiguousarray(w, dtype=np.float64)
    out = [np.zeros(2 * nbins + 2) for _ in range(4)]
    L = lib()
    L.orc_auc_builder.restype = C.c_int
    L.orc_auc_builder.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong] + [C.c_void_p] * 4
    n = L.orc_auc_builder(int(nbins), ptr(pred), ptr(act), ptr(w), len(pred), *[ptr(o) for o in out])
    return tuple(o[:n] for o in out)


def score_metrics(family, Fk, y, w=None):
    """hex.tree.Score's metric sums over Tree_k columns Fk ([K][n], link space incl. init_f). Returns dict(count, wcount, mse, logloss)."""
    Fk = [np.ascontiguousarray(f, dtype=np.float64) for f in Fk]
    y = np.ascontiguousarray(y, dtype=np.float64)
    w = None if w is None else np.ascontiguousarray(w, dtype=np.float64)
    ptrs = (C.c_void_p * len(Fk))(*[f.ctypes.data for f in Fk])
    out = np.zeros(4)
    L = lib()
    L.orc_score_metrics.restype = None
    L.orc_score_metrics.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p]
    L.orc_score_metrics(int(family), len(Fk), ptrs, ptr(y), ptr(w), len(y), ptr(out))
    return dict(count=out[0], wcount=out[1], mse=out[2] / out[1] if out[1] else float("nan"), logloss=out[3] / out[1] if out[1] else float("nan"))
