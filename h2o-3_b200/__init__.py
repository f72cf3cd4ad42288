"""h2o-3_b200 — Python (ctypes) binding of libh2otree_b200.so, the B200-native engine behind H2O-3's
hex.tree ScoreBuildHistogram2 seam.

Import with ``importlib.import_module("h2o-3_b200")`` (the directory name is not a Python identifier).

Two layers, both thin wrappers over the C ABI declared in include/h2ot.h and include/h2ot_gbm.h:

* ``Engine``  — the per-level drop-in (h2ot_* entry points): what a Java ``ScoreBuildHistogramGPU`` binds.
* ``GBM``     — the host-side model builder (h2ot_gbm_* entry points); parameter names are those of
  ``H2OGradientBoostingEstimator`` / GBMV3 (ntrees, max_depth, nbins, nbins_cats, min_rows, sample_rate,
  col_sample_rate, learn_rate, distribution, seed ...).

There is no CPU fallback: loading fails loudly when the shared library is missing, and engine creation
fails when no CUDA device is present.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libh2otree_b200.so")

OK = 0
FAMILY = {"gaussian": 0, "bernoulli": 1, "multinomial": 2, "poisson": 3, "gamma": 4, "tweedie": 5, "laplace": 6, "quantile": 7}
NA_NONE, NA_VS_REST, NA_LEFT, NA_RIGHT, LEFT, RIGHT = 0, 1, 2, 3, 4, 5
DECIDED_ROW, OUT_OF_BAG = -1, -2
STORE_F32, STORE_F64, STORE_U8, STORE_U16 = 0, 1, 2, 3
DESC_ACTIVE, DESC_ISINT, DESC_ISCAT, DESC_INTOPT, DESC_CHECKFLOAT, DESC_INITNA, DESC_SPLITPTS, DESC_ABSOLUTE = 1, 2, 4, 8, 16, 32, 64, 128
MAX_BITSET_BYTES = 512
COMM_ID_BYTES = 128


class H2OTError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("h2ot error %d: %s" % (code, msg))
        self.code = code


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("ncols", C.c_int32), ("nrows", C.c_int64), ("row_offset", C.c_int64),
                ("nrows_global", C.c_int64), ("nclass_trees", C.c_int32), ("nclass", C.c_int32), ("family", C.c_int32),
                ("has_weights", C.c_int32), ("has_offset", C.c_int32), ("reserved", C.c_int32), ("tweedie_power", C.c_double)]


class HistDesc(C.Structure):
    _fields_ = [("min", C.c_double), ("maxEx", C.c_double), ("step", C.c_double), ("nbin", C.c_int32),
                ("order", C.c_int16), ("flags", C.c_uint8), ("pad", C.c_uint8),
                ("splitpts_off", C.c_int32), ("splitpts_n", C.c_int32)]


class DecidedNode(C.Structure):
    _fields_ = [("nid", C.c_int32), ("pid", C.c_int32), ("col", C.c_int32), ("equal", C.c_int32),
                ("nasplit", C.c_int32), ("splat", C.c_float), ("child", C.c_int32 * 2), ("bs_off", C.c_int32),
                ("bs_nbits", C.c_int32), ("bs_bitoff", C.c_int32), ("pad", C.c_int32)]


class Split(C.Structure):
    _fields_ = [("col", C.c_int32), ("bin", C.c_int32), ("nasplit", C.c_int32), ("equal", C.c_int32),
                ("splat", C.c_float), ("bs_nbits", C.c_int32), ("bs_bitoff", C.c_int32), ("bs_nbytes", C.c_int32),
                ("se", C.c_double), ("se0", C.c_double), ("se1", C.c_double), ("n0", C.c_double), ("n1", C.c_double),
                ("p0", C.c_double), ("p1", C.c_double), ("tree_p0", C.c_double), ("tree_p1", C.c_double),
                ("bs", C.c_uint8 * MAX_BITSET_BYTES)]


class ColumnInfo(C.Structure):
    _fields_ = [("min", C.c_double), ("max", C.c_double), ("nacnt", C.c_int64), ("is_int", C.c_int32),
                ("storage", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("ms_score", C.c_double), ("ms_hist", C.c_double), ("ms_finalize", C.c_double),
                ("ms_split", C.c_double), ("ms_residuals", C.c_double), ("ms_gamma", C.c_double),
                ("ms_addtree", C.c_double), ("ms_sample", C.c_double), ("ms_ingest", C.c_double),
                ("launches", C.c_int64), ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64),
                ("hist_rows_cols", C.c_int64), ("hist_algo_bytes", C.c_int64), ("hist_launches", C.c_int64),
                ("sparse_launches", C.c_int64), ("ms_exchange", C.c_double)]


class GbmParams(C.Structure):
    _fields_ = [("ntrees", C.c_int32), ("max_depth", C.c_int32), ("nbins", C.c_int32), ("nbins_cats", C.c_int32),
                ("nbins_top_level", C.c_int32), ("family", C.c_int32), ("nclass", C.c_int32), ("algo", C.c_int32),
                ("mtries", C.c_int32), ("histogram_type", C.c_int32),
                ("min_rows", C.c_double), ("min_split_improvement", C.c_double), ("sample_rate", C.c_double),
                ("col_sample_rate", C.c_double), ("col_sample_rate_per_tree", C.c_double),
                ("col_sample_rate_change_per_level", C.c_double), ("learn_rate", C.c_double),
                ("learn_rate_annealing", C.c_double), ("max_abs_leafnode_pred", C.c_double), ("seed", C.c_int64),
                ("tweedie_power", C.c_double)]


class TreeNode(C.Structure):
    _fields_ = [("kind", C.c_int32), ("pid", C.c_int32), ("nid", C.c_int32), ("nid0", C.c_int32), ("nid1", C.c_int32),
                ("splat", C.c_float), ("pred", C.c_float), ("has_split", C.c_int32), ("col", C.c_int32),
                ("bin", C.c_int32), ("nasplit", C.c_int32), ("equal", C.c_int32),
                ("se", C.c_double), ("se0", C.c_double), ("se1", C.c_double), ("n0", C.c_double), ("n1", C.c_double),
                ("p0", C.c_double), ("p1", C.c_double), ("tree_p0", C.c_double), ("tree_p1", C.c_double),
                ("bs_nbits", C.c_int32), ("bs_bitoff", C.c_int32), ("bs_nbytes", C.c_int32),
                ("bs", C.c_uint8 * MAX_BITSET_BYTES)]


class GbmTiming(C.Structure):
    _fields_ = [("total_s", C.c_double), ("level_s", C.c_double), ("hist_rows_cols", C.c_double), ("levels", C.c_int64)]


# every symbol declared in include/h2ot.h and include/h2ot_gbm.h (tests check that the library exports all of them)
ENGINE_SYMBOLS = [
    "h2ot_leaf_quantiles", "h2ot_tree_finish_enqueue", "h2ot_tree_levels_end2", "h2ot_compute_residuals_async", "h2ot_residual_bounds_wait",
    "h2ot_version", "h2ot_last_error", "h2ot_device_count", "h2ot_create", "h2ot_destroy", "h2ot_load_column",
    "h2ot_load_response", "h2ot_load_weights", "h2ot_load_offset", "h2ot_finalize_frame", "h2ot_column_info_get",
    "h2ot_init_model", "h2ot_initial_stats", "h2ot_set_predictions", "h2ot_get_predictions", "h2ot_get_work", "h2ot_set_work",
    "h2ot_get_nids", "h2ot_set_nids", "h2ot_compute_residuals", "h2ot_get_gamma_bounds", "h2ot_set_gamma_bounds", "h2ot_begin_tree", "h2ot_build_level",
    "h2ot_level_buffers", "h2ot_allreduce_level", "h2ot_finish_level", "h2ot_find_splits", "h2ot_level_step", "h2ot_set_tree",
    "h2ot_gamma_pass", "h2ot_add_tree", "h2ot_add_tree_scaled", "h2ot_set_sample_rate_per_class", "h2ot_set_lazy_sync", "h2ot_set_timers", "h2ot_score_metrics", "h2ot_drf_init_model", "h2ot_sample", "h2ot_drf_collect_preds", "h2ot_get_oobt",
    "h2ot_comm_unique_id", "h2ot_comm_init", "h2ot_allreduce_host_f64", "h2ot_reserve_exchange",
    "h2ot_get_weight_bound", "h2ot_set_weight_bound", "h2ot_debug_bin_codes", "h2ot_stats_get", "h2ot_stats_reset",
    "h2ot_set_hist_mode", "h2ot_auc_builders", "h2ot_hist_column_noalloc", "h2ot_load_chunk", "h2ot_debug_get_column", "h2ot_set_split_points", "h2ot_score_tree", "h2ot_level_step_begin", "h2ot_level_step_end", "h2ot_tree_levels_begin", "h2ot_tree_levels_end",
]
GBM_SYMBOLS = [
    "h2ot_gbm_flush", "h2ot_gbm_predict", "h2ot_gbm_create", "h2ot_gbm_destroy", "h2ot_gbm_set_column", "h2ot_gbm_load_chunk", "h2ot_gbm_chunks_done", "h2ot_gbm_set_response", "h2ot_gbm_set_weights",
    "h2ot_gbm_set_offset", "h2ot_gbm_comm_init", "h2ot_gbm_init", "h2ot_gbm_train", "h2ot_gbm_capture_levels",
    "h2ot_gbm_init_f", "h2ot_gbm_ntrees", "h2ot_gbm_tree_len", "h2ot_gbm_tree_leaves_start", "h2ot_gbm_tree_nodes",
    "h2ot_gbm_tree_compress", "h2ot_gbm_predictions", "h2ot_gbm_oobt", "h2ot_gbm_score", "h2ot_gbm_auc", "h2ot_gbm_set_pred_noise_bandwidth", "h2ot_gbm_set_sample_rate_per_class", "h2ot_gbm_varimp", "h2ot_gbm_timing_get", "h2ot_gbm_engine", "h2ot_gbm_ncaptures",
    "h2ot_gbm_capture_info", "h2ot_gbm_capture_desc", "h2ot_gbm_capture_vals", "h2ot_gbm_capture_nids",
    "h2ot_hostcheck_hist", "h2ot_hostcheck_find_max_ex", "h2ot_hostcheck_pcg_next_long", "h2ot_hostcheck_pcg_bounded",
    "h2ot_hostcheck_score_cols",
]

_lib = None


def load():
    """Loads libh2otree_b200.so. Raises if it was not built (``python h2o-3_b200/build.py``): no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError("libh2otree_b200.so is missing (%s); build it with `python h2o-3_b200/build.py` — "
                          "there is no CPU or PyTorch fallback" % SO_PATH)
    L = C.CDLL(SO_PATH)
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    L.h2ot_last_error.restype = C.c_char_p
    L.h2ot_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.h2ot_destroy.argtypes = [vp]
    L.h2ot_device_count.argtypes = [C.POINTER(i32)]
    for n in ("h2ot_load_response", "h2ot_load_weights", "h2ot_load_offset"):
        getattr(L, n).argtypes = [vp, i64, i64, vp]
    L.h2ot_load_column.argtypes = [vp, i32, i64, i64, vp]
    L.h2ot_load_chunk.argtypes = [vp, i32, i64, i64, i32, vp, i64]
    L.h2ot_set_split_points.argtypes = [vp, i32, vp, i64]
    L.h2ot_score_tree.argtypes = [vp, i32, i32, i32, vp]
    L.h2ot_gbm_predict.argtypes = [vp, i64, vp, vp]
    L.h2ot_debug_get_column.argtypes = [vp, i32, i64, i64, vp]
    L.h2ot_gbm_load_chunk.argtypes = [vp, i32, i64, i64, i32, vp, i64]
    L.h2ot_gbm_chunks_done.argtypes = [vp, i32, i32, i32]
    L.h2ot_finalize_frame.argtypes = [vp]
    L.h2ot_column_info_get.argtypes = [vp, i32, C.POINTER(ColumnInfo)]
    L.h2ot_init_model.argtypes = [vp, vp, C.POINTER(i64)]
    L.h2ot_initial_stats.argtypes = [vp, C.POINTER(dbl), C.POINTER(dbl), vp, C.POINTER(dbl)]
    for n in ("h2ot_set_predictions", "h2ot_get_predictions", "h2ot_get_work", "h2ot_get_nids", "h2ot_set_nids"):
        getattr(L, n).argtypes = [vp, i32, vp]
    L.h2ot_set_work.argtypes = [vp, i32, vp, dbl]
    L.h2ot_compute_residuals.argtypes = [vp, vp]
    L.h2ot_get_gamma_bounds.argtypes = [vp, vp]
    L.h2ot_set_gamma_bounds.argtypes = [vp, vp]
    L.h2ot_begin_tree.argtypes = [vp, i32, i64, dbl, dbl]
    L.h2ot_build_level.argtypes = [vp, i32, i32, i32, i32, vp, vp, i64, vp]
    L.h2ot_level_buffers.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(i64), C.POINTER(vp), C.POINTER(i64)]
    L.h2ot_allreduce_level.argtypes = [vp, i32]
    L.h2ot_finish_level.argtypes = [vp, i32, vp, i64, vp, vp, vp, vp, vp]
    L.h2ot_find_splits.argtypes = [vp, i32, dbl, dbl, vp]
    L.h2ot_level_step.argtypes = [vp, i32, i32, i32, i32, vp, vp, i64, vp, dbl, dbl, vp, vp, vp, vp, vp]
    L.h2ot_level_step_begin.argtypes = [vp, i32, i32, i32, i32, vp, vp, i64, vp, dbl, dbl]
    L.h2ot_level_step_end.argtypes = [vp, i32, vp, vp, vp, vp, vp]
    L.h2ot_tree_levels_begin.argtypes = [vp, i32, i32, vp, dbl, dbl, i32, i32]
    L.h2ot_tree_levels_end.argtypes = [vp, i32, vp, vp]
    L.h2ot_set_tree.argtypes = [vp, i32, i32, vp, vp, i64]
    L.h2ot_gamma_pass.argtypes = [vp, i32, i32, i32, vp, vp]
    L.h2ot_add_tree.argtypes = [vp, i32, i32, vp]
    L.h2ot_add_tree_scaled.argtypes = [vp, i32, i32, vp, vp]
    L.h2ot_set_sample_rate_per_class.argtypes = [vp, vp, i32]
    L.h2ot_set_lazy_sync.argtypes = [vp, i32]
    L.h2ot_set_timers.argtypes = [vp, i32]
    L.h2ot_gbm_set_pred_noise_bandwidth.argtypes = [vp, dbl]
    L.h2ot_gbm_set_sample_rate_per_class.argtypes = [vp, vp, i32]
    L.h2ot_drf_init_model.argtypes = [vp]
    L.h2ot_sample.argtypes = [vp, i32, i64, dbl]
    L.h2ot_drf_collect_preds.argtypes = [vp, i32, i32, vp, i32, i32]
    L.h2ot_get_oobt.argtypes = [vp, vp]
    L.h2ot_comm_unique_id.argtypes = [vp]
    L.h2ot_comm_init.argtypes = [vp, i32, i32, vp]
    L.h2ot_allreduce_host_f64.argtypes = [vp, vp, i64, i32]
    L.h2ot_reserve_exchange.argtypes = [vp, i64, i64, i64]
    L.h2ot_get_weight_bound.argtypes = [vp, C.POINTER(dbl)]
    L.h2ot_set_weight_bound.argtypes = [vp, dbl]
    L.h2ot_debug_bin_codes.argtypes = [vp, i32, C.POINTER(HistDesc), vp]
    L.h2ot_stats_get.argtypes = [vp, C.POINTER(Stats)]
    L.h2ot_stats_reset.argtypes = [vp]
    L.h2ot_set_hist_mode.argtypes = [vp, i32]
    L.h2ot_hist_column_noalloc.argtypes = [vp, vp, vp, i64, i32, vp, dbl, vp, vp]
    L.h2ot_gbm_create.argtypes = [C.POINTER(GbmParams), i32, i64, i32, i64, i64, i32, i32, C.POINTER(vp)]
    L.h2ot_gbm_destroy.argtypes = [vp]
    L.h2ot_gbm_set_column.argtypes = [vp, i32, vp, i32, i32]
    for n in ("h2ot_gbm_set_response", "h2ot_gbm_set_weights", "h2ot_gbm_set_offset"):
        getattr(L, n).argtypes = [vp, vp]
    L.h2ot_gbm_comm_init.argtypes = [vp, i32, i32, vp]
    L.h2ot_gbm_init.argtypes = [vp]
    L.h2ot_gbm_train.argtypes = [vp, i32]
    L.h2ot_gbm_flush.argtypes = [vp]
    L.h2ot_gbm_capture_levels.argtypes = [vp, i32]
    L.h2ot_gbm_init_f.argtypes = [vp]
    L.h2ot_gbm_init_f.restype = dbl
    L.h2ot_gbm_ntrees.argtypes = [vp]
    L.h2ot_gbm_tree_len.argtypes = [vp, i32, i32]
    L.h2ot_gbm_tree_leaves_start.argtypes = [vp, i32, i32]
    L.h2ot_gbm_tree_nodes.argtypes = [vp, i32, i32, vp]
    L.h2ot_gbm_predictions.argtypes = [vp, i32, vp]
    L.h2ot_gbm_tree_compress.argtypes = [vp, i32, i32, vp, i64, C.POINTER(i64), vp, i64, C.POINTER(i64)]
    L.h2ot_gbm_varimp.argtypes = [vp, vp]
    L.h2ot_gbm_oobt.argtypes = [vp, vp]
    L.h2ot_gbm_score.argtypes = [vp, C.POINTER(dbl), C.POINTER(dbl)]
    L.h2ot_gbm_auc.argtypes = [vp, C.POINTER(dbl), C.POINTER(i32), vp, vp, vp, C.POINTER(i64)]
    L.h2ot_auc_builders.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(i64), vp]
    L.h2ot_score_metrics.argtypes = [vp, vp]
    L.h2ot_gbm_timing_get.argtypes = [vp, C.POINTER(GbmTiming)]
    L.h2ot_gbm_engine.argtypes = [vp]
    L.h2ot_gbm_engine.restype = vp
    L.h2ot_gbm_ncaptures.argtypes = [vp]
    L.h2ot_gbm_capture_info.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
    L.h2ot_gbm_capture_desc.argtypes = [vp, i32, i32, i32, C.POINTER(HistDesc), C.POINTER(dbl), C.POINTER(dbl),
                                        C.POINTER(i32), C.POINTER(i64)]
    L.h2ot_gbm_capture_vals.argtypes = [vp, i32, i32, i32, vp]
    L.h2ot_gbm_capture_nids.argtypes = [vp, i32, vp]
    L.h2ot_hostcheck_hist.argtypes = [i32, i32, i32, dbl, dbl, C.POINTER(i32), C.POINTER(dbl)]
    L.h2ot_hostcheck_find_max_ex.argtypes = [dbl, i32]
    L.h2ot_hostcheck_find_max_ex.restype = dbl
    L.h2ot_hostcheck_pcg_next_long.argtypes = [i64, i32]
    L.h2ot_hostcheck_pcg_next_long.restype = i64
    L.h2ot_hostcheck_pcg_bounded.argtypes = [i64, i32, i32, vp]
    L.h2ot_hostcheck_pcg_bounded.restype = None
    L.h2ot_hostcheck_score_cols.argtypes = [i32, i32, i32, i64, dbl, i32, vp, i32, vp, vp]
    L.h2ot_hostcheck_score_cols.restype = None
    _lib = L
    return L


def _check(rc):
    if rc != OK:
        raise H2OTError(rc, load().h2ot_last_error().decode("utf-8", "replace"))


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def device_count():
    n = C.c_int32(0)
    _check(load().h2ot_device_count(C.byref(n)))
    return n.value


def comm_unique_id():
    buf = np.zeros(COMM_ID_BYTES, dtype=np.uint8)
    _check(load().h2ot_comm_unique_id(_p(buf)))
    return buf


class Engine:
    """One model build on one GPU: resident frame + per-level / per-tree passes (include/h2ot.h)."""

    def __init__(self, nrows, ncols, family="gaussian", nclass=1, nclass_trees=1, has_weights=False,
                 has_offset=False, device=0, row_offset=0, nrows_global=0, _handle=None):
        self.L = load()
        self.nrows, self.ncols, self.K = int(nrows), int(ncols), int(nclass_trees)
        self._owned = _handle is None
        self._last_descs = ()
        if _handle is not None:
            self.h = C.c_void_p(_handle)
            return
        cfg = Config(device=device, ncols=ncols, nrows=nrows, row_offset=row_offset, nrows_global=nrows_global,
                     nclass_trees=nclass_trees, nclass=nclass, family=FAMILY[family] if isinstance(family, str) else family,
                     has_weights=int(has_weights), has_offset=int(has_offset), reserved=0, tweedie_power=1.5)
        self.h = C.c_void_p()
        _check(self.L.h2ot_create(C.byref(cfg), C.byref(self.h)))

    def close(self):
        if self._owned and self.h:
            self.L.h2ot_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- frame
    def load_column(self, col, values, row0=0):
        v = _f64(values)
        _check(self.L.h2ot_load_column(self.h, col, row0, v.size, _p(v)))

    def load_chunk(self, col, codec, mem, row0, n):
        """One H2O chunk (its _mem bytes) decoded on the device into rows [row0, row0 + n) of column col (-1 response,
        -2 weights, -3 offset)."""
        m = np.ascontiguousarray(mem, dtype=np.uint8)
        _check(self.L.h2ot_load_chunk(self.h, col, row0, n, codec, _p(m), m.size))

    def get_column(self, col, row0=0, n=None):
        n = self.nrows - row0 if n is None else n
        out = np.empty(n, dtype=np.float64)
        _check(self.L.h2ot_debug_get_column(self.h, col, row0, n, _p(out)))
        return out

    def load_response(self, y, row0=0):
        v = _f64(y)
        _check(self.L.h2ot_load_response(self.h, row0, v.size, _p(v)))

    def load_weights(self, w, row0=0):
        v = _f64(w)
        _check(self.L.h2ot_load_weights(self.h, row0, v.size, _p(v)))

    def load_offset(self, o, row0=0):
        v = _f64(o)
        _check(self.L.h2ot_load_offset(self.h, row0, v.size, _p(v)))

    def finalize_frame(self):
        _check(self.L.h2ot_finalize_frame(self.h))

    def column_info(self, col):
        ci = ColumnInfo()
        _check(self.L.h2ot_column_info_get(self.h, col, C.byref(ci)))
        return ci

    # ---- model state
    def init_model(self, init_f):
        f = _f64(np.atleast_1d(init_f))
        skipped = C.c_int64(0)
        _check(self.L.h2ot_init_model(self.h, _p(f), C.byref(skipped)))
        return skipped.value

    def get_predictions(self, k=0):
        out = np.empty(self.nrows, dtype=np.float64)
        _check(self.L.h2ot_get_predictions(self.h, k, _p(out)))
        return out

    def set_predictions(self, f, k=0):
        v = _f64(f)
        _check(self.L.h2ot_set_predictions(self.h, k, _p(v)))

    def get_work(self, k=0):
        out = np.empty(self.nrows, dtype=np.float64)
        _check(self.L.h2ot_get_work(self.h, k, _p(out)))
        return out

    def set_work(self, work, abs_bound=None, k=0):
        v = _f64(work)
        if abs_bound is None:
            abs_bound = float(np.nanmax(np.abs(v))) if v.size else 1.0
        _check(self.L.h2ot_set_work(self.h, k, _p(v), abs_bound))

    def get_nids(self, k=0):
        out = np.empty(self.nrows, dtype=np.int32)
        _check(self.L.h2ot_get_nids(self.h, k, _p(out)))
        return out

    def set_nids(self, nids, k=0):
        v = np.ascontiguousarray(nids, dtype=np.int32)
        _check(self.L.h2ot_set_nids(self.h, k, _p(v)))

    # ---- per tree / per level
    def compute_residuals(self):
        out = np.zeros(self.K + 2, dtype=np.float64)
        _check(self.L.h2ot_compute_residuals(self.h, _p(out)))
        return out

    def begin_tree(self, k=0, tree_seed=0, sample_rate=1.0, work_abs_bound=0.0):
        _check(self.L.h2ot_begin_tree(self.h, k, tree_seed, sample_rate, work_abs_bound))

    def build_level(self, descs, leaf_offset=0, decided=None, bitset_pool=None, k=0):
        """descs: ctypes array (HistDesc * (n_leaves*ncols)); decided: ctypes array of DecidedNode or None."""
        n_leaves = len(descs) // self.ncols
        nd = len(decided) if decided is not None else 0
        pool = np.ascontiguousarray(bitset_pool, dtype=np.uint8) if bitset_pool is not None else None
        _check(self.L.h2ot_build_level(self.h, k, leaf_offset, n_leaves, nd,
                                       C.cast(decided, C.c_void_p) if nd else None, _p(pool),
                                       pool.size if pool is not None else 0, C.cast(descs, C.c_void_p)))
        self._last_descs = descs
        return n_leaves

    def allreduce_level(self, k=0):
        _check(self.L.h2ot_allreduce_level(self.h, k))

    def finish_level(self, k=0, want_vals=True):
        descs = self._last_descs
        LC = len(descs)
        nvals = sum((d.nbin + 1) * 3 for d in descs if d.flags & DESC_ACTIVE)
        vals = np.zeros(max(1, nvals), dtype=np.float64) if want_vals else None
        offs = np.zeros(LC, dtype=np.int64)
        min2 = np.zeros(LC, dtype=np.float64)
        max_in = np.zeros(LC, dtype=np.float64)
        wna = np.zeros(LC, dtype=np.float64)
        has = np.zeros(LC, dtype=np.uint8)
        _check(self.L.h2ot_finish_level(self.h, k, _p(vals), vals.size if want_vals else 0, _p(offs), _p(min2),
                                        _p(max_in), _p(wna), _p(has)))
        return dict(vals=vals, offsets=offs, min2=min2, maxIn=max_in, wNA=wna, has_vals=has)

    def find_splits(self, min_rows, min_split_improvement, k=0):
        n_leaves = max(1, len(self._last_descs) // self.ncols)
        out = (Split * n_leaves)()
        _check(self.L.h2ot_find_splits(self.h, k, min_rows, min_split_improvement, C.cast(out, C.c_void_p)))
        return out

    def set_tree(self, nodes, bitset_pool=None, k=0):
        pool = np.ascontiguousarray(bitset_pool, dtype=np.uint8) if bitset_pool is not None else None
        _check(self.L.h2ot_set_tree(self.h, k, len(nodes), C.cast(nodes, C.c_void_p), _p(pool),
                                    pool.size if pool is not None else 0))

    def gamma_pass(self, leaf_start, n_leaf_nodes, k=0):
        num = np.zeros(n_leaf_nodes, dtype=np.float64)
        den = np.zeros(n_leaf_nodes, dtype=np.float64)
        _check(self.L.h2ot_gamma_pass(self.h, k, leaf_start, n_leaf_nodes, _p(num), _p(den)))
        return num, den

    def add_tree(self, node_pred, k=0):
        v = np.ascontiguousarray(node_pred, dtype=np.float32)
        _check(self.L.h2ot_add_tree(self.h, k, v.size, _p(v)))

    # ---- multi-GPU
    def comm_init(self, world_size, rank, uid):
        u = np.ascontiguousarray(uid, dtype=np.uint8)
        _check(self.L.h2ot_comm_init(self.h, world_size, rank, _p(u)))

    # ---- diagnostics
    def bin_codes(self, col, desc):
        out = np.empty(self.nrows, dtype=np.uint16)
        _check(self.L.h2ot_debug_bin_codes(self.h, col, C.byref(desc), _p(out)))
        return out

    def set_split_points(self, points, k=0):
        """Pool of split points for the next build_level / level_step (descriptors refer to it by offset and count)."""
        v = _f64(points)
        _check(self.L.h2ot_set_split_points(self.h, k, _p(v), v.size))

    def stats(self):
        s = Stats()
        _check(self.L.h2ot_stats_get(self.h, C.byref(s)))
        return s

    def stats_reset(self):
        _check(self.L.h2ot_stats_reset(self.h))

    def set_timers(self, mode):
        """2: every stage timed (CUDA events), 1: the histogram launches only, 0: none."""
        _check(self.L.h2ot_set_timers(self.h, int(mode)))

    def set_hist_mode(self, mode):
        _check(self.L.h2ot_set_hist_mode(self.h, mode))


NOALLOC_MAX_LEAVES, NOALLOC_MAX_BINS = 64, 65536


def hist_column_noalloc(d_x, d_work, d_leaf_code, n, descs, work_abs_bound, d_out_vals, stream=0):
    """h2ot_hist_column_noalloc: all d_* are raw device addresses (int) on the current device; descs a ctypes HistDesc array."""
    _check(load().h2ot_hist_column_noalloc(C.c_void_p(d_x), C.c_void_p(d_work), C.c_void_p(d_leaf_code) if d_leaf_code else None,
                                           n, len(descs), C.cast(descs, C.c_void_p), work_abs_bound, C.c_void_p(d_out_vals),
                                           C.c_void_p(stream) if stream else None))


def make_desc(mn, max_ex, nbin, step=None, is_int=0, int_opt=False, check_float=False, init_na=False, order=0, active=True,
              splitpts=None, absolute=False):
    """HistDesc. step defaults to nbin/(maxEx-min) (T/DHistogram.java:253). splitpts=(offset, count) places the node's
    _splitPts in the pool of Engine.set_split_points (histogram_type Random / QuantilesGlobal / UniformRobust)."""
    if step is None:
        step = nbin / (max_ex - mn)
    flags = (DESC_ACTIVE if active else 0) | (DESC_ISINT if is_int > 0 else 0) | (DESC_ISCAT if is_int == 2 else 0) | \
        (DESC_INTOPT if int_opt else 0) | (DESC_CHECKFLOAT if check_float else 0) | (DESC_INITNA if init_na else 0) | \
        (DESC_SPLITPTS if splitpts is not None else 0) | (DESC_ABSOLUTE if absolute else 0)
    off, cnt = splitpts if splitpts is not None else (0, 0)
    return HistDesc(min=mn, maxEx=max_ex, step=step, nbin=nbin, order=order, flags=flags, pad=0, splitpts_off=off, splitpts_n=cnt)


class GBM:
    """Gradient boosting on the B200 engine. Parameter names follow H2OGradientBoostingEstimator
    (h2o-py/h2o/estimators/gbm.py) / GBMV3; defaults follow T/SharedTreeModel.java:51-98 and T/gbm/GBMModel.java:24-41."""

    def __init__(self, ntrees=50, max_depth=5, min_rows=10.0, nbins=20, nbins_cats=1024, nbins_top_level=1024,
                 min_split_improvement=1e-5, sample_rate=1.0, col_sample_rate=1.0, col_sample_rate_per_tree=1.0,
                 col_sample_rate_change_per_level=1.0, learn_rate=0.1, learn_rate_annealing=1.0,
                 max_abs_leafnode_pred=float(np.finfo(np.float64).max), distribution="bernoulli", nclass=2, seed=1234,
                 device=0, algo=0, mtries=-1, tweedie_power=1.5, histogram_type="AUTO", pred_noise_bandwidth=0.0,
                 sample_rate_per_class=None, quantile_alpha=0.5):
        if distribution == "quantile":  # (the parameter struct has one slot for the family's parameter)
            tweedie_power = quantile_alpha
        self.params = GbmParams(ntrees=ntrees, max_depth=max_depth, nbins=nbins, nbins_cats=nbins_cats,
                                nbins_top_level=nbins_top_level, family=FAMILY[distribution], nclass=nclass, algo=algo,
                                mtries=mtries, histogram_type={"AUTO": 0, "UniformAdaptive": 0, "Random": 1}[histogram_type],
                                min_rows=min_rows, min_split_improvement=min_split_improvement, sample_rate=sample_rate,
                                col_sample_rate=col_sample_rate, col_sample_rate_per_tree=col_sample_rate_per_tree,
                                col_sample_rate_change_per_level=col_sample_rate_change_per_level, learn_rate=learn_rate,
                                learn_rate_annealing=learn_rate_annealing, max_abs_leafnode_pred=max_abs_leafnode_pred,
                                seed=seed, tweedie_power=tweedie_power)
        self.device = device
        self.pred_noise_bandwidth = pred_noise_bandwidth
        self.sample_rate_per_class = None if sample_rate_per_class is None else [float(r) for r in sample_rate_per_class]
        self.L = load()
        self.g = None
        self.K = 1 if distribution != "multinomial" else nclass

    def close(self):
        if self.g:
            self.L.h2ot_gbm_destroy(self.g)
        self.g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load(self, columns, y, weights=None, offset=None, categorical=None, row_offset=0, nrows_global=0, comm=None):
        """columns: sequence of 1-D arrays (NaN == NA); categorical: {col: cardinality}.
        comm: (world_size, rank, unique_id) for the row-sharded multi-GPU build."""
        self.close()
        nrows, ncols = len(y), len(columns)
        self.nrows, self.ncols = nrows, ncols
        self.g = C.c_void_p()
        _check(self.L.h2ot_gbm_create(C.byref(self.params), self.device, nrows, ncols, row_offset, nrows_global,
                                      int(weights is not None), int(offset is not None), C.byref(self.g)))
        if comm is not None:
            u = np.ascontiguousarray(comm[2], dtype=np.uint8)
            _check(self.L.h2ot_gbm_comm_init(self.g, comm[0], comm[1], _p(u)))
        categorical = categorical or {}
        for c, col in enumerate(columns):
            v = _f64(col)
            _check(self.L.h2ot_gbm_set_column(self.g, c, _p(v), int(c in categorical), int(categorical.get(c, 0))))
        v = _f64(y)
        _check(self.L.h2ot_gbm_set_response(self.g, _p(v)))
        if weights is not None:
            v = _f64(weights)
            _check(self.L.h2ot_gbm_set_weights(self.g, _p(v)))
        if offset is not None:
            v = _f64(offset)
            _check(self.L.h2ot_gbm_set_offset(self.g, _p(v)))
        self._extra_params()
        _check(self.L.h2ot_gbm_init(self.g))
        return self

    def _extra_params(self):
        if self.pred_noise_bandwidth:
            _check(self.L.h2ot_gbm_set_pred_noise_bandwidth(self.g, float(self.pred_noise_bandwidth)))
        if self.sample_rate_per_class is not None:
            r = np.ascontiguousarray(self.sample_rate_per_class, dtype=np.float64)
            _check(self.L.h2ot_gbm_set_sample_rate_per_class(self.g, _p(r), r.size))

    def load_chunks(self, nrows, columns, y, weights=None, offset=None, categorical=None):
        """Like load(), from H2O chunks: every column (and y / weights / offset) is a list of (codec, mem_bytes, n_rows) in row
        order -- what a Java caller has in Vec.chunkForChunkIdx(i)._mem. Decoded on the device (h2ot_load_chunk)."""
        self.close()
        ncols = len(columns)
        self.nrows, self.ncols = nrows, ncols
        self.g = C.c_void_p()
        _check(self.L.h2ot_gbm_create(C.byref(self.params), self.device, nrows, ncols, 0, 0,
                                      int(weights is not None), int(offset is not None), C.byref(self.g)))
        categorical = categorical or {}

        def push(col, chunks, is_cat=0, card=0):
            row0 = 0
            for codec, mem, n in chunks:
                m = np.ascontiguousarray(mem, dtype=np.uint8)
                _check(self.L.h2ot_gbm_load_chunk(self.g, col, row0, n, codec, _p(m), m.size))
                row0 += n
            if row0 != nrows:
                raise ValueError("chunks of column %d cover %d of %d rows" % (col, row0, nrows))
            _check(self.L.h2ot_gbm_chunks_done(self.g, col, is_cat, card))

        for c, chunks in enumerate(columns):
            push(c, chunks, int(c in categorical), int(categorical.get(c, 0)))
        push(-1, y)
        if weights is not None:
            push(-2, weights)
        if offset is not None:
            push(-3, offset)
        self._extra_params()
        _check(self.L.h2ot_gbm_init(self.g))
        return self

    def capture_levels(self, on=True):
        _check(self.L.h2ot_gbm_capture_levels(self.g, int(on)))

    def train(self, n_iterations=None):
        n = self.params.ntrees if n_iterations is None else n_iterations
        _check(self.L.h2ot_gbm_train(self.g, n))
        return self

    def flush(self):
        """Waits for the trees still in flight on the device (pipelined builds) and mirrors them."""
        _check(self.L.h2ot_gbm_flush(self.g))
        return self

    def engine(self):
        return Engine(self.nrows, self.ncols, nclass_trees=self.K, _handle=self.L.h2ot_gbm_engine(self.g))

    @property
    def init_f(self):
        return self.L.h2ot_gbm_init_f(self.g)

    @property
    def ntrees(self):
        return self.L.h2ot_gbm_ntrees(self.g)

    def tree(self, t, k=0):
        n = self.L.h2ot_gbm_tree_len(self.g, t, k)
        out = (TreeNode * n)()
        _check(self.L.h2ot_gbm_tree_nodes(self.g, t, k, C.cast(out, C.c_void_p)))
        return out, self.L.h2ot_gbm_tree_leaves_start(self.g, t, k)

    def compressed_tree(self, t, k=0):
        """(CompressedTree._bits, auxiliary bytes) of tree (t, k): DTree.compress, the stream SharedTreeMojoModel.scoreTree reads."""
        nb, na = C.c_int64(0), C.c_int64(0)
        _check(self.L.h2ot_gbm_tree_compress(self.g, t, k, None, 0, C.byref(nb), None, 0, C.byref(na)))
        bits = np.zeros(max(1, nb.value), dtype=np.uint8)
        aux = np.zeros(max(1, na.value), dtype=np.uint8)
        _check(self.L.h2ot_gbm_tree_compress(self.g, t, k, _p(bits), bits.size, C.byref(nb), _p(aux), aux.size, C.byref(na)))
        return bits[:nb.value].tobytes(), aux[:na.value].tobytes()

    def predictions(self, k=0, out=None):
        """Tree_k column (link space). out: optional float64[nrows] buffer to fill (e.g. pinned host memory)."""
        if out is None:
            out = np.empty(self.nrows, dtype=np.float64)
        assert out.dtype == np.float64 and out.size == self.nrows and out.flags["C_CONTIGUOUS"]
        _check(self.L.h2ot_gbm_predictions(self.g, k, _p(out)))
        return out

    def predict(self, columns):
        """Link-space scores of NEW rows: array [K, nrows] (GBM: init_f + sum of the trees; DRF: sum of the trees' votes)."""
        cols = [_f64(c) for c in columns]
        n = cols[0].size
        ptrs = (C.c_void_p * len(cols))(*[c.ctypes.data for c in cols])
        out = np.empty((self.K, n), dtype=np.float64)
        _check(self.L.h2ot_gbm_predict(self.g, n, C.cast(ptrs, C.c_void_p), _p(out)))
        return out

    def oobt(self):
        out = np.empty(self.nrows, dtype=np.float64)
        _check(self.L.h2ot_gbm_oobt(self.g, _p(out)))
        return out

    def score(self):
        """Training MSE / logloss computed on the device from the resident Tree_k columns (hex.tree.Score)."""
        mse, ll = C.c_double(), C.c_double()
        _check(self.L.h2ot_gbm_score(self.g, C.byref(mse), C.byref(ll)))
        return dict(mse=mse.value, logloss=ll.value)

    def auc(self):
        """Training AUC as hex.AUC2 computes it: per-chunk AUCBuilders on the device, folded with AUCBuilder.reduce.
        Returns dict(auc, ths, tps, fps, chunk_rows)."""
        a, n, cr = C.c_double(), C.c_int32(), C.c_int64()
        ths, tps, fps = (np.zeros(400) for _ in range(3))
        _check(self.L.h2ot_gbm_auc(self.g, C.byref(a), C.byref(n), _p(ths), _p(tps), _p(fps), C.byref(cr)))
        return dict(auc=a.value, ths=ths[:n.value], tps=tps[:n.value], fps=fps[:n.value], chunk_rows=cr.value)

    def varimp(self):
        out = np.zeros(self.ncols, dtype=np.float32)
        _check(self.L.h2ot_gbm_varimp(self.g, _p(out)))
        return out

    def timing(self):
        t = GbmTiming()
        _check(self.L.h2ot_gbm_timing_get(self.g, C.byref(t)))
        return t

    # captured levels (histogram monitor analogue)
    def ncaptures(self):
        return self.L.h2ot_gbm_ncaptures(self.g)

    def capture_info(self, i):
        a, b, c, d = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        _check(self.L.h2ot_gbm_capture_info(self.g, i, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return dict(tree=a.value, k=b.value, leaf=c.value, nleaves=d.value)

    def capture_desc(self, i, n, c):
        d = HistDesc()
        mn, mx, has, nv = C.c_double(), C.c_double(), C.c_int32(), C.c_int64()
        _check(self.L.h2ot_gbm_capture_desc(self.g, i, n, c, C.byref(d), C.byref(mn), C.byref(mx), C.byref(has), C.byref(nv)))
        return d, mn.value, mx.value, has.value, nv.value

    def capture_vals(self, i, n, c, nvals):
        out = np.empty(nvals, dtype=np.float64)
        _check(self.L.h2ot_gbm_capture_vals(self.g, i, n, c, _p(out)))
        return out

    def capture_nids(self, i):
        out = np.empty(self.nrows, dtype=np.int32)
        _check(self.L.h2ot_gbm_capture_nids(self.g, i, _p(out)))
        return out


class DRF(GBM):
    """Distributed random forest on the same engine. Parameter names and defaults follow H2ORandomForestEstimator /
    DRFModel.DRFParameters (ntrees 50, max_depth 20, min_rows 1, nbins 20, sample_rate 0.632, mtries -1).
    distribution: "bernoulli" == binomial (one tree per iteration), "multinomial", "gaussian" == regression.
    predictions(k) holds the summed out-of-bag votes (Tree_k); divide by oobt() to score (T/drf/DRF.java:343-364)."""

    def __init__(self, ntrees=50, max_depth=20, min_rows=1.0, nbins=20, sample_rate=0.632, mtries=-1, **kw):
        super().__init__(ntrees=ntrees, max_depth=max_depth, min_rows=min_rows, nbins=nbins, sample_rate=sample_rate,
                         algo=1, mtries=mtries, **kw)
