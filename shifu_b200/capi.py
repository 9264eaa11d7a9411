"""ctypes binding of include/shifu_b200.h (the C ABI of the B200 tree-training hot path).

This is the binding a Python host would use; the Java host binds the same symbols over JNI.
Nothing here computes: every call goes to libshifu_b200.so, and loading fails loudly if the library is missing.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libshifu_b200.so")

SB_IMPURITY_VARIANCE, SB_IMPURITY_FRIEDMANMSE, SB_IMPURITY_ENTROPY, SB_IMPURITY_GINI = 0, 1, 2, 3
SB_LOSS_SQUARED, SB_LOSS_HALFGRADSQUARED, SB_LOSS_ABSOLUTE, SB_LOSS_LOG = 0, 1, 2, 3
SB_SPLIT_CONTINUOUS, SB_SPLIT_CATEGORICAL = 1, 2
SB_ROOT_INDEX, SB_INVALID_INDEX = 1, -1
SB_MAX_CAT_BYTES = 4096
SB_STATIC_MAX_ROWS = 67108864


class SbConfig(C.Structure):
    _fields_ = [
        ("n_features", C.c_int32),
        ("n_bins", C.POINTER(C.c_int32)),
        ("is_categorical", C.POINTER(C.c_uint8)),
        ("column_num", C.POINTER(C.c_int32)),
        ("bin_boundary", C.POINTER(C.c_double)),
        ("impurity", C.c_int32),
        ("num_classes", C.c_int32),
        ("loss", C.c_int32),
        ("is_gbdt", C.c_int32),
        ("tree_num", C.c_int32),
        ("weights_per_row", C.c_int32),
        ("max_depth", C.c_int32),
        ("min_instances_per_node", C.c_int32),
        ("min_info_gain", C.c_double),
        ("learning_rate", C.c_double),
        ("device", C.c_int32),
        ("flags", C.c_uint32),
    ]


class SbNode(C.Structure):
    _fields_ = [
        ("id", C.c_int32),
        ("left", C.c_int32),
        ("right", C.c_int32),
        ("feature", C.c_int32),
        ("split_type", C.c_int32),
        ("left_mask_offset", C.c_int32),
        ("left_mask_bytes", C.c_int32),
        ("has_predict", C.c_int32),
        ("predict", C.c_double),
        ("threshold", C.c_double),
        ("is_left", C.c_int32),
        ("cat_offset", C.c_int32),
        ("cat_bytes", C.c_int32),
        ("reserved", C.c_int32),
    ]


class SbTree(C.Structure):
    _fields_ = [
        ("tree_id", C.c_int32),
        ("n_nodes", C.c_int32),
        ("nodes", C.POINTER(SbNode)),
        ("mask_pool", C.POINTER(C.c_uint8)),
        ("mask_pool_bytes", C.c_int32),
    ]


class SbTodoNode(C.Structure):
    _fields_ = [
        ("key", C.c_int32),
        ("tree_id", C.c_int32),
        ("node_id", C.c_int32),
        ("n_features", C.c_int32),
        ("features", C.POINTER(C.c_int32)),
    ]


class SbIterIn(C.Structure):
    _fields_ = [
        ("n_trees", C.c_int32),
        ("trees", C.POINTER(SbTree)),
        ("n_todo", C.c_int32),
        ("todo", C.POINTER(SbTodoNode)),
        ("switch_to_next_tree", C.c_int32),
        ("is_first_tree", C.c_int32),
        ("continuous_running_start", C.c_int32),
        ("new_weights", C.POINTER(C.c_float)),
        ("dropout", C.POINTER(C.c_uint8)),
    ]


class SbIterOut(C.Structure):
    _fields_ = [
        ("train_count", C.c_double),
        ("validation_count", C.c_double),
        ("train_error", C.c_double),
        ("validation_error", C.c_double),
        ("stats", C.POINTER(C.c_double)),
        ("stats_len", C.c_int64),
    ]


class SbGain(C.Structure):
    _fields_ = [
        ("found", C.c_int32),
        ("feature", C.c_int32),
        ("split_type", C.c_int32),
        ("split_bin", C.c_int32),
        ("is_left", C.c_int32),
        ("cat_bytes", C.c_int32),
        ("left_class", C.c_int32),
        ("right_class", C.c_int32),
        ("node_class", C.c_int32),
        ("reserved", C.c_int32),
        ("gain", C.c_double),
        ("impurity", C.c_double),
        ("predict", C.c_double),
        ("left_impurity", C.c_double),
        ("right_impurity", C.c_double),
        ("left_predict", C.c_double),
        ("right_predict", C.c_double),
        ("wgt_cnt", C.c_double),
    ]


class SbScoreNode(C.Structure):
    _fields_ = [("left", C.c_int32), ("right", C.c_int32), ("column", C.c_int32), ("split_type", C.c_int32),
                ("threshold", C.c_double), ("is_left", C.c_int32), ("cat_offset", C.c_int32), ("cat_bytes", C.c_int32),
                ("class_value", C.c_int32), ("predict", C.c_double)]


class SbScoreModel(C.Structure):
    _fields_ = [("n_bags", C.c_int32), ("bag_begin", C.POINTER(C.c_int32)), ("n_trees", C.c_int32),
                ("tree_root", C.POINTER(C.c_int32)), ("weights", C.POINTER(C.c_double)), ("n_nodes", C.c_int32),
                ("nodes", C.POINTER(SbScoreNode)), ("cat_pool", C.POINTER(C.c_uint8)), ("cat_pool_bytes", C.c_int32),
                ("is_gbdt", C.c_int32), ("is_classification", C.c_int32), ("convert", C.c_int32),
                ("n_columns", C.c_int32), ("categorical_size", C.POINTER(C.c_int32))]


SCORE_NODE_DTYPE = np.dtype([("left", "<i4"), ("right", "<i4"), ("column", "<i4"), ("split_type", "<i4"),
                             ("threshold", "<f8"), ("is_left", "<i4"), ("cat_offset", "<i4"), ("cat_bytes", "<i4"),
                             ("class_value", "<i4"), ("predict", "<f8")], align=True)
assert SCORE_NODE_DTYPE.itemsize == C.sizeof(SbScoreNode)
CONVERT = {"RAW": 0, "OLD_SIGMOID": 1, "SIGMOID": 2, "CUTOFF": 3}


class ScoreModel:
    """Flattened IndependentTreeModel (sb_score_model) plus the numpy arrays it points at.

    bags: list of bags, each a list of (root, weight); a node is a dict with "split" (None or {"column", "feature_type",
    "threshold" | "categories" + "is_left"}), "left", "right" (dicts or None) and "predict" ((value, class) or None);
    split["column"] is the position in a data row.  categorical_size: per data column, 0 for numeric."""

    def __init__(self, bags, n_columns, categorical_size=None, is_gbdt=True, is_classification=False, convert="RAW"):
        recs, pool, roots, weights, bag_begin = [], bytearray(), [], [], [0]

        def add(n):
            i = len(recs)
            recs.append(None)
            sp = n.get("split")
            col, st, thr, isl, co, cb = -1, 0, 0.0, 0, 0, 0
            if sp is not None:
                col, st = sp["column"], sp["feature_type"]
                if st == SB_SPLIT_CATEGORICAL:
                    isl = 1 if sp["is_left"] else 0
                    cats = sp.get("categories") or b""
                    co, cb = len(pool), len(cats)
                    pool.extend(cats)
                else:
                    thr = sp["threshold"]
            left = add(n["left"]) if n.get("left") is not None else -1
            right = add(n["right"]) if n.get("right") is not None else -1
            pr = n.get("predict") or (0.0, 0)
            recs[i] = (left, right, col, st, thr, isl, co, cb, int(pr[1]), float(pr[0]))
            return i

        for bag in bags:
            for root, w in bag:
                roots.append(add(root))
                weights.append(w)
            bag_begin.append(len(roots))
        self.nodes = np.array(recs, dtype=SCORE_NODE_DTYPE)
        self.pool = np.frombuffer(bytes(pool) + b"\0", dtype=np.uint8).copy()
        self.roots = np.array(roots, dtype=np.int32)
        self.weights = np.array(weights, dtype=np.float64)
        self.bag_begin = np.array(bag_begin, dtype=np.int32)
        self.cat_size = np.zeros(n_columns, dtype=np.int32) if categorical_size is None else np.asarray(categorical_size, dtype=np.int32)
        assert len(self.cat_size) == n_columns
        m = SbScoreModel()
        m.n_bags, m.bag_begin = len(bags), self.bag_begin.ctypes.data_as(C.POINTER(C.c_int32))
        m.n_trees, m.tree_root = len(roots), self.roots.ctypes.data_as(C.POINTER(C.c_int32))
        m.weights = self.weights.ctypes.data_as(C.POINTER(C.c_double))
        m.n_nodes, m.nodes = len(recs), C.cast(self.nodes.ctypes.data, C.POINTER(SbScoreNode))
        m.cat_pool, m.cat_pool_bytes = self.pool.ctypes.data_as(C.POINTER(C.c_uint8)), len(self.pool) - 1
        m.is_gbdt, m.is_classification, m.convert = int(bool(is_gbdt)), int(bool(is_classification)), CONVERT[convert]
        m.n_columns, m.categorical_size = n_columns, self.cat_size.ctypes.data_as(C.POINTER(C.c_int32))
        self.struct = m
        self.n_trees, self.n_columns, self.is_classification = len(roots), n_columns, bool(is_classification)

    def out_per_row(self):
        return self.n_trees if self.is_classification else 1


class Scorer:
    """sb_scorer: the model resident on one GPU; score(data) -> IndependentTreeModel.compute for every row."""

    def __init__(self, model: ScoreModel, device=0):
        self.lib = load_library()
        self.model = model
        self.h = C.c_void_p()
        check(self.lib, self.lib.sb_scorer_create(C.byref(model.struct), device, C.byref(self.h)))

    def close(self):
        if self.h:
            self.lib.sb_scorer_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def score(self, data):
        data = np.ascontiguousarray(data, dtype=np.float64)
        assert data.ndim == 2 and data.shape[1] == self.model.n_columns
        out = np.zeros((data.shape[0], self.model.out_per_row()), dtype=np.float64)
        check(self.lib, self.lib.sb_scorer_score(self.h, data.shape[0], data.ctypes.data, 0, out.ctypes.data))
        return out[:, 0] if not self.model.is_classification else out

    def last_kernel_ms(self):
        ms = C.c_float()
        check(self.lib, self.lib.sb_scorer_last_kernel_ms(self.h, C.byref(ms)))
        return ms.value


class SbhParams(C.Structure):
    _fields_ = [
        ("tree_num", C.c_int32),
        ("max_depth", C.c_int32),
        ("is_gbdt", C.c_int32),
        ("feature_subset_strategy", C.c_int32),
        ("feature_subset_rate", C.c_double),
        ("learning_rate", C.c_double),
        ("feature_sampling_seed", C.c_int64),
        ("max_leaves", C.c_int32),
        ("max_batch_split_size", C.c_int32),
        ("max_stats_memory", C.c_int64),
        ("worker_number", C.c_int32),
        ("reserved", C.c_int32),
    ]


SBH_FSS = {"ALL": 0, "HALF": 1, "ONETHIRD": 2, "TWOTHIRDS": 3, "AUTO": 4, "SQRT": 5, "LOG2": 6, None: 7}


# every symbol include/shifu_b200.h declares (tests check that the library exports all of them)
EXPORTS = [
    "sb_create", "sb_destroy", "sb_last_error", "sb_version", "sb_load_rows", "sb_load_rows_device", "sb_read_rows",
    "sb_stats_len", "sb_compute", "sb_compute_begin", "sb_scale_buffer", "sb_compute_stats", "sb_raw_buffer",
    "sb_acc_buffer", "sb_compute_finish", "sb_compute_finish_split", "sb_set_stream", "sb_best_splits", "sb_time_stats_pass", "sb_counters",
    "sb_load_synthetic", "sb_read_bins", "sb_read_labels", "sb_set_profiling", "sb_hist_time",
    "sb_error_sums_static", "sb_pass_info", "sb_set_weights", "sb_sample_weights", "sb_bin_values",
    "sb_dist_unique_id", "sb_dist_init", "sbh_grow_tree",
    "sb_sample_weights_stratified", "sb_sample_poisson", "sb_scorer_create", "sb_scorer_destroy", "sb_scorer_score", "sb_scorer_last_kernel_ms",
    "sbh_master_create", "sbh_master_destroy", "sbh_master_initial", "sbh_master_do_compute", "sbh_master_tree_count",
    "sbh_master_halted", "sbh_master_export_tree",
    "sb_rows_streamed", "sb_msg_pack", "sb_msg_unpack", "sb_msg_iter_in", "sb_msg_free",
    "sb_category_map_create", "sb_category_map_lookup", "sb_category_map_destroy",
]

_lib = None


def load_library(path=None):
    """dlopen libshifu_b200.so; raises if it is absent (there is no CPU fallback)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(
            "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` or ./build.sh "
            "(shifu_b200 has no CPU fallback)" % p)
    lib = C.CDLL(p)
    H = C.c_void_p
    lib.sb_create.argtypes = [C.POINTER(SbConfig), C.POINTER(H)]
    lib.sb_destroy.argtypes = [H]
    lib.sb_destroy.restype = None
    lib.sb_last_error.restype = C.c_char_p
    lib.sb_version.restype = C.c_char_p
    lib.sb_load_rows.argtypes = [H, C.c_int32, C.c_int64, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p, C.c_void_p]
    lib.sb_load_rows_device.argtypes = [H, C.c_int32, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.sb_read_rows.argtypes = [H, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.sb_stats_len.argtypes = [H, C.POINTER(SbIterIn)]
    lib.sb_stats_len.restype = C.c_int64
    lib.sb_compute.argtypes = [H, C.POINTER(SbIterIn), C.POINTER(SbIterOut)]
    lib.sb_compute_begin.argtypes = [H, C.POINTER(SbIterIn)]
    lib.sb_scale_buffer.argtypes = [H, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    lib.sb_compute_stats.argtypes = [H]
    lib.sb_raw_buffer.argtypes = [H, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    lib.sb_acc_buffer.argtypes = [H, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    lib.sb_compute_finish.argtypes = [H, C.POINTER(SbIterOut)]
    lib.sb_set_stream.argtypes = [H, C.c_void_p]
    lib.sb_best_splits.argtypes = [H, C.POINTER(SbGain), C.c_void_p]
    lib.sb_compute_finish_split.argtypes = [H, C.POINTER(SbIterOut), C.POINTER(SbGain), C.c_void_p]
    lib.sb_time_stats_pass.argtypes = [H, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.sb_counters.argtypes = [H, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.sb_load_synthetic.argtypes = [H, C.c_int32, C.c_int64, C.c_int64, C.c_uint64, C.c_int32, C.c_int32]
    lib.sb_read_bins.argtypes = [H, C.c_int32, C.c_int32, C.c_int64, C.c_int64, C.c_void_p]
    lib.sb_read_labels.argtypes = [H, C.c_int32, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
    lib.sb_set_profiling.argtypes = [H, C.c_int32]
    lib.sb_hist_time.argtypes = [H, C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.sb_set_weights.argtypes = [H, C.c_int32, C.c_void_p]
    lib.sb_sample_weights.argtypes = [H, C.c_int32, C.c_int64, C.c_double, C.c_int64]
    lib.sb_bin_values.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]
    lib.sb_pass_info.argtypes = [H, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.sbh_master_create.argtypes = [C.POINTER(SbConfig), C.POINTER(SbhParams), C.POINTER(H)]
    lib.sbh_master_destroy.argtypes = [H]
    lib.sbh_master_destroy.restype = None
    lib.sbh_master_initial.argtypes = [H]
    lib.sbh_master_initial.restype = C.POINTER(SbIterIn)
    lib.sbh_master_do_compute.argtypes = [H, C.POINTER(SbGain), C.c_void_p]
    lib.sb_dist_unique_id.argtypes = [C.c_void_p]
    lib.sb_sample_weights_stratified.argtypes = [H, C.c_int32, C.c_int64, C.c_double, C.c_int32]
    lib.sb_sample_poisson.argtypes = [H, C.c_int32, C.c_uint64, C.c_double, C.c_int64]
    lib.sb_scorer_create.argtypes = [C.POINTER(SbScoreModel), C.c_int32, C.POINTER(H)]
    lib.sb_scorer_destroy.argtypes = [H]
    lib.sb_scorer_destroy.restype = None
    lib.sb_scorer_score.argtypes = [H, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p]
    lib.sb_scorer_last_kernel_ms.argtypes = [H, C.POINTER(C.c_float)]
    lib.sb_dist_init.argtypes = [H, C.c_void_p, C.c_int32, C.c_int32]
    lib.sbh_grow_tree.argtypes = [H, H, C.POINTER(SbIterIn), C.POINTER(C.POINTER(SbIterIn)), C.POINTER(C.c_int32),
                                  C.POINTER(SbIterOut)]
    lib.sbh_master_do_compute.restype = C.POINTER(SbIterIn)
    lib.sbh_master_tree_count.argtypes = [H]
    lib.sbh_master_halted.argtypes = [H]
    lib.sbh_master_export_tree.argtypes = [H, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.sb_error_sums_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]
    lib.sb_rows_streamed.argtypes = [H, C.POINTER(C.c_int64), C.c_int32]
    lib.sb_msg_pack.argtypes = [C.POINTER(SbIterIn), C.c_int64, C.c_void_p, C.c_int64]
    lib.sb_msg_pack.restype = C.c_int64
    lib.sb_msg_unpack.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]
    lib.sb_msg_iter_in.argtypes = [C.c_void_p]
    lib.sb_msg_iter_in.restype = C.POINTER(SbIterIn)
    lib.sb_msg_free.argtypes = [C.c_void_p]
    lib.sb_msg_free.restype = None
    lib.sb_category_map_create.argtypes = [C.POINTER(C.c_char_p), C.c_int32, C.POINTER(C.c_void_p)]
    lib.sb_category_map_lookup.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int64, C.c_void_p, C.c_int32]
    lib.sb_category_map_destroy.argtypes = [C.c_void_p]
    lib.sb_category_map_destroy.restype = None
    if path is None:
        _lib = lib
    return lib


class SbError(RuntimeError):
    """Raised for a non-zero sb_status; the reference throws RuntimeException / IllegalStateException here."""

    def __init__(self, code, msg):
        super().__init__("shifu_b200 error %d: %s" % (code, msg))
        self.code = code


def check(lib, rc):
    if rc != 0:
        raise SbError(rc, lib.sb_last_error().decode())


def _ptr(a, ctype):
    return a.ctypes.data_as(C.POINTER(ctype))


class ConfigArrays:
    """Keeps the numpy arrays an SbConfig points at alive."""

    def __init__(self, n_bins, is_categorical=None, column_num=None, bin_boundary=None, impurity=SB_IMPURITY_VARIANCE,
                 num_classes=2, loss=SB_LOSS_SQUARED, is_gbdt=True, tree_num=1, weights_per_row=1, max_depth=8,
                 min_instances_per_node=1, min_info_gain=0.0, learning_rate=0.05, device=0, flags=0):
        self.n_bins = np.ascontiguousarray(n_bins, dtype=np.int32)
        F = len(self.n_bins)
        self.is_categorical = np.ascontiguousarray(
            is_categorical if is_categorical is not None else np.zeros(F), dtype=np.uint8)
        self.column_num = np.ascontiguousarray(column_num if column_num is not None else np.arange(F), dtype=np.int32)
        if bin_boundary is None:
            # default boundaries [-inf, 1, 2, ...] for numeric slots
            bb = []
            for f in range(F):
                if not self.is_categorical[f]:
                    bb.extend([-np.inf] + [float(b) for b in range(1, int(self.n_bins[f]))])
            bin_boundary = bb
        self.bin_boundary = np.ascontiguousarray(bin_boundary, dtype=np.float64)
        c = SbConfig()
        c.n_features = F
        c.n_bins = _ptr(self.n_bins, C.c_int32)
        c.is_categorical = _ptr(self.is_categorical, C.c_uint8)
        c.column_num = _ptr(self.column_num, C.c_int32)
        c.bin_boundary = _ptr(self.bin_boundary, C.c_double)
        c.impurity, c.num_classes, c.loss = impurity, num_classes, loss
        c.is_gbdt, c.tree_num, c.weights_per_row = int(bool(is_gbdt)), tree_num, weights_per_row
        c.max_depth, c.min_instances_per_node = max_depth, min_instances_per_node
        c.min_info_gain, c.learning_rate = min_info_gain, learning_rate
        c.device, c.flags = device, flags
        self.struct = c
        self.stats_size = 3 if impurity in (SB_IMPURITY_VARIANCE, SB_IMPURITY_FRIEDMANMSE) else num_classes

    def boundaries_of(self, f):
        if not hasattr(self, "_bb_off"):
            sizes = np.where(self.is_categorical != 0, 0, self.n_bins).astype(np.int64)
            self._bb_off = np.concatenate([[0], np.cumsum(sizes)])
        off = int(self._bb_off[f])
        return self.bin_boundary[off:off + int(self.n_bins[f])]


class Worker:
    """Thin object wrapper over an sb_handle: the DTWorker of one GPU."""

    def __init__(self, cfg: ConfigArrays):
        self.lib = load_library()
        self.cfg = cfg
        self.h = C.c_void_p()
        check(self.lib, self.lib.sb_create(C.byref(cfg.struct), C.byref(self.h)))

    def close(self):
        if self.h:
            self.lib.sb_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_rows(self, bins_cols, label, significance=None, weights=None, is_validation=False):
        """bins_cols: list of 1-D arrays, one per feature slot (copied to uint8, or uint16 for features with more than
        256 bins); weights: [weights_per_row, n] float32."""
        n = len(label)
        self._keep = [np.ascontiguousarray(b, dtype=np.uint16 if self.cfg.n_bins[f] > 256 else np.uint8)
                      for f, b in enumerate(bins_cols)]
        ptrs = (C.c_void_p * len(self._keep))(*[b.ctypes.data for b in self._keep])
        label = np.ascontiguousarray(label, dtype=np.float32)
        sig = None if significance is None else np.ascontiguousarray(significance, dtype=np.float32)
        w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float32)
        check(self.lib, self.lib.sb_load_rows(self.h, int(is_validation), n, ptrs, label.ctypes.data,
                                              None if sig is None else sig.ctypes.data,
                                              None if w is None else w.ctypes.data))

    def set_weights(self, w_index, weights):
        w = np.ascontiguousarray(weights, dtype=np.float32)
        check(self.lib, self.lib.sb_set_weights(self.h, w_index, w.ctypes.data))

    def sample_weights(self, w_index, seed, rate, first_draw=0):
        check(self.lib, self.lib.sb_sample_weights(self.h, w_index, seed, rate, first_draw))

    def load_synthetic(self, n_rows, seed, row0=0, label_kind=0, weighted=False, is_validation=False):
        check(self.lib, self.lib.sb_load_synthetic(self.h, int(is_validation), row0, n_rows, seed, label_kind,
                                                   int(weighted)))

    def read_bins(self, feature, row0, n_rows, is_validation=False):
        out = np.empty(n_rows, dtype=np.uint16 if self.cfg.n_bins[feature] > 256 else np.uint8)
        check(self.lib, self.lib.sb_read_bins(self.h, int(is_validation), feature, row0, n_rows, out.ctypes.data))
        return out

    def read_labels(self, row0, n_rows, is_validation=False):
        lab = np.empty(n_rows, dtype=np.float32)
        sig = np.empty(n_rows, dtype=np.float32)
        check(self.lib, self.lib.sb_read_labels(self.h, int(is_validation), row0, n_rows, lab.ctypes.data,
                                                sig.ctypes.data))
        return lab, sig

    def read_rows(self, n, is_validation=False, tree_id=0, w_index=0, want_weight=True):
        predict = np.empty(n, dtype=np.float32)
        output = np.empty(n, dtype=np.float32)
        node = np.empty(n, dtype=np.int32)
        weight = np.empty(n, dtype=np.float32) if (want_weight and not is_validation) else None
        check(self.lib, self.lib.sb_read_rows(self.h, int(is_validation), tree_id, w_index, predict.ctypes.data,
                                              output.ctypes.data, node.ctypes.data,
                                              None if weight is None else weight.ctypes.data))
        return predict, output, node, weight

    def stats_len(self, it):
        n = self.lib.sb_stats_len(self.h, C.byref(it.struct))
        if n < 0:
            check(self.lib, int(n))
        return int(n)

    def compute(self, it, want_stats=True):
        """DTWorker.doCompute: returns (train_count, validation_count, train_error, validation_error, stats)."""
        n = self.stats_len(it)
        out = SbIterOut()
        stats = np.zeros(max(n, 1), dtype=np.float64)
        if want_stats:
            out.stats = _ptr(stats, C.c_double)
            out.stats_len = n
        check(self.lib, self.lib.sb_compute(self.h, C.byref(it.struct), C.byref(out)))
        return out.train_count, out.validation_count, out.train_error, out.validation_error, stats[:n]

    def compute_begin(self, it):
        check(self.lib, self.lib.sb_compute_begin(self.h, C.byref(it.struct)))

    def pass_info(self):
        """(scales_fresh, acc_fresh) of the last compute_begin: which small buffers need an all-reduce this pass."""
        a, b = C.c_int32(), C.c_int32()
        check(self.lib, self.lib.sb_pass_info(self.h, C.byref(a), C.byref(b)))
        return bool(a.value), bool(b.value)

    def compute_stats(self):
        check(self.lib, self.lib.sb_compute_stats(self.h))

    def compute_finish(self, n_stats=0, want_stats=False):
        out = SbIterOut()
        stats = np.zeros(max(n_stats, 1), dtype=np.float64)
        if want_stats:
            out.stats = _ptr(stats, C.c_double)
            out.stats_len = n_stats
        check(self.lib, self.lib.sb_compute_finish(self.h, C.byref(out)))
        return out.train_count, out.validation_count, out.train_error, out.validation_error, stats[:n_stats]

    def compute_finish_split(self, n_todo, with_cat=False, out=None):
        """compute_finish + best_splits behind one synchronisation: ((scalars), gains, bits)."""
        out = out or SbIterOut()
        gains = (SbGain * max(n_todo, 1))()
        bits = np.zeros(max(n_todo, 1) * SB_MAX_CAT_BYTES, dtype=np.uint8) if with_cat else None
        check(self.lib, self.lib.sb_compute_finish_split(self.h, C.byref(out), gains, None if bits is None else bits.ctypes.data))
        return (out.train_count, out.validation_count, out.train_error, out.validation_error), gains, bits

    def sample_weights_stratified(self, w_index, seed, rate, n_classes):
        check(self.lib, self.lib.sb_sample_weights_stratified(self.h, w_index, seed, rate, n_classes))

    def sample_poisson(self, w_index, seed, rate, row0=0):
        check(self.lib, self.lib.sb_sample_poisson(self.h, w_index, seed, rate, row0))

    def dist_init(self, id128, rank, world_size):
        """Join the library's own NCCL communicator (id128: the 128 bytes of dist_unique_id() made on rank 0)."""
        buf = (C.c_uint8 * 128).from_buffer_copy(bytes(id128))
        check(self.lib, self.lib.sb_dist_init(self.h, buf, rank, world_size))

    def device_buffer(self, which):
        """(ptr, bytes) of one of the "scale" / "raw" / "acc" buffers."""
        p, b = C.c_void_p(), C.c_int64()
        fn = {"scale": self.lib.sb_scale_buffer, "raw": self.lib.sb_raw_buffer, "acc": self.lib.sb_acc_buffer}[which]
        check(self.lib, fn(self.h, C.byref(p), C.byref(b)))
        return p.value, b.value

    def device_buffers(self):
        """(scale_ptr, scale_bytes, raw_ptr, raw_bytes, acc_ptr, acc_bytes): what a multi-GPU driver all-reduces."""
        p, b = C.c_void_p(), C.c_int64()
        res = []
        for fn in (self.lib.sb_scale_buffer, self.lib.sb_raw_buffer, self.lib.sb_acc_buffer):
            check(self.lib, fn(self.h, C.byref(p), C.byref(b)))
            res += [p.value, b.value]
        return tuple(res)

    def best_splits(self, n_todo, with_cat=False):
        gains = (SbGain * max(n_todo, 1))()
        bits = np.zeros(max(n_todo, 1) * SB_MAX_CAT_BYTES, dtype=np.uint8) if with_cat else None
        check(self.lib, self.lib.sb_best_splits(self.h, gains, None if bits is None else bits.ctypes.data))
        return gains, bits  # gains: ctypes array, entries [0, n_todo) are valid

    def time_stats_pass(self, n_repeat=3):
        a, b = C.c_float(), C.c_float()
        check(self.lib, self.lib.sb_time_stats_pass(self.h, n_repeat, C.byref(a), C.byref(b)))
        return a.value, b.value

    def counters(self):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        check(self.lib, self.lib.sb_counters(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def rows_streamed(self, reset=False):
        """Rows the statistics kernels read since creation / the last reset."""
        r = C.c_int64()
        check(self.lib, self.lib.sb_rows_streamed(self.h, C.byref(r), int(reset)))
        return r.value

    def set_profiling(self, on=True):
        check(self.lib, self.lib.sb_set_profiling(self.h, int(on)))

    def hist_time(self):
        """(ms_total, n_passes, n_launches) of the histogram kernel since the last call."""
        ms, a, b = C.c_float(), C.c_int32(), C.c_int32()
        check(self.lib, self.lib.sb_hist_time(self.h, C.byref(ms), C.byref(a), C.byref(b)))
        return ms.value, a.value, b.value

    def set_stream(self, stream_ptr):
        check(self.lib, self.lib.sb_set_stream(self.h, C.c_void_p(stream_ptr)))


def dist_unique_id():
    """128 bytes for Worker.dist_init, to be made on one rank and handed to all."""
    lib = load_library()
    buf = (C.c_uint8 * 128)()
    check(lib, lib.sb_dist_unique_id(buf))
    return bytes(buf)


class NativeMessage:
    """A DTMasterParams message owned by a NativeMaster (valid until its next call); duck-types messages.IterIn."""

    def __init__(self, ptr):
        self.struct = ptr.contents
        self.n_todo = self.struct.n_todo

    @property
    def todo(self):
        n = self.n_todo
        return {"node_id": np.array([self.struct.todo[i].node_id for i in range(n)], dtype=np.int32)}


class NativeMaster:
    """The C++ DTMaster mirror (csrc/host_master.cpp) behind the sbh_* entry points."""

    def __init__(self, cfg: ConfigArrays, tree_num, max_depth, is_gbdt=True, learning_rate=0.05,
                 feature_subset_strategy="ALL", feature_subset_rate=1.0, feature_sampling_seed=0, max_leaves=-1,
                 max_batch_split_size=32, max_stats_memory=256 << 20, worker_number=1):
        self.lib = load_library()
        self.cfg = cfg
        p = SbhParams()
        p.tree_num, p.max_depth, p.is_gbdt = tree_num, max_depth, int(bool(is_gbdt))
        p.feature_subset_strategy = SBH_FSS[feature_subset_strategy]
        p.feature_subset_rate, p.learning_rate = feature_subset_rate, learning_rate
        p.feature_sampling_seed, p.max_leaves = feature_sampling_seed, max_leaves
        p.max_batch_split_size, p.max_stats_memory, p.worker_number = max_batch_split_size, max_stats_memory, worker_number
        self.h = C.c_void_p()
        check(self.lib, self.lib.sbh_master_create(C.byref(cfg.struct), C.byref(p), C.byref(self.h)))

    def close(self):
        if self.h:
            self.lib.sbh_master_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def build_initial_master_params(self):
        return NativeMessage(self.lib.sbh_master_initial(self.h))

    def do_compute(self, gains, cat_bits=None):
        """gains: the (SbGain * n) array or list returned by Worker.best_splits."""
        if isinstance(gains, list):
            gains = (SbGain * max(len(gains), 1))(*gains)
        ptr = self.lib.sbh_master_do_compute(self.h, gains, None if cat_bits is None else cat_bits.ctypes.data)
        return NativeMessage(ptr) if ptr else None

    def grow_tree(self, worker, message):
        """The iteration loop of one tree in native code: (next message or None, passes, (4 scalars))."""
        nxt = C.POINTER(SbIterIn)()
        n = C.c_int32()
        out = SbIterOut()
        check(self.lib, self.lib.sbh_grow_tree(self.h, worker.h, C.byref(message.struct), C.byref(nxt), C.byref(n), C.byref(out)))
        return (NativeMessage(nxt) if nxt else None, n.value,
                (out.train_count, out.validation_count, out.train_error, out.validation_error))

    @property
    def n_trees(self):
        return self.lib.sbh_master_tree_count(self.h)

    def export_tree(self, index):
        """(nodes structured array in pre-order, mask pool bytes, extras [n, 7])."""
        nn, pb = C.c_int(), C.c_int()
        check(self.lib, self.lib.sbh_master_export_tree(self.h, index, None, 0, None, 0, None, C.byref(nn), C.byref(pb)))
        nodes = (SbNode * max(nn.value, 1))()
        pool = np.zeros(max(pb.value, 1), dtype=np.uint8)
        extras = np.zeros((max(nn.value, 1), 7), dtype=np.float64)
        check(self.lib, self.lib.sbh_master_export_tree(self.h, index, nodes, nn.value, pool.ctypes.data, pb.value,
                                                        extras.ctypes.data, C.byref(nn), C.byref(pb)))
        return list(nodes)[:nn.value], pool[:pb.value], extras[:nn.value]


class CategoryMap:
    """ColumnConfig.binCategory -> bin indices (sb_category_map_*): groups "a@^b" share an index, empty and unknown
    values get the missing-value bin len(bin_category)."""

    def __init__(self, bin_category):
        self.lib = load_library()
        self.n = len(bin_category)
        arr = (C.c_char_p * max(self.n, 1))(*[c.encode("utf-8") for c in bin_category])
        self.h = C.c_void_p()
        check(self.lib, self.lib.sb_category_map_create(arr, self.n, C.byref(self.h)))

    def lookup(self, values):
        n = len(values)
        width = 2 if self.n + 1 > 256 else 1
        out = np.zeros(n, dtype=np.uint16 if width == 2 else np.uint8)
        arr = (C.c_char_p * max(n, 1))(*[None if v is None else v.encode("utf-8") for v in values])
        check(self.lib, self.lib.sb_category_map_lookup(self.h, arr, n, out.ctypes.data, width))
        return out

    def close(self):
        if self.h:
            self.lib.sb_category_map_destroy(self.h)
            self.h = C.c_void_p()
