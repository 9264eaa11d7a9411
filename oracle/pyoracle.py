"""ctypes binding of oracle/liboracle.so (the C++ restatement of DTWorker / DTMaster / Impurity / Loss).

TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED: see the header of shifu_oracle.cpp.
The message structs (sb_iter_in, sb_gain, ...) are shared with the product's C ABI so the same message can be
handed to both sides.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from shifu_b200.capi import SB_MAX_CAT_BYTES, SbConfig, SbGain, SbIterIn, SbIterOut

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    lib = C.CDLL(LIB_PATH)
    V = C.c_void_p
    lib.orc_last_error.restype = C.c_char_p
    lib.orc_create.argtypes = [C.POINTER(SbConfig), C.POINTER(V)]
    lib.orc_destroy.argtypes = [V]
    lib.orc_destroy.restype = None
    lib.orc_load_rows.argtypes = [V, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.orc_read_rows.argtypes = [V, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.orc_stats_len.argtypes = [V, C.POINTER(SbIterIn)]
    lib.orc_stats_len.restype = C.c_int64
    lib.orc_compute.argtypes = [V, C.POINTER(SbIterIn), C.POINTER(SbIterOut), C.c_int]
    lib.orc_compute_impurity.argtypes = [V, C.c_int, C.c_void_p, C.c_int, C.POINTER(SbGain), C.c_void_p]
    lib.orc_best_splits.argtypes = [V, C.POINTER(SbIterIn), C.c_void_p, C.POINTER(SbGain), C.c_void_p]
    lib.orc_hashmap_order.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    lib.orc_get_bin_index.argtypes = [C.c_float, C.c_void_p, C.c_int]
    lib.orc_loss_gradient.argtypes = [C.c_int, C.c_float, C.c_float]
    lib.orc_loss_gradient.restype = C.c_float
    lib.orc_loss_error.argtypes = [C.c_int, C.c_float, C.c_float]
    lib.orc_loss_error.restype = C.c_float
    lib.orc_feature_update.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float]
    lib.orc_feature_update.restype = None
    lib.orc_java_random_doubles.argtypes = [C.c_int64, C.c_int, C.c_void_p]
    lib.orc_score_rows.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    lib.orc_java_random_doubles.restype = None
    lib.orc_index_to_level.argtypes = [C.c_int]
    lib.orc_flatten_cat_val_grp.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    lib.orc_category_index.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.c_char_p]
    from shifu_b200.capi import SbhParams
    lib.orc_master_create.argtypes = [C.POINTER(SbConfig), C.POINTER(SbhParams), C.POINTER(V)]
    lib.orc_master_destroy.argtypes = [V]
    lib.orc_master_destroy.restype = None
    lib.orc_master_initial.argtypes = [V]
    lib.orc_master_initial.restype = C.POINTER(SbIterIn)
    lib.orc_master_do_compute.argtypes = [V, C.POINTER(SbGain), C.c_void_p]
    lib.orc_master_do_compute.restype = C.POINTER(SbIterIn)
    lib.orc_master_tree_count.argtypes = [V]
    lib.orc_master_dump_tree.argtypes = [V, C.c_int, C.c_void_p, C.c_int]
    _lib = lib
    return lib


class OracleWorker:
    """DTWorker restated on the CPU; rows are kept in the reference's own layout (row-major short[] inputs)."""

    def __init__(self, cfg):
        self.lib = load()
        self.cfg = cfg
        self.h = C.c_void_p()
        rc = self.lib.orc_create(C.byref(cfg.struct), C.byref(self.h))
        assert rc == 0

    def close(self):
        if self.h:
            self.lib.orc_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_rows(self, bins_cols, label, significance=None, weights=None, is_validation=False):
        """Same arguments as shifu_b200.capi.Worker.load_rows (column list; weights [wpr, n])."""
        n = len(label)
        inputs = np.ascontiguousarray(np.stack([np.asarray(b) for b in bins_cols], axis=1).astype(np.int16))
        label = np.ascontiguousarray(label, dtype=np.float32)
        sig = None if significance is None else np.ascontiguousarray(significance, dtype=np.float32)
        w = None if weights is None else np.ascontiguousarray(np.asarray(weights, dtype=np.float32).T)
        rc = self.lib.orc_load_rows(self.h, int(is_validation), n, inputs.ctypes.data, label.ctypes.data,
                                    None if sig is None else sig.ctypes.data, None if w is None else w.ctypes.data)
        assert rc == 0

    def read_rows(self, n, is_validation=False):
        predict = np.empty(n, dtype=np.float32)
        output = np.empty(n, dtype=np.float32)
        rc = self.lib.orc_read_rows(self.h, int(is_validation), predict.ctypes.data, output.ctypes.data, None)
        assert rc == 0
        return predict, output

    def compute(self, it, n_threads=1):
        n = int(self.lib.orc_stats_len(self.h, C.byref(it.struct)))
        out = SbIterOut()
        stats = np.zeros(max(n, 1), dtype=np.float64)
        out.stats = stats.ctypes.data_as(C.POINTER(C.c_double))
        out.stats_len = n
        rc = self.lib.orc_compute(self.h, C.byref(it.struct), C.byref(out), n_threads)
        if rc != 0:
            raise RuntimeError(self.lib.orc_last_error().decode())
        return out.train_count, out.validation_count, out.train_error, out.validation_error, stats[:n]

    def best_splits(self, it, stats, with_cat=False):
        gains = (SbGain * max(it.n_todo, 1))()
        bits = np.zeros(max(it.n_todo, 1) * SB_MAX_CAT_BYTES, dtype=np.uint8) if with_cat else None
        stats = np.ascontiguousarray(stats, dtype=np.float64)
        rc = self.lib.orc_best_splits(self.h, C.byref(it.struct), stats.ctypes.data, gains,
                                      None if bits is None else bits.ctypes.data)
        assert rc == 0
        return list(gains)[:it.n_todo], bits

    def compute_impurity(self, feature, stats, master_quirk=True):
        g = SbGain()
        bits = np.zeros(SB_MAX_CAT_BYTES, dtype=np.uint8)
        stats = np.ascontiguousarray(stats, dtype=np.float64)
        rc = self.lib.orc_compute_impurity(self.h, feature, stats.ctypes.data, int(master_quirk), C.byref(g),
                                           bits.ctypes.data)
        assert rc == 0
        return g, bits


def score_rows(model, data):
    """IndependentTreeModel.compute restated (orc_score_rows) over the rows of `data`; model: capi.ScoreModel."""
    import numpy as np
    lib = load()
    data = np.ascontiguousarray(data, dtype=np.float64)
    out = np.zeros((data.shape[0], model.out_per_row()), dtype=np.float64)
    rc = lib.orc_score_rows(C.byref(model.struct), data.shape[0], data.ctypes.data, out.ctypes.data)
    if rc != 0:
        raise RuntimeError("NullPointerException in predictNode: a split node is missing a child")
    return out if model.is_classification else out[:, 0]


class OracleMaster:
    """DTMaster's growth rules restated on the CPU (oracle_master.cpp); same constructor and calls as
    shifu_b200.capi.NativeMaster, so a test can drive both with one stream of gain records."""

    def __init__(self, cfg, tree_num, max_depth, is_gbdt=True, learning_rate=0.05, feature_subset_strategy="ALL",
                 feature_subset_rate=1.0, feature_sampling_seed=0, max_leaves=-1, max_batch_split_size=32,
                 max_stats_memory=256 << 20, worker_number=1):
        from shifu_b200.capi import SBH_FSS, NativeMessage, SbhParams
        self._message = NativeMessage
        self.lib = load()
        self.cfg = cfg
        p = SbhParams()
        p.tree_num, p.max_depth, p.is_gbdt = tree_num, max_depth, int(bool(is_gbdt))
        p.feature_subset_strategy = SBH_FSS[feature_subset_strategy]
        p.feature_subset_rate, p.learning_rate = feature_subset_rate, learning_rate
        p.feature_sampling_seed, p.max_leaves = feature_sampling_seed, max_leaves
        p.max_batch_split_size, p.max_stats_memory, p.worker_number = max_batch_split_size, max_stats_memory, worker_number
        self.h = C.c_void_p()
        rc = self.lib.orc_master_create(C.byref(cfg.struct), C.byref(p), C.byref(self.h))
        if rc != 0:
            raise RuntimeError("orc_master_create failed: %d" % rc)

    def close(self):
        if self.h:
            self.lib.orc_master_destroy(self.h)
            self.h = C.c_void_p()

    def build_initial_master_params(self):
        return self._message(self.lib.orc_master_initial(self.h))

    def do_compute(self, gains, cat_bits=None):
        if isinstance(gains, list):
            gains = (SbGain * max(len(gains), 1))(*gains)
        ptr = self.lib.orc_master_do_compute(self.h, gains, None if cat_bits is None else cat_bits.ctypes.data)
        return self._message(ptr) if ptr else None

    @property
    def n_trees(self):
        return self.lib.orc_master_tree_count(self.h)

    def dump_tree(self, index):
        """[n_nodes, 13] in pre-order: id, isLeaf, hasSplit, feature, gain, impurity, leftImpurity, rightImpurity,
        leftPredict, rightPredict, wgtCnt, wgtCntRatio, predict."""
        n = self.lib.orc_master_dump_tree(self.h, index, None, 0)
        out = np.zeros((max(n, 1), 13), dtype=np.float64)
        self.lib.orc_master_dump_tree(self.h, index, out.ctypes.data, n)
        return out[:n]
