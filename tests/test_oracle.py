"""Pins the CPU oracle (oracle/shifu_oracle.cpp) without the reference: independent numpy re-derivations of the
restated formulas, published known answers of java.util.Random, self-consistency, and the cancer-judgement fixture
binned from the reference's own ColumnConfig.json.  PARITY UNPINNED: no reference-produced golden values exist
for this path."""
import ctypes as C
import os

import numpy as np
import pytest

from shifu_b200 import capi
from shifu_b200.capi import ConfigArrays
from shifu_b200.messages import IterIn, Node

from helpers import make_data

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def orc():
    from oracle import pyoracle
    pyoracle.build()
    return pyoracle


def test_java_random_known_answers(orc):
    lib = orc.load()
    out = np.zeros(3)
    lib.orc_java_random_doubles(42, 3, out.ctypes.data)
    assert out[0] == 0.7275636800328681  # new java.util.Random(42).nextDouble()
    lib.orc_java_random_doubles(0, 1, out.ctypes.data)
    assert out[0] == 0.730967787376657   # new java.util.Random(0).nextDouble()
    from java_random import JavaRandom
    r = JavaRandom(42)
    lib.orc_java_random_doubles(42, 3, out.ctypes.data)
    assert [r.next_double() for _ in range(3)] == list(out)


def test_feature_update_float_semantics(orc):
    lib = orc.load()
    rng = np.random.default_rng(1)
    for _ in range(200):
        y, s, w = (np.float32(x) for x in rng.normal(size=3) * 3)
        st = np.zeros(9)
        lib.orc_feature_update(capi.SB_IMPURITY_VARIANCE, 3, st.ctypes.data, 1, C.c_float(y), C.c_float(s), C.c_float(w))
        assert st[3] == np.float64(np.float32(s * w))
        assert st[4] == np.float64(np.float32(np.float32(y * s) * w))
        assert st[5] == np.float64(np.float32(np.float32(np.float32(y * y) * s) * w))
    st = np.zeros(6)
    lib.orc_feature_update(capi.SB_IMPURITY_GINI, 2, st.ctypes.data, 2, C.c_float(0.9999999), C.c_float(2.0), C.c_float(3.0))
    assert st[5] == 6.0  # (int)(0.9999999f + 1e-6f) == 1


def test_loss_functions(orc):
    lib = orc.load()
    f = np.float32
    assert lib.orc_loss_gradient(capi.SB_LOSS_SQUARED, f(0.25), f(1.0)) == f(2.0) * f(f(0.25) - f(1.0))
    assert lib.orc_loss_gradient(capi.SB_LOSS_HALFGRADSQUARED, f(0.25), f(1.0)) == f(-0.75)
    assert lib.orc_loss_gradient(capi.SB_LOSS_ABSOLUTE, f(0.25), f(1.0)) == -1.0  # label > predict
    assert lib.orc_loss_gradient(capi.SB_LOSS_ABSOLUTE, f(1.5), f(1.0)) == 1.0
    p, y = f(0.3), f(1.0)
    g = f(f(2) - f(4) * y) / f(np.exp(np.float64(f(f(f(4) * y) * p) - f(f(2) * p))))
    assert lib.orc_loss_gradient(capi.SB_LOSS_LOG, p, y) == g
    e = f(np.log1p(1.0 + np.exp(np.float64(f(f(2) * p) - f(f(f(4) * p) * y)))))
    assert lib.orc_loss_error(capi.SB_LOSS_LOG, p, y) == e
    assert lib.orc_loss_error(capi.SB_LOSS_SQUARED, f(0.25), f(1.0)) == f(0.5625)


def test_get_bin_index_against_fixture(orc):
    lib = orc.load()
    z = np.load(os.path.join(HERE, "golden", "cancer_judgement.npz"))
    for f in range(z["bins"].shape[0]):
        bb = np.ascontiguousarray(z["boundaries"][f])
        for r in range(0, z["bins"].shape[1], 7):
            assert lib.orc_get_bin_index(C.c_float(z["raw"][f, r]), bb.ctypes.data, len(bb)) == z["bins"][f, r]
    bb = np.array([-np.inf, 1.0, 1.0, 2.0])
    assert lib.orc_get_bin_index(C.c_float(np.inf), bb.ctypes.data, 4) == 3
    assert lib.orc_get_bin_index(C.c_float(-np.inf), bb.ctypes.data, 4) == 0
    assert lib.orc_get_bin_index(C.c_float(1.0), bb.ctypes.data, 4) == 2  # duplicate boundaries: the later bin
    assert lib.orc_get_bin_index(C.c_float(5.0), bb.ctypes.data, 1) == 0


def test_hashmap_order(orc):
    lib = orc.load()
    keys = np.array([5, 3, 9, 1], dtype=np.int32)  # cap 4: buckets 1,3,1,1 -> 5, 9, 1 (insertion order), then 3
    out = np.zeros(4, dtype=np.int32)
    lib.orc_hashmap_order(keys.ctypes.data, 4, out.ctypes.data)
    assert list(keys[out]) == [5, 9, 1, 3]
    keys = np.arange(3, 203, dtype=np.int32)  # 200 columns, cap 256, all < cap: ascending
    out = np.zeros(200, dtype=np.int32)
    lib.orc_hashmap_order(keys.ctypes.data, 200, out.ctypes.data)
    assert list(out) == list(range(200))


def variance_gain_numpy(st, min_inst, min_gain):
    """Independent statement of Variance.computeImpurity for a numeric feature (dt/Impurity.java:120-211)."""
    c, s, q = st[:, 0], st[:, 1], st[:, 2]
    C_, S_, Q_ = c.sum(), s.sum(), q.sum()
    imp = (Q_ - S_ * S_ / C_) / C_ if C_ != 0 else 0.0
    best = None
    for i in range(len(c) - 1):
        lc, ls, lq = c[:i + 1].sum(), s[:i + 1].sum(), q[:i + 1].sum()
        rc, rs, rq = C_ - lc, S_ - ls, Q_ - lq
        if lc <= min_inst or rc <= min_inst:
            continue
        li = (lq - ls * ls / lc) / lc
        ri = (rq - rs * rs / rc) / rc
        g = imp - lc / C_ * li - rc / C_ * ri
        if g <= min_gain:
            continue
        if best is None or g > best[1]:
            best = (i, g, ls / lc, rs / rc)
    return best


def gini_entropy_gain_numpy(st, entropy, min_inst):
    tot = st.sum(axis=0)

    def imp(v):
        n = v.sum()
        if n == 0:
            return 0.0
        r = v / n
        if entropy:
            r = r[r > 0]
            return float(-(r * np.log(r) / np.log(2)).sum())
        return float(-(r * r).sum())  # the reference stores gini as -sum(p^2)
    best = None
    for i in range(st.shape[0] - 1):
        l = st[:i + 1].sum(axis=0)
        r = tot - l
        if l.sum() <= min_inst or r.sum() <= min_inst:
            continue
        g = imp(tot) - l.sum() / tot.sum() * imp(l) - r.sum() / tot.sum() * imp(r)
        if g <= 0:
            continue
        if best is None or g > best[1]:
            best = (i, g)
    return best


def test_compute_impurity_against_numpy(orc):
    rng = np.random.default_rng(3)
    cfg = ConfigArrays([12] * 2, min_instances_per_node=2)
    o = orc.OracleWorker(cfg)
    for _ in range(50):
        cnt = rng.integers(0, 40, size=12).astype(np.float64)
        mean = rng.normal(size=12)
        st = np.stack([cnt, cnt * mean, cnt * (mean ** 2 + 0.1)], axis=1)
        g, _ = o.compute_impurity(0, st.ravel())
        ref = variance_gain_numpy(st, 2, 0.0)
        if ref is None:
            assert g.found == 0
            continue
        assert g.found == 1 and g.split_bin == ref[0]
        assert np.isclose(g.gain, ref[1], rtol=1e-9, atol=1e-15)
        assert np.isclose(g.left_predict, ref[2], rtol=1e-12) and np.isclose(g.right_predict, ref[3], rtol=1e-12)
    for imp_kind, ent in ((capi.SB_IMPURITY_GINI, False), (capi.SB_IMPURITY_ENTROPY, True)):
        cfg = ConfigArrays([10] * 2, impurity=imp_kind, num_classes=3, is_gbdt=False, min_instances_per_node=1)
        o = orc.OracleWorker(cfg)
        for _ in range(50):
            st = rng.integers(0, 30, size=(10, 3)).astype(np.float64)
            g, _ = o.compute_impurity(1, st.ravel())
            ref = gini_entropy_gain_numpy(st, ent, 1)
            if ref is None:
                assert g.found == 0
                continue
            assert g.found == 1 and g.split_bin == ref[0] and np.isclose(g.gain, ref[1], rtol=1e-9)


def test_categorical_order_and_bitset(orc):
    # 4 categories + missing bin; means 3, (empty), 1, 2, 0.5 -> order: empty(MIN_VALUE), missing(.5), c2(1), c3(2), c0(3)
    cfg = ConfigArrays([5], is_categorical=[1], min_instances_per_node=0)
    o = orc.OracleWorker(cfg)
    cnt = np.array([10, 0, 10, 10, 10.0])
    mean = np.array([3.0, 0.0, 1.0, 2.0, 0.5])
    st = np.stack([cnt, cnt * mean, cnt * (mean ** 2 + 0.01)], axis=1)
    g, bits = o.compute_impurity(0, st.ravel())
    assert g.found == 1 and g.split_type == capi.SB_SPLIT_CATEGORICAL
    # i = 0 is the empty bin: leftCount 0 <= minInstances(0) -> skipped and NOT added to the left set (reference quirk)
    # best split separates {missing, c2} from {c3, c0}: prefix i = 2
    assert g.split_bin == 2
    left = {v for v in range(5) if (bits[v >> 3] >> (v % 8)) & 1}
    assert g.is_left == 1 and left == {4, 2}  # missing bin is stored as index nCats (= 4)
    assert g.cat_bytes == 5 // 8 + 1


def test_do_compute_threads_agree_and_match_numpy(orc):
    n, n_bins = 5000, [16] * 6 + [5]
    cols, y, sig, _ = make_data(n, n_bins, seed=5, weighted=True)
    cfg = ConfigArrays(n_bins, min_instances_per_node=1)
    it = IterIn(cfg, [(0, Node(1))], [(0, 0, 1, [])], switch_to_next_tree=True, is_first_tree=True)
    res = []
    for th in (1, 4, 7):
        o = orc.OracleWorker(cfg)
        o.load_rows(cols, y, sig, None)
        res.append(o.compute(it, n_threads=th))
    for r in res[1:]:
        assert np.allclose(r[4], res[0][4], rtol=1e-12, atol=1e-12)
        assert r[0] == res[0][0]
    # independent numpy histogram of feature 2
    off = 2 * 16 * 3
    st = res[0][4][off:off + 48].reshape(16, 3)
    w = sig.astype(np.float32)
    for b in range(16):
        m = cols[2] == b
        assert np.isclose(st[b, 0], w[m].astype(np.float64).sum(), rtol=1e-12)
        assert np.isclose(st[b, 1], (y[m] * w[m]).astype(np.float32).astype(np.float64).sum(), rtol=1e-12, atol=1e-12)
    # iteration 1: predict == label, so the training error is exactly 0 (dt/DTWorker.java:681)
    assert res[0][2] == 0.0 and np.isclose(res[0][0], sig.astype(np.float64).sum())


def test_cancer_judgement_fixture_first_split(orc):
    """C0: the reference's own example data.  The root split of the first tree must separate malignant from benign
    on one of the strong 'worst' features and the master rule must accept it."""
    z = np.load(os.path.join(HERE, "golden", "cancer_judgement.npz"))
    bins, label = z["bins"], z["label"]
    cfg = ConfigArrays([10] * 30, column_num=z["column_num"], bin_boundary=z["boundaries"].ravel(), max_depth=6,
                       min_instances_per_node=1)
    o = orc.OracleWorker(cfg)
    o.load_rows(list(bins), label, None, None)
    it = IterIn(cfg, [(0, Node(1))], [(0, 0, 1, [])], switch_to_next_tree=True, is_first_tree=True)
    tc, vc, te, ve, stats = o.compute(it, n_threads=4)
    assert tc == 429.0 and te == 0.0
    assert stats.reshape(30, 10, 3)[:, :, 0].sum(axis=1).tolist() == [429.0] * 30
    assert stats.reshape(30, 10, 3)[:, :, 1].sum(axis=1).tolist() == [154.0] * 30  # 154 malignant rows
    g, _ = o.best_splits(it, stats)
    assert g[0].found == 1 and g[0].wgt_cnt == 429.0
    assert abs(g[0].predict - 154.0 / 429.0) < 1e-15
    assert 0.1 < g[0].gain < 0.25  # impurity of the root is p(1-p) = 0.230; readablespec/model0.gbt has root gain 0.158
    lp, rp = g[0].left_predict, g[0].right_predict
    assert lp < 0.2 and rp > 0.8
