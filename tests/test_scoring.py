"""f2: batch scoring (IndependentTreeModel.compute, dt/IndependentTreeModel.java:352-446, predictNode :516-560).

CPU part: the oracle's restatement against the plain-Python scorer of model_io on the reference's own model fixture
(readablespec/model0.gbt) and on hand-built trees that exercise every branch of predictNode.  GPU part: the CUDA
scorer through the C ABI against the oracle, bit for bit for raw / cut-off scores and class values, 2 ulp for the
sigmoid converts (exp)."""
import os

import numpy as np
import pytest

from oracle import pyoracle
from shifu_b200 import model_io
from shifu_b200.capi import SB_SPLIT_CATEGORICAL, SB_SPLIT_CONTINUOUS, ScoreModel, Scorer

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURE = os.path.join(HERE, "golden", "readablespec_model0.gbt")


def leaf(v, cls=0):
    return {"split": None, "left": None, "right": None, "predict": (v, cls)}


def num(col, thr, left, right):
    return {"split": {"column": col, "feature_type": SB_SPLIT_CONTINUOUS, "threshold": thr}, "left": left, "right": right,
            "predict": None}


def cat(col, cats, is_left, left, right):
    bits = bytearray(max(cats, default=0) // 8 + 1)
    for c in cats:
        bits[c >> 3] |= 1 << (c % 8)
    return {"split": {"column": col, "feature_type": SB_SPLIT_CATEGORICAL, "categories": bytes(bits), "is_left": is_left},
            "left": left, "right": right, "predict": None}


def hand_built_bags():
    # column 0, 1 numeric; column 2 categorical with 5 categories (index 5 = missing), column 3 categorical with 12
    t0 = num(0, 0.5, cat(2, [1, 3], True, leaf(0.25, 1), leaf(-0.5, 0)), num(1, -1.0, leaf(1.5, 1), leaf(0.75, 0)))
    t1 = cat(3, [0, 11, 12], False, num(0, 2.0, leaf(0.125, 0), leaf(0.0625, 1)), leaf(-2.0, 1))
    t2 = leaf(0.3, 1)                                   # a tree that is only its root
    t3 = num(1, 0.0, leaf(3.0, 0), cat(2, [5], True, leaf(7.0, 1), leaf(-7.0, 0)))   # the missing category on the left
    return [[(t0, 1.0), (t1, 0.1), (t2, 0.1)], [(t3, 0.5), (t0, 0.25)]]


def edge_rows(rng, n, n_columns, cat_cols):
    x = rng.normal(size=(n, n_columns)) * 2.0
    for c, size in cat_cols.items():
        x[:, c] = rng.integers(-2, size + 3, size=n).astype(np.float64)
        frac = rng.random(n) < 0.2
        x[frac, c] += rng.choice([0.9999999, 0.5, -0.05, 0.95], size=int(frac.sum()))
    specials = np.array([np.nan, np.inf, -np.inf, -0.0, 0.0, 0.5, -1.0, 2.0, 1e300, -1e300])
    for c in range(n_columns):
        idx = rng.choice(n, size=min(n, 40), replace=False)
        x[idx, c] = rng.choice(specials, size=len(idx))
    return x


def python_score_rows(bags, cat_size, x, is_gbdt=True):
    """Plain-Python restatement, independent of the C++ oracle (different code, same source lines)."""
    def walk(n, row):
        while n["split"] is not None and (n["left"] is not None or n["right"] is not None):
            s = n["split"]
            v = float(row[s["column"]])
            if s["feature_type"] == SB_SPLIT_CONTINUOUS:
                n = n["left"] if v < s["threshold"] else n["right"]
            else:
                size = cat_size[s["column"]]
                neg = v < 0 or (v == 0 and np.signbit(v))
                ge = np.isnan(v) or v >= size
                idx = size if (neg or ge) else int(np.int16(int(v + 0.1)))
                bits = s["categories"]
                has = 0 <= idx and (idx >> 3) < len(bits) and (bits[idx >> 3] >> (idx % 8)) & 1
                n = n["left"] if bool(has) == bool(s["is_left"]) else n["right"]
        return n
    out = np.zeros(len(x))
    for r, row in enumerate(x):
        fin = 0.0
        for bag in bags:
            if is_gbdt:
                p = 0.0
                for root, w in bag:
                    p += walk(root, row)["predict"][0] * w
                fin += p
            else:
                ps = ws = 0.0
                for root, w in bag:
                    ws += w
                    ps += walk(root, row)["predict"][0] * w
                fin += ps / ws
        out[r] = fin / len(bags)
    return out


def test_oracle_scoring_matches_python_on_hand_built_trees():
    rng = np.random.default_rng(3)
    bags, cat_size = hand_built_bags(), [0, 0, 5, 12]
    x = edge_rows(rng, 3000, 4, {2: 5, 3: 12})
    for is_gbdt in (True, False):
        m = ScoreModel(bags, 4, cat_size, is_gbdt=is_gbdt)
        got = pyoracle.score_rows(m, x)
        ref = python_score_rows(bags, cat_size, x, is_gbdt)
        assert np.array_equal(got, ref)
    assert len(np.unique(got)) > 8  # the rows reach many different leaves


def test_oracle_scoring_on_the_reference_model_fixture():
    m = model_io.read_model(FIXTURE)
    sm = model_io.to_score_model(m)
    assert sm.n_trees == sum(len(b) for b in m["bags"]) == 100
    rng = np.random.default_rng(5)
    x = rng.normal(size=(400, sm.n_columns)) * np.array([30.0] * sm.n_columns) + 10.0
    x[::7, 3] = np.nan
    got = pyoracle.score_rows(sm, x)
    ref = np.array([model_io.score_gbt(m, row) for row in x])
    assert np.array_equal(got, ref)
    # converts
    raw = got
    for name, fn in (("OLD_SIGMOID", lambda s: 1 / (1 + np.minimum(1e19, np.exp(-s)))),
                     ("SIGMOID", lambda s: 1 / (1 + np.minimum(1e19, np.exp(-20 * s)))),
                     ("CUTOFF", lambda s: np.clip(s, 0.0, 1.0))):
        conv = pyoracle.score_rows(model_io.to_score_model(m, convert=name), x)
        assert np.allclose(conv, fn(raw), rtol=4e-16, atol=0)


def test_oracle_reports_a_missing_child():
    broken = num(0, 0.0, leaf(1.0), None)
    m = ScoreModel([[(broken, 1.0)]], 1)
    assert pyoracle.score_rows(m, np.array([[-1.0]]))[0] == 1.0
    with pytest.raises(RuntimeError):
        pyoracle.score_rows(m, np.array([[1.0]]))


@pytest.mark.gpu
def test_device_scoring_bit_exact_on_hand_built_trees():
    rng = np.random.default_rng(8)
    bags, cat_size = hand_built_bags(), [0, 0, 5, 12]
    x = edge_rows(rng, 70001, 4, {2: 5, 3: 12})  # ragged last block
    for kw in (dict(is_gbdt=True), dict(is_gbdt=False), dict(is_gbdt=True, convert="CUTOFF"),
               dict(is_gbdt=False, is_classification=True)):
        m = ScoreModel(bags, 4, cat_size, **kw)
        sc = Scorer(m)
        got = sc.score(x)
        assert np.array_equal(got, pyoracle.score_rows(m, x)), kw
        assert sc.last_kernel_ms() > 0
        sc.close()
    for conv in ("OLD_SIGMOID", "SIGMOID"):
        m = ScoreModel(bags, 4, cat_size, convert=conv)
        sc = Scorer(m)
        assert np.allclose(sc.score(x), pyoracle.score_rows(m, x), rtol=5e-16, atol=0)
        sc.close()


@pytest.mark.gpu
def test_device_scoring_reference_fixture_and_wide_rows():
    m = model_io.read_model(FIXTURE)
    sm = model_io.to_score_model(m)
    rng = np.random.default_rng(9)
    x = rng.normal(size=(200000, sm.n_columns)) * 30.0 + 10.0
    x[::11, 5] = np.nan
    sc = Scorer(sm)
    got = sc.score(x)
    assert np.array_equal(got, pyoracle.score_rows(sm, x))
    assert np.array_equal(sc.score(x[:1]), got[:1]) and len(sc.score(x[:0])) == 0
    sc.close()
    # more columns than the shared-memory staging takes: the direct-read variant of the kernel
    wide = 260
    trees = [[(num(int(c), 0.1 * k, leaf(1.0 + k), num(int((c * 7) % wide), -0.3, leaf(-1.0), leaf(0.5 * k))), 0.3)
              for k, c in enumerate(rng.integers(0, wide, size=40))]]
    wm = ScoreModel(trees, wide, None)
    xs = rng.normal(size=(5000, wide))
    sc = Scorer(wm)
    assert np.array_equal(sc.score(xs), pyoracle.score_rows(wm, xs))
    sc.close()


@pytest.mark.gpu
def test_device_scoring_missing_child_is_an_error():
    m = ScoreModel([[(num(0, 0.0, leaf(1.0), None), 1.0)]], 1)
    sc = Scorer(m)
    assert sc.score(np.array([[-1.0]]))[0] == 1.0
    with pytest.raises(RuntimeError):
        sc.score(np.array([[1.0]]))
    sc.close()
