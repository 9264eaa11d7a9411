"""Shared test helpers: random binned data, tree growth driven by gain records, comparisons."""
import numpy as np

from shifu_b200.capi import SB_SPLIT_CATEGORICAL, SB_SPLIT_CONTINUOUS
from shifu_b200.messages import Node


def make_data(n, n_bins, seed, is_categorical=None, label="binary", weighted=False, wpr=1, poisson=False):
    rng = np.random.default_rng(seed)
    F = len(n_bins)
    cols = [rng.integers(0, n_bins[f], size=n, dtype=np.int64).astype(np.uint16 if n_bins[f] > 256 else np.uint8)
            for f in range(F)]
    z = sum((cols[f].astype(np.float64) / n_bins[f] - 0.5) * (3.0 if f % 2 == 0 else -2.0) for f in range(min(F, 5)))
    if label == "binary":
        y = (rng.random(n) < 1.0 / (1.0 + np.exp(-z))).astype(np.float32)
    elif label == "real":
        y = (z + rng.normal(0, 0.3, n)).astype(np.float32)
    else:
        raise ValueError(label)
    sig = (0.5 + 1.5 * rng.random(n)).astype(np.float32) if weighted else None
    if poisson:
        w = rng.poisson(1.0, size=(wpr, n)).astype(np.float32)
    elif wpr > 1:
        w = np.ones((wpr, n), dtype=np.float32)
    else:
        w = None
    return cols, y, sig, w


def apply_gain(cfg, node, g, cat_bits=None, node_idx=0):
    """populateGainInfoToNode (dt/DTMaster.java:710-727) + child creation (:568-605) for tests."""
    node.predict = g.predict
    node.feature = g.feature
    node.split_type = g.split_type
    node.gain = g.gain
    node.impurity = g.impurity
    node.left_impurity, node.right_impurity = g.left_impurity, g.right_impurity
    node.left_predict, node.right_predict = g.left_predict, g.right_predict
    node.wgt_cnt = g.wgt_cnt
    if g.split_type == SB_SPLIT_CONTINUOUS:
        node.threshold = float(cfg.boundaries_of(g.feature)[g.split_bin + 1])
    else:
        from shifu_b200.capi import SB_MAX_CAT_BYTES
        node.is_left = bool(g.is_left)
        node.cat_bits = bytes(cat_bits[node_idx * SB_MAX_CAT_BYTES: node_idx * SB_MAX_CAT_BYTES + g.cat_bytes])
    node.left = Node(2 * node.id, predict=g.left_predict)
    node.right = Node(2 * node.id + 1, predict=g.right_predict)
    node.is_leaf = False
    return node.left, node.right


def rel_close(a, b, rtol=1e-9, scale=None):
    """|a-b| <= rtol*|b| + 1e-12*scale.  `scale` = sum of |contributions|: two fp64 summation orders of the same
    numbers already differ by ~1e-16*scale, and a sum that cancels to ~0 has no meaningful relative error."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    s = np.max(np.abs(b)) if scale is None else scale
    return np.abs(a - b) <= rtol * np.abs(b) + 1e-12 * s


def gain_fields(g):
    return dict(found=g.found, feature=g.feature, split_type=g.split_type, split_bin=g.split_bin, is_left=g.is_left,
                gain=g.gain, impurity=g.impurity, predict=g.predict, left_impurity=g.left_impurity,
                right_impurity=g.right_impurity, left_predict=g.left_predict, right_predict=g.right_predict,
                wgt_cnt=g.wgt_cnt, cat_bytes=g.cat_bytes, left_class=g.left_class, right_class=g.right_class,
                node_class=g.node_class)
