"""GPU parity for the branches of the reference path that the main parity file does not take: dropout, the
"friedmanmse" master-side quirk, MinInfoGain > 0, a wide dynamic range of contributions, weight changes between the
levels of a tree, RF over adopted rows without a weight array, and a pass in which this GPU has no active row."""
import numpy as np
import pytest

from shifu_b200 import capi
from shifu_b200.capi import ConfigArrays, Worker
from shifu_b200.messages import IterIn, Node

from helpers import apply_gain, make_data
from test_gpu_parity import index_to_level, run_gbt

pytestmark = pytest.mark.gpu


def test_gbt_dropout_rows_keep_their_predict():
    """dt/DTWorker.java:634-638: a dropped row does not take the finished tree's contribution; its output is still
    recomputed from the unchanged predict.  run_gbt compares predict / output row by row after every pass."""
    assert run_gbt(9000, [32] * 12, seed=31, n_trees=4, max_depth=3, dropout_rate=0.3) >= 8
    run_gbt(7000, [16] * 8, seed=32, n_trees=3, max_depth=3, dropout_rate=0.5, loss=capi.SB_LOSS_LOG, weighted=True)


def test_gbt_friedmanmse_scans_as_variance_on_the_master():
    """dt/DTMaster.java:1026-1032: the master creates a Variance impurity for "friedmanmse" (the worker's FriedmanMSE
    only differs in computeImpurity, which the worker never calls).  Statistics and splits against the oracle."""
    assert run_gbt(12000, [32] * 16 + [7, 13], seed=33, n_trees=2, max_depth=4, impurity=capi.SB_IMPURITY_FRIEDMANMSE,
                   is_categorical=[0] * 16 + [1, 1], weighted=True) >= 6


def test_gbt_min_info_gain_skips_candidates():
    """dt/Impurity.java:166 (`gain <= minInfoGain` candidates are skipped): with a threshold inside the range of the
    gains some nodes find no split at all and the tree stops early -- on both sides alike."""
    seen = []

    def collect(it, d, o, gains):
        seen.extend(gains[k].gain for k in range(it.n_todo) if gains[k].found)

    run_gbt(15000, [32] * 10, seed=34, n_trees=2, max_depth=5, min_info_gain=0.0, on_pass=collect)
    threshold = float(np.median(seen))  # half of the winning gains of the unrestricted run lie below it
    kept = []

    def collect_kept(it, d, o, gains):
        kept.extend(gains[k].gain for k in range(it.n_todo) if gains[k].found)

    run_gbt(15000, [32] * 10, seed=34, n_trees=2, max_depth=5, min_info_gain=threshold, on_pass=collect_kept)
    assert 0 < len(kept) < len(seen) and min(kept) > threshold, (len(kept), len(seen), threshold)


def test_wide_dynamic_range_of_contributions():
    """Log loss deep into the boosting run (residuals of well-classified rows shrink by orders of magnitude) with
    significance spread log-uniformly over 2^-12 .. 2^12: every bin within 1e-9 of the oracle's bin, no global slack
    (assert_stats_close_per_bin).  This is the proven envelope of the 44-bit fixed point with one scale per tree:
    a contribution is rounded at max|v| * 2^-45, so a bin stays within 1e-9 as long as its mean |contribution| is not
    below ~max|v| / 35000."""
    run_gbt(400_000, [32] * 10, seed=35, n_trees=32, max_depth=4, loss=capi.SB_LOSS_LOG, sig_log2_range=12,
            per_bin_strict=True, with_validation=False, learning_rate=0.3, oracle_threads=8)


def grow_levels(worker, cfg, n_levels, poke=None):
    """Level-wise growth of one tree through sb_compute; poke(level) runs between two levels.  Returns the statistics."""
    root = Node(1)
    todo_nodes, switch, out = [root], True, []
    for level in range(n_levels):
        if not todo_nodes:
            break
        todo = [(k, 0, nd.id, []) for k, nd in enumerate(todo_nodes)]
        it = IterIn(cfg, [(0, root)], todo, switch_to_next_tree=switch, is_first_tree=True)
        res = worker.compute(it)
        out.append(res[4].copy())
        gains, bits = worker.best_splits(len(todo))
        nxt = []
        for k, nd in enumerate(todo_nodes):
            g = gains[k]
            if not g.found or g.gain <= 0:
                continue
            left, right = apply_gain(cfg, nd, g, bits, k)
            nxt += [left, right]
        todo_nodes, switch = nxt, False
        if poke is not None:
            poke(level)
    return out


def test_weights_changed_between_levels_do_not_reuse_the_parent():
    """A weight change inside a tree invalidates the parent's accumulators: the children must be accumulated directly,
    not derived as parent - sibling (the parent was built with the old weights).  Same statistics as with
    SB_FLAG_NO_SUBTRACTION, bit for bit."""
    n, n_bins = 20000, [32] * 12
    cols, y, sig, _ = make_data(n, n_bins, seed=36, weighted=False)
    rng = np.random.default_rng(5)
    new_w = [(rng.random(n) < 0.6).astype(np.float32) for _ in range(3)]
    runs = []
    for flags in (0, 4):  # 4 = SB_FLAG_NO_SUBTRACTION
        cfg = ConfigArrays(n_bins, max_depth=5, min_instances_per_node=1, flags=flags)
        w = Worker(cfg)
        w.load_rows(cols, y, sig, np.ones((1, n), np.float32))
        runs.append(grow_levels(w, cfg, 4, poke=lambda level: w.set_weights(0, new_w[level]) if level < 3 else None))
        w.close()
    assert len(runs[0]) == len(runs[1]) == 4
    for a, b in zip(runs[0], runs[1]):
        assert np.array_equal(a, b)


def test_rf_over_adopted_rows_without_weights():
    """sb_load_rows_device with weights == NULL: every sample weight is 1 and no row is out of bag; the RF error pass
    used to dereference the missing array."""
    import torch
    from oracle.pyoracle import OracleWorker
    n, n_bins = 6000, [16] * 6
    cols, y, _, _ = make_data(n, n_bins, seed=37)
    cfg = ConfigArrays(n_bins, impurity=capi.SB_IMPURITY_GINI, num_classes=2, is_gbdt=False, tree_num=2,
                       weights_per_row=2, max_depth=3)
    pitch = ((n + 255) // 256) * 256
    mat = np.zeros((len(n_bins), pitch), np.uint8)
    for f, c in enumerate(cols):
        mat[f, :n] = c
    d_bins, d_y = torch.from_numpy(mat).cuda(), torch.from_numpy(y).cuda()
    d_sig = torch.ones(n, dtype=torch.float32, device="cuda")
    dev, orc = Worker(cfg), OracleWorker(cfg)
    capi.check(dev.lib, dev.lib.sb_load_rows_device(dev.h, 0, n, d_bins.data_ptr(), pitch, d_y.data_ptr(), d_sig.data_ptr(), None))
    orc.load_rows(cols, y, None, np.ones((2, n), np.float32))
    roots = [Node(1), Node(1)]
    it = IterIn(cfg, [(0, roots[0]), (1, roots[1])], [(0, 0, 1, []), (1, 1, 1, [])])
    d, o = dev.compute(it), orc.compute(it, n_threads=2)
    assert np.array_equal(d[4], o[4])
    for a, b in zip(d[:4], o[:4]):
        assert abs(a - b) <= 1e-9 * abs(b) + 1e-12
    gains, bits = dev.best_splits(2)
    for k in range(2):
        apply_gain(cfg, roots[k], gains[k], bits, k)
    it2 = IterIn(cfg, [(0, roots[0]), (1, roots[1])], [(0, 0, 2, []), (1, 0, 3, []), (2, 1, 2, [])])
    d2, o2 = dev.compute(it2), orc.compute(it2, n_threads=2)  # the error pass now walks real trees
    assert np.array_equal(d2[4], o2[4])
    for a, b in zip(d2[:4], o2[:4]):
        assert abs(a - b) <= 1e-9 * abs(b) + 1e-12
    dev.close()
    orc.close()


def test_pass_without_any_active_row_on_this_gpu():
    """A shard can hold no row of any todo node (deep levels, sharded rows; or weight 0 on all rows of the nodes): the
    row lists are empty, hist_rows_kernel must do nothing, and the statistics of the pass are all zero."""
    n, n_bins = 5000, [64] * 8
    cols, y, sig, _ = make_data(n, n_bins, seed=38)
    cfg = ConfigArrays(n_bins, max_depth=4, min_instances_per_node=1)
    w = Worker(cfg)
    wts = np.ones((1, n), np.float32)
    w.load_rows(cols, y, sig, wts)
    root = Node(1)
    it = IterIn(cfg, [(0, root)], [(0, 0, 1, [])], switch_to_next_tree=True, is_first_tree=True)
    w.compute(it)
    gains, bits = w.best_splits(1)
    left, right = apply_gain(cfg, root, gains[0], bits, 0)
    # every row gets weight 0: the children exist, but none of their rows counts
    w.set_weights(0, np.zeros(n, np.float32))
    it2 = IterIn(cfg, [(0, root)], [(0, 0, 2, []), (1, 0, 3, [])])
    res = w.compute(it2)
    assert np.all(res[4] == 0.0)
    # and a todo node that no row reaches at all (heap id 12 does not exist in the tree: its slot stays empty)
    w.set_weights(0, np.ones(n, np.float32))
    res3 = w.compute(IterIn(cfg, [(0, root)], [(0, 0, 2, []), (1, 0, 3, [])]))
    assert res3[4].reshape(2, -1, 3)[:, :, 0].sum() == float(n) * len(n_bins)
    w.close()
