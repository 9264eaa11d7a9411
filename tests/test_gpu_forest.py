"""RF passes as ONE list for all trees (plan_forest / forest_* kernels / hist_rows_class_kernel) against the oracle and
against the per-tree launch sequence they replace (SB_FLAG_NO_FOREST_PASS): integer sums, so bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from shifu_b200 import capi  # noqa: E402
from shifu_b200.capi import ConfigArrays, Worker  # noqa: E402
from shifu_b200.messages import IterIn, Node  # noqa: E402

from helpers import apply_gain, make_data  # noqa: E402
from test_gpu_parity import index_to_level, is_pos_zero, run_rf  # noqa: E402

NO_FOREST = 16  # SB_FLAG_NO_FOREST_PASS


def test_forest_many_slot_groups_gini():
    """6 trees to depth 8: up to 6 x 128 todo nodes in a pass -> more than a hundred slot groups, groups that span trees."""
    assert run_rf(30000, [32] * 40, seed=61, n_trees=6, max_depth=8) >= 7


def test_forest_partly_filled_tile_and_two_tiles():
    run_rf(9000, [16] * 37, seed=62, n_trees=3, max_depth=5)          # one tile, 2 features per lane populated
    run_rf(9000, [8] * 150, seed=63, n_trees=3, max_depth=4)          # two tiles of 128, the second partly filled
    run_rf(5000, [200] * 12, seed=64, n_trees=3, max_depth=4)         # wide histograms: fewer features per lane


def test_forest_general_operands():
    """Non-integer weights (no ClassOps), more than two classes, and a regression forest (variance statistics)."""
    run_rf(9000, [16] * 20 + [8] * 5, seed=65, n_trees=4, max_depth=6, impurity=capi.SB_IMPURITY_ENTROPY, num_classes=3, weighted=True)
    run_rf(9000, [32] * 24, seed=66, n_trees=4, max_depth=5, impurity=capi.SB_IMPURITY_GINI, num_classes=5)
    run_rf(9000, [32] * 24, seed=67, n_trees=4, max_depth=5, impurity=capi.SB_IMPURITY_VARIANCE)
    run_rf(9000, [32] * 24, seed=68, n_trees=3, max_depth=5, impurity=capi.SB_IMPURITY_VARIANCE, weighted=True)


def test_forest_wide_features():
    """Features with more than 256 bins (uint16 columns): tiles of 8 in hist_rows_class_wide_kernel over the same pair list."""
    run_rf(12000, [32] * 6 + [501, 300, 11, 1001], seed=73, n_trees=4, max_depth=5, is_categorical=[0] * 6 + [1, 1, 1, 1])
    run_rf(8000, [64] * 3 + [257] * 9 + [201], seed=74, n_trees=3, max_depth=5, impurity=capi.SB_IMPURITY_ENTROPY,
           is_categorical=[0] * 3 + [1] * 9 + [1])                                   # two wide tiles, one of them partly filled
    run_rf(6000, [400, 1001, 300], seed=75, n_trees=3, max_depth=4, is_categorical=[0, 1, 1])   # uint16 columns only
    # non-integer weights: no 8-byte operands, the pass falls back to one launch sequence per tree
    run_rf(6000, [32] * 4 + [300, 600], seed=76, n_trees=3, max_depth=4, is_categorical=[0] * 4 + [1, 1], weighted=True)


def test_forest_in_batches_of_trees(monkeypatch):
    """An operand buffer that holds the pairs of 4 trees only: the 7 trees of the pass go through in two batches."""
    monkeypatch.setenv("SB_FOREST_OPS_MB", "1")  # 2^20 bytes / 8 per pair = 131072 pairs = 4 trees of 30000 rows
    assert run_rf(30000, [32] * 20, seed=69, n_trees=7, max_depth=6) >= 5


def test_per_tree_launches_still_match_the_oracle():
    run_rf(12000, [32] * 40, seed=21, n_trees=4, max_depth=5, flags=NO_FOREST)
    run_rf(9000, [16] * 20 + [8] * 5, seed=22, n_trees=3, max_depth=4, impurity=capi.SB_IMPURITY_ENTROPY, num_classes=3,
           weighted=True, flags=NO_FOREST)


def _grow_both(n, n_bins, seed, n_trees, max_depth, impurity, weighted):
    # 0: forest passes, full (every valid feature accumulated, siblings paired); 16: one launch sequence per tree (full as
    # well); 4: SB_FLAG_NO_SUBTRACTION -- forest passes that accumulate every listed (node, feature) directly
    cfgs = [ConfigArrays(n_bins, impurity=impurity, num_classes=2, loss=capi.SB_LOSS_SQUARED, is_gbdt=False, tree_num=n_trees,
                         weights_per_row=n_trees, max_depth=max_depth, min_instances_per_node=1, flags=f) for f in (0, NO_FOREST, 4)]
    cols, y, sig, w = make_data(n, n_bins, seed, weighted=weighted, wpr=n_trees, poisson=True)
    vcols, vy, vsig, _ = make_data(n // 3, n_bins, seed + 5, weighted=weighted)
    workers = [Worker(c) for c in cfgs]
    for wk in workers:
        wk.load_rows(cols, y, sig, w)
        wk.load_rows(vcols, vy, vsig, None, is_validation=True)
    rng = np.random.default_rng(seed + 1)
    F = len(n_bins)
    roots = [Node(1) for _ in range(n_trees)]
    todo_nodes = [(t, roots[t]) for t in range(n_trees)]
    passes = 0
    while todo_nodes:
        todo = [(k, t, nd.id, rng.choice(F, size=max(1, (2 * F) // 3), replace=False).tolist()) for k, (t, nd) in enumerate(todo_nodes)]
        it = IterIn(cfgs[0], [(t, roots[t]) for t in range(n_trees)], todo)
        a, b, c = workers[0].compute(it), workers[1].compute(it), workers[2].compute(it)
        assert np.array_equal(a[4], b[4]), "forest pass and per-tree passes must give the same statistics, bit for bit"
        assert np.array_equal(a[4], c[4]), "paired (parent - sibling) and directly accumulated statistics must be the same, bit for bit"
        for x, z in zip(a[:4], b[:4]):  # fp64 sums taken in a different order
            assert abs(x - z) <= 1e-12 * abs(z) + 1e-300
        ga, _ = workers[0].best_splits(len(todo))
        gb, _ = workers[1].best_splits(len(todo))
        workers[2].best_splits(len(todo))
        assert all(p.found == q.found and p.feature == q.feature and p.gain == q.gain for p, q in zip(ga, gb))
        passes += 1
        nxt = []
        for k, (t, nd) in enumerate(todo_nodes):
            g = ga[k]
            if not g.found:
                continue
            level = index_to_level(nd.id)
            nd.predict = g.predict
            if g.gain <= 0 or level == max_depth:
                continue
            left, right = apply_gain(cfgs[0], nd, g, None, k)
            if not (level + 1 == max_depth or is_pos_zero(g.left_impurity)):
                nxt.append((t, left))
            if not (level + 1 == max_depth or is_pos_zero(g.right_impurity)):
                nxt.append((t, right))
        todo_nodes = nxt
    for wk in workers:
        wk.close()
    return passes


def test_forest_pass_equals_per_tree_passes_bit_for_bit():
    assert _grow_both(40000, [32] * 30, 71, n_trees=5, max_depth=7, impurity=capi.SB_IMPURITY_GINI, weighted=False) >= 6
    assert _grow_both(20000, [32] * 30, 72, n_trees=4, max_depth=6, impurity=capi.SB_IMPURITY_VARIANCE, weighted=True) >= 5


def _one_pass(cfg, cols, y, sig, w, todo, roots, vcols=None, vy=None):
    from oracle.pyoracle import OracleWorker
    dev, orc = Worker(cfg), OracleWorker(cfg)
    for wk in (dev, orc):
        wk.load_rows(cols, y, sig, w)
        if vcols is not None:
            wk.load_rows(vcols, vy, None, None, is_validation=True)
    it = IterIn(cfg, roots, todo)
    d, o = dev.compute(it), orc.compute(it, n_threads=2)
    dev.close()
    orc.close()
    return d, o


def test_forest_ragged_and_empty_inputs():
    """Rows 0, 1, 119, 120, 121, 4095, 4096, 4097 (tile of 120 pairs, range of 4096 rows), a tree none of whose rows is in
    its bag, and a pass whose only active tree is the last one."""
    T, n_bins = 3, [32] * 9 + [7]
    cfg = ConfigArrays(n_bins, impurity=capi.SB_IMPURITY_GINI, num_classes=2, loss=capi.SB_LOSS_SQUARED, is_gbdt=False, tree_num=T,
                       weights_per_row=T, max_depth=3, min_instances_per_node=1)
    roots = [(t, Node(1)) for t in range(T)]
    todo = [(t, t, 1, []) for t in range(T)]
    for n in (0, 1, 119, 120, 121, 4095, 4096, 4097):
        cols, y, sig, w = make_data(n, n_bins, 300 + n, weighted=False, wpr=T, poisson=True)
        d, o = _one_pass(cfg, cols, y, sig, w, todo, roots)
        assert np.array_equal(d[4], o[4]), n
        for a, b in zip(d[:4], o[:4]):
            assert abs(a - b) <= 1e-9 * abs(b) + 1e-12, (n, d[:4], o[:4])
    cols, y, sig, w = make_data(5000, n_bins, 77, weighted=False, wpr=T, poisson=True)
    w[1, :] = 0.0                      # tree 1: every row out of bag
    d, o = _one_pass(cfg, cols, y, sig, w, todo, roots)
    assert np.array_equal(d[4], o[4]) and np.all(d[4].reshape(T, -1)[1] == 0)
    w[0, :] = 0.0                      # only the last tree has pairs
    d, o = _one_pass(cfg, cols, y, sig, w, todo, roots)
    assert np.array_equal(d[4], o[4])
    w[2, :] = 0.0                      # no pair at all: the kernels must do nothing
    d, o = _one_pass(cfg, cols, y, sig, w, todo, roots)
    assert np.array_equal(d[4], o[4]) and not d[4].any()


def test_forest_only_some_trees_have_todo_nodes():
    """Trees 0 and 3 of five are in the pass, each with a node list of its own; the others are walked for the error sums only."""
    T, n_bins = 5, [16] * 20
    cfg = ConfigArrays(n_bins, impurity=capi.SB_IMPURITY_ENTROPY, num_classes=2, loss=capi.SB_LOSS_SQUARED, is_gbdt=False, tree_num=T,
                       weights_per_row=T, max_depth=3, min_instances_per_node=1)
    cols, y, sig, w = make_data(7000, n_bins, 81, weighted=False, wpr=T, poisson=True)
    vcols, vy, _, _ = make_data(900, n_bins, 82, weighted=False)
    roots = [(t, Node(1)) for t in range(T)]
    todo = [(0, 0, 1, [0, 3, 5, 19]), (1, 3, 1, list(range(2, 17)))]
    d, o = _one_pass(cfg, cols, y, sig, w, todo, roots, vcols, vy)
    assert np.array_equal(d[4], o[4])
    for a, b in zip(d[:4], o[:4]):
        assert abs(a - b) <= 1e-9 * abs(b) + 1e-12, (d[:4], o[:4])


def test_gbt_feature_lists_full_passes_equal_direct_accumulation():
    """GBT nodes with feature lists of their own (FeatureSubsetStrategy): the full pass (internal all-features layout,
    sibling subtraction, root count cache, gather into the caller's layout) against SB_FLAG_NO_SUBTRACTION, which
    accumulates exactly the listed (node, feature) blocks -- bit for bit, over two trees, with lists that differ between
    siblings and a root that takes every feature."""
    n, n_bins = 30000, [64] * 20 + [16] * 5 + [300]
    is_cat = [0] * 25 + [1]
    cols, y, sig, _ = make_data(n, n_bins, seed=91, weighted=True, is_categorical=is_cat)
    F = len(n_bins)
    runs = []
    for flags in (0, 4):
        cfg = ConfigArrays(n_bins, is_categorical=is_cat, max_depth=5, tree_num=2, min_instances_per_node=1, flags=flags)
        w = Worker(cfg)
        w.load_rows(cols, y, sig, None)
        rng = np.random.default_rng(7)
        stats_all, roots = [], []
        for t in range(2):
            root = Node(1)
            roots.append(root)
            todo_nodes, switch = [root], True
            while todo_nodes:
                todo = [(k, t, nd.id, [] if nd.id == 1 else sorted(rng.choice(F, size=F // 2, replace=False).tolist()))
                        for k, nd in enumerate(todo_nodes)]
                it = IterIn(cfg, [(i, r) for i, r in enumerate(roots)], todo, switch_to_next_tree=switch, is_first_tree=(t == 0))
                res = w.compute(it)
                stats_all.append(res[4].copy())
                gains, bits = w.best_splits(len(todo), with_cat=True)
                nxt = []
                for k, nd in enumerate(todo_nodes):
                    g = gains[k]
                    nd.predict = g.predict if g.found else nd.predict
                    if not g.found or g.gain <= 0 or index_to_level(nd.id) == 5:
                        continue
                    left, right = apply_gain(cfg, nd, g, bits, k)
                    if index_to_level(nd.id) + 1 < 5:
                        nxt += [left, right]
                todo_nodes, switch = nxt, False
        runs.append(stats_all)
        w.close()
    assert len(runs[0]) == len(runs[1]) >= 8
    for a, b in zip(runs[0], runs[1]):
        assert np.array_equal(a, b)
