"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on identical seeded inputs.

Bar (BASELINE.json north_star): per-bin counts bit-exact; weighted sums, gains and error sums within 1e-9
relative; identical chosen split feature / bin; per-row predict / output identical floats.
"""
import numpy as np
import pytest

from shifu_b200 import capi
from shifu_b200.capi import ConfigArrays, Worker
from shifu_b200.messages import IterIn, Node

from helpers import apply_gain, gain_fields, make_data, rel_close

pytestmark = pytest.mark.gpu


def index_to_level(i):
    return int(i).bit_length()


def is_pos_zero(x):
    return x == 0.0 and not np.signbit(x)


def assert_stats_close(cfg, it, dev, orc, contributions_scale, exact_counts):
    assert dev.shape == orc.shape
    ss = cfg.stats_size
    if ss == 3 and cfg.struct.impurity in (capi.SB_IMPURITY_VARIANCE, capi.SB_IMPURITY_FRIEDMANMSE):
        d, o = dev.reshape(-1, 3), orc.reshape(-1, 3)
        if exact_counts:
            assert np.array_equal(d[:, 0], o[:, 0]), "per-bin counts must be bit-exact"
        else:
            assert rel_close(d[:, 0], o[:, 0], scale=contributions_scale[0]).all()
        assert rel_close(d[:, 1], o[:, 1], scale=contributions_scale[1]).all()
        assert rel_close(d[:, 2], o[:, 2], scale=contributions_scale[2]).all()
    else:
        if exact_counts:
            assert np.array_equal(dev, orc), "per-bin class counts must be bit-exact"
        else:
            assert rel_close(dev, orc, scale=contributions_scale[0]).all()


def assert_same_split(gd, go):
    """Device split (from device statistics) against the oracle's split from the ORACLE's statistics: same decision;
    the gain agrees to 1e-9 of the node impurity (a gain is a difference of impurities and an impurity of a nearly pure
    child is itself a cancelling difference, so its own magnitude is no yardstick)."""
    assert gd.found == go.found
    if gd.found:
        assert (gd.feature, gd.split_bin, gd.split_type) == (go.feature, go.split_bin, go.split_type)
        assert abs(gd.gain - go.gain) <= 1e-9 * max(abs(go.gain), abs(go.impurity))
        assert abs(gd.predict - go.predict) <= 1e-9 * abs(go.predict) + 1e-12


def assert_gain_equal(gd, go, bits_d=None, bits_o=None, idx=0):
    """Device computeImpurity / max-gain against the oracle's, both fed the SAME statistics: every field must agree
    (the scan order is the reference's, so the doubles are the same up to log() in the entropy)."""
    d, o = gain_fields(gd), gain_fields(go)
    for k in ("found", "feature", "split_type", "split_bin", "is_left", "cat_bytes", "left_class", "right_class",
              "node_class"):
        assert d[k] == o[k], (k, d, o)
    if not d["found"]:
        return
    for k in ("gain", "impurity", "predict", "left_impurity", "right_impurity", "left_predict", "right_predict",
              "wgt_cnt"):
        assert abs(d[k] - o[k]) <= 1e-12 * abs(o[k]) + 1e-300, (k, d[k], o[k])
    if d["cat_bytes"]:
        a = bits_d[idx * capi.SB_MAX_CAT_BYTES: idx * capi.SB_MAX_CAT_BYTES + d["cat_bytes"]]
        b = bits_o[idx * capi.SB_MAX_CAT_BYTES: idx * capi.SB_MAX_CAT_BYTES + d["cat_bytes"]]
        assert np.array_equal(a, b)


def assert_stats_close_per_bin(dev, orc):
    """The strict reading of the bar: every statistic of every bin within 1e-9 of the ORACLE'S OWN value for that bin,
    with no allowance tied to the size of the whole data set.  The signed sum may cancel inside a bin, so its yardstick
    is the bin's sum of |contributions|, bounded from the bin's other two statistics by Cauchy-Schwarz:
    sum |y s| <= sqrt(sum s * sum y^2 s)."""
    d, o = dev.reshape(-1, 3), orc.reshape(-1, 3)
    assert np.all(np.abs(d[:, 0] - o[:, 0]) <= 1e-9 * np.abs(o[:, 0]))
    assert np.all(np.abs(d[:, 2] - o[:, 2]) <= 1e-9 * np.abs(o[:, 2]))
    yard = np.maximum(np.abs(o[:, 1]), np.sqrt(np.abs(o[:, 0] * o[:, 2])))
    assert np.all(np.abs(d[:, 1] - o[:, 1]) <= 1e-9 * yard)


def run_gbt(n, n_bins, seed, n_trees=2, max_depth=4, weighted=False, loss=capi.SB_LOSS_SQUARED, label="binary",
            is_categorical=None, feature_subsets=False, with_validation=True, bagging=False, min_inst=1,
            oracle_threads=3, flags=0, impurity=capi.SB_IMPURITY_VARIANCE, min_info_gain=0.0, dropout_rate=0.0,
            sig_log2_range=0, per_bin_strict=False, learning_rate=0.1, on_pass=None):
    from oracle.pyoracle import OracleWorker
    cfg = ConfigArrays(n_bins, is_categorical=is_categorical, loss=loss, is_gbdt=True, tree_num=n_trees,
                       max_depth=max_depth, min_instances_per_node=min_inst, learning_rate=learning_rate,
                       column_num=np.arange(len(n_bins)) + 3, flags=flags, impurity=impurity,
                       min_info_gain=min_info_gain)
    cols, y, sig, _ = make_data(n, n_bins, seed, label=label, weighted=weighted)
    rng = np.random.default_rng(seed + 1)
    if sig_log2_range:  # significance spread log-uniformly over 2^-r .. 2^r
        sig = np.exp2(rng.uniform(-sig_log2_range, sig_log2_range, n)).astype(np.float32)
        weighted = True
    w0 = (rng.random(n) < 0.7).astype(np.float32)[None, :] if bagging else None
    dev, orc = Worker(cfg), OracleWorker(cfg)
    dev.load_rows(cols, y, sig, w0)
    orc.load_rows(cols, y, sig, w0)
    if with_validation:
        vcols, vy, vsig, _ = make_data(max(n // 5, 1), n_bins, seed + 7, label=label, weighted=weighted)
        dev.load_rows(vcols, vy, vsig, None, is_validation=True)
        orc.load_rows(vcols, vy, vsig, None, is_validation=True)
    has_cat = is_categorical is not None and any(is_categorical)
    F = len(n_bins)
    trees = []
    n_iter = 0
    for t in range(n_trees):
        root = Node(1)
        trees.append(root)
        todo_nodes = [root]
        switch = True
        while todo_nodes:
            todo = []
            for k, nd in enumerate(todo_nodes):
                feats = []
                if feature_subsets:
                    feats = sorted(rng.choice(F, size=max(1, F // 2), replace=False).tolist())
                    rng.shuffle(feats)
                todo.append((k, t, nd.id, feats))
            sent = [(t - 1, trees[t - 1]), (t, root)] if (switch and t >= 1) else [(t, root)]
            new_w = (rng.random(n) < 0.7).astype(np.float32) if (bagging and switch and t >= 1) else None
            # isFirstTree as the master sets it: trees.size() == 1, evaluated before a new root is added
            # (dt/DTMaster.java:418,475)
            first = (t == 0) or (switch and t == 1)
            # the reference draws dropOutRandom.nextDouble() < dropOutRate per row at a tree switch after the first tree
            # (dt/DTWorker.java:634-638); the draws are injected as a mask so both sides drop the same rows
            drop = (rng.random(n) < dropout_rate).astype(np.uint8) if (dropout_rate > 0 and switch and not first) else None
            it = IterIn(cfg, sent, todo, switch_to_next_tree=switch, is_first_tree=first, new_weights=new_w, dropout=drop)
            d = dev.compute(it)
            o = orc.compute(it, n_threads=oracle_threads)
            n_iter += 1
            # scalars
            for a, b in zip(d[:4], o[:4]):
                assert abs(a - b) <= 1e-9 * abs(b) + 1e-12, (d[:4], o[:4])
            # per-row state after phase A
            dp, dout, dnode, dw = dev.read_rows(n)
            op, oout = orc.read_rows(n)
            if loss == capi.SB_LOSS_LOG:
                assert np.allclose(dp, op, rtol=3e-7, atol=0) and np.allclose(dout, oout, rtol=3e-7, atol=1e-37)
            else:
                assert np.array_equal(dp, op) and np.array_equal(dout, oout)
            wv = np.ones(n, np.float32) if dw is None else dw
            sv = np.ones(n, np.float32) if sig is None else sig
            scale = (np.sum(np.abs(sv * wv)), np.sum(np.abs(dout * sv * wv)), np.sum(np.abs(dout * dout * sv * wv)))
            exact_counts = not weighted
            assert_stats_close(cfg, it, d[4], o[4], scale, exact_counts)
            if per_bin_strict:
                assert_stats_close_per_bin(d[4], o[4])
            gd, bits_d = dev.best_splits(len(todo), with_cat=has_cat)
            if on_pass is not None:
                on_pass(it, d, o, gd)
            go, bits_o = orc.best_splits(it, d[4], with_cat=has_cat)   # oracle scan of the device's statistics
            go2, _ = orc.best_splits(it, o[4], with_cat=has_cat)      # oracle scan of its own statistics
            next_nodes = []
            for k, nd in enumerate(todo_nodes):
                assert_gain_equal(gd[k], go[k], bits_d, bits_o, k)
                assert_same_split(gd[k], go2[k])
                g = gd[k]
                if not g.found:
                    continue
                level = index_to_level(nd.id)
                is_leaf = g.gain <= 0 or level == max_depth
                nd.predict = g.predict
                if is_leaf:
                    continue
                left, right = apply_gain(cfg, nd, g, bits_d, k)
                # Double.compare(x, 0d) == 0 is true for +0.0 only (dt/DTMaster.java:570-571,586-587)
                if not (level + 1 == max_depth or is_pos_zero(g.left_impurity)):
                    next_nodes.append(left)
                if not (level + 1 == max_depth or is_pos_zero(g.right_impurity)):
                    next_nodes.append(right)
            todo_nodes = next_nodes
            switch = False
    dev.close()
    orc.close()
    return n_iter


def test_gbt_variance_unweighted_counts_bit_exact():
    assert run_gbt(20000, [64] * 40, seed=11, n_trees=2, max_depth=5) >= 8


def test_root_count_cache_matches_direct_accumulation():
    # trees 2 and 3 take the root's bin counts from the first root pass (weights unchanged); flag 8 accumulates them
    # every time.  Both are checked against the oracle pass by pass (counts bit-exact).
    assert run_gbt(20000, [64] * 40, seed=13, n_trees=3, max_depth=3) >= 6
    assert run_gbt(20000, [64] * 40, seed=13, n_trees=3, max_depth=3, flags=8) >= 6
    # redrawn bags invalidate the cache
    run_gbt(20000, [64] * 40, seed=14, n_trees=3, max_depth=3, bagging=True)


def test_gbt_deep_levels_row_lists_and_tile_groups():
    # depth 7: up to 32 todo nodes per pass -> several shared-memory slot groups.  Default: per-pass row lists over the
    # row-major bin copy; SB_FLAG_NO_ROW_LISTS: the column-tile kernel visits every tile once per group.
    run_gbt(30000, [64] * 40, seed=18, n_trees=1, max_depth=7, weighted=True)
    run_gbt(30000, [64] * 40, seed=18, n_trees=1, max_depth=7, weighted=True, flags=2)
    run_gbt(30000, [64] * 40, seed=19, n_trees=2, max_depth=6, bagging=True, feature_subsets=True)


def test_histogram_subtraction_is_exact():
    """sibling = parent - child on the integer accumulators must give the very same statistics as accumulating every
    node directly (SB_FLAG_NO_SUBTRACTION), bit for bit, weighted or not."""
    n, n_bins = 25000, [64] * 30 + [16] * 6
    cols, y, sig, _ = make_data(n, n_bins, seed=41, weighted=True)
    runs = []
    for flags in (0, 4):
        cfg = ConfigArrays(n_bins, max_depth=6, min_instances_per_node=1, flags=flags)
        w = Worker(cfg)
        w.load_rows(cols, y, sig, None)
        root = Node(1)
        todo_nodes, switch, stats_all = [root], True, []
        while todo_nodes:
            todo = [(k, 0, nd.id, []) for k, nd in enumerate(todo_nodes)]
            it = IterIn(cfg, [(0, root)], todo, switch_to_next_tree=switch, is_first_tree=True)
            res = w.compute(it)
            stats_all.append(res[4].copy())
            gains, bits = w.best_splits(len(todo))
            nxt = []
            for k, nd in enumerate(todo_nodes):
                g = gains[k]
                if not g.found or g.gain <= 0 or index_to_level(nd.id) == 6:
                    continue
                left, right = apply_gain(cfg, nd, g, bits, k)
                if index_to_level(nd.id) + 1 < 6:
                    nxt += [left, right]
            todo_nodes, switch = nxt, False
        runs.append(stats_all)
        w.close()
    assert len(runs[0]) == len(runs[1]) >= 5
    for a, b in zip(runs[0], runs[1]):
        assert np.array_equal(a, b)


def test_finish_split_equals_finish_then_best_splits():
    """sb_compute_finish_split is sb_compute_finish + sb_best_splits behind one synchronisation: same bytes out."""
    import ctypes as C
    n, n_bins = 20000, [64] * 12 + [9] * 4
    is_cat = [False] * 12 + [True] * 4
    cols, y, sig, _ = make_data(n, n_bins, seed=77, weighted=True)
    outs = []
    for fused in (False, True):
        cfg = ConfigArrays(n_bins, is_categorical=is_cat, max_depth=4, min_instances_per_node=1)
        w = Worker(cfg)
        w.load_rows(cols, y, sig, None)
        root = Node(1)
        todo_nodes, switch, rec = [root], True, []
        while todo_nodes:
            todo = [(k, 0, nd.id, []) for k, nd in enumerate(todo_nodes)]
            it = IterIn(cfg, [(0, root)], todo, switch_to_next_tree=switch, is_first_tree=True)
            w.compute_begin(it)
            w.compute_stats()
            if fused:
                scal, gains, bits = w.compute_finish_split(len(todo), with_cat=True)
            else:
                scal = w.compute_finish()[:4]
                gains, bits = w.best_splits(len(todo), with_cat=True)
            rec.append((tuple(scal), C.string_at(gains, C.sizeof(capi.SbGain) * len(todo)), bits.tobytes()))
            nxt = []
            for k, nd in enumerate(todo_nodes):
                g = gains[k]
                if not g.found or g.gain <= 0 or index_to_level(nd.id) == 4:
                    continue
                left, right = apply_gain(cfg, nd, g, bits, k)
                if index_to_level(nd.id) + 1 < 4:
                    nxt += [left, right]
            todo_nodes, switch = nxt, False
        outs.append(rec)
        w.close()
    assert len(outs[0]) == len(outs[1]) >= 3
    for a, b in zip(outs[0], outs[1]):
        assert a == b


def test_gbt_weighted_mixed_bins():
    run_gbt(15000, [64] * 20 + [32] * 10 + [10] * 7 + [2, 1, 255], seed=12, n_trees=2, max_depth=4, weighted=True)


def test_gbt_feature_subsets_and_bagging():
    run_gbt(12000, [32] * 50, seed=13, n_trees=3, max_depth=4, feature_subsets=True, bagging=True)


def test_gbt_log_loss_real_labels_absolute():
    run_gbt(8000, [16] * 12, seed=14, n_trees=3, max_depth=3, loss=capi.SB_LOSS_LOG)
    run_gbt(8000, [16] * 12, seed=15, n_trees=2, max_depth=3, loss=capi.SB_LOSS_ABSOLUTE, label="real")
    run_gbt(8000, [16] * 12, seed=16, n_trees=2, max_depth=3, loss=capi.SB_LOSS_HALFGRADSQUARED, label="real")


def test_gbt_categorical_features():
    n_bins = [64] * 6 + [11, 51, 201, 31]
    is_cat = [0] * 6 + [1, 1, 1, 1]
    run_gbt(10000, n_bins, seed=17, n_trees=2, max_depth=4, is_categorical=is_cat)


def test_gbt_wide_categorical_uint16_columns():
    """Config C3's column mix in small: numeric 64-bin features plus categorical features with up to 1000 categories
    (1001 statistics bins -> uint16 bin indices, 8-feature tiles in hist_rows_wide_kernel), GBT with log loss."""
    n_bins = [64] * 10 + [301, 1001, 11, 257] + [64] * 3
    is_cat = [0] * 10 + [1, 1, 1, 1] + [0] * 3
    run_gbt(20000, n_bins, seed=51, n_trees=2, max_depth=5, is_categorical=is_cat, loss=capi.SB_LOSS_LOG)
    run_gbt(9000, n_bins, seed=52, n_trees=2, max_depth=4, is_categorical=is_cat, weighted=True, feature_subsets=True)
    # only wide features, and a numeric feature with more than 256 bins
    run_gbt(6000, [400, 1001, 300], seed=53, n_trees=1, max_depth=4, is_categorical=[0, 1, 1])


def test_rf_entropy_wide_categorical():
    run_rf(8000, [32] * 6 + [501, 300], seed=54, n_trees=3, max_depth=4, impurity=capi.SB_IMPURITY_ENTROPY,
           is_categorical=[0] * 6 + [1, 1], feature_subsets=False)


def test_gbt_ragged_sizes():
    for n in (1, 2, 127, 128, 129, 1000):
        run_gbt(n, [8] * 5, seed=100 + n, n_trees=2, max_depth=3, with_validation=(n > 4))


def test_empty_training_set():
    cfg = ConfigArrays([8] * 3)
    w = Worker(cfg)
    w.load_rows([np.zeros(0, np.uint8)] * 3, np.zeros(0, np.float32))
    it = IterIn(cfg, [(0, Node(1))], [(0, 0, 1, [])], switch_to_next_tree=True, is_first_tree=True)
    tc, vc, te, ve, stats = w.compute(it)
    assert tc == 0 and te == 0 and np.all(stats == 0) and len(stats) == 3 * 8 * 3
    w.close()


def test_bin_out_of_range_is_rejected():
    cfg = ConfigArrays([8] * 2)
    w = Worker(cfg)
    with pytest.raises(capi.SbError) as e:
        w.load_rows([np.array([1, 9], np.uint8), np.array([0, 0], np.uint8)], np.zeros(2, np.float32))
    assert e.value.code == -4
    w.close()


def run_rf(n, n_bins, seed, n_trees=4, max_depth=4, impurity=capi.SB_IMPURITY_GINI, num_classes=2, weighted=False,
           is_categorical=None, feature_subsets=True, oracle_threads=3, flags=0):
    """RF: all trees advance together, Poisson bag counts per (row, tree), per-node feature subsets, OOB error."""
    from oracle.pyoracle import OracleWorker
    cfg = ConfigArrays(n_bins, is_categorical=is_categorical, impurity=impurity, num_classes=num_classes,
                       loss=capi.SB_LOSS_SQUARED, is_gbdt=False, tree_num=n_trees, weights_per_row=n_trees,
                       max_depth=max_depth, min_instances_per_node=1, flags=flags)
    cols, y, sig, w = make_data(n, n_bins, seed, weighted=weighted, wpr=n_trees, poisson=True)
    if num_classes > 2:
        rng0 = np.random.default_rng(seed + 3)
        y = ((y + rng0.integers(0, num_classes, n)) % num_classes).astype(np.float32)
    dev, orc = Worker(cfg), OracleWorker(cfg)
    dev.load_rows(cols, y, sig, w)
    orc.load_rows(cols, y, sig, w)
    vcols, vy, vsig, _ = make_data(max(n // 4, 1), n_bins, seed + 9, weighted=weighted)
    dev.load_rows(vcols, vy, vsig, None, is_validation=True)
    orc.load_rows(vcols, vy, vsig, None, is_validation=True)
    has_cat = is_categorical is not None and any(is_categorical)
    rng = np.random.default_rng(seed + 1)
    F = len(n_bins)
    roots = [Node(1) for _ in range(n_trees)]
    todo_nodes = [(t, roots[t]) for t in range(n_trees)]
    n_iter = 0
    while todo_nodes:
        todo = []
        for k, (t, nd) in enumerate(todo_nodes):
            feats = []
            if feature_subsets:
                feats = rng.choice(F, size=max(1, (2 * F) // 3), replace=False).tolist()
            todo.append((k, t, nd.id, feats))
        it = IterIn(cfg, [(t, roots[t]) for t in range(n_trees)], todo)
        d = dev.compute(it)
        o = orc.compute(it, n_threads=oracle_threads)
        n_iter += 1
        for a, b in zip(d[:4], o[:4]):
            assert abs(a - b) <= 1e-9 * abs(b) + 1e-12, (d[:4], o[:4])
        scale = (float(np.sum(np.abs((1.0 if sig is None else sig) * w.max(axis=0)))),) * 3
        assert_stats_close(cfg, it, d[4], o[4], scale, exact_counts=not weighted)
        gd, bits_d = dev.best_splits(len(todo), with_cat=has_cat)
        go, bits_o = orc.best_splits(it, d[4], with_cat=has_cat)
        go2, _ = orc.best_splits(it, o[4], with_cat=has_cat)
        nxt = []
        for k, (t, nd) in enumerate(todo_nodes):
            assert_gain_equal(gd[k], go[k], bits_d, bits_o, k)
            assert_same_split(gd[k], go2[k])
            g = gd[k]
            if not g.found:
                continue
            level = index_to_level(nd.id)
            nd.predict = g.predict
            if g.gain <= 0 or level == max_depth:
                continue
            left, right = apply_gain(cfg, nd, g, bits_d, k)
            if not (level + 1 == max_depth or is_pos_zero(g.left_impurity)):
                nxt.append((t, left))
            if not (level + 1 == max_depth or is_pos_zero(g.right_impurity)):
                nxt.append((t, right))
        todo_nodes = nxt
    dev.close()
    orc.close()
    return n_iter


def test_rf_gini_poisson_bags():
    assert run_rf(12000, [32] * 40, seed=21, n_trees=4, max_depth=5) >= 4


def test_rf_entropy_weighted_multiclass():
    run_rf(9000, [16] * 20 + [8] * 5, seed=22, n_trees=3, max_depth=4, impurity=capi.SB_IMPURITY_ENTROPY,
           num_classes=3, weighted=True)


def test_rf_gini_categorical():
    run_rf(8000, [32] * 5 + [13, 41, 7], seed=23, n_trees=3, max_depth=4,
           is_categorical=[0] * 5 + [1, 1, 1], feature_subsets=False)


def test_synthetic_generator_matches_host_twin():
    from shifu_b200 import synthetic
    n_bins = [64, 32, 7, 200, 1001]
    cfg = ConfigArrays(n_bins)
    w = Worker(cfg)
    w.load_synthetic(10000, seed=77, row0=12345)
    cols, y, sig = synthetic.make(77, n_bins, 12345, 10000)
    for f in range(5):
        assert np.array_equal(w.read_bins(f, 0, 10000), cols[f])
    lab, s = w.read_labels(0, 10000)
    assert np.mean(lab != y) < 1e-3 and np.array_equal(s, sig)
    w.close()


def test_large_shape_size_independent_properties():
    """At a size the CPU oracle cannot finish quickly (8M rows x 200 features x 64 bins, the C2 column shape): properties
    that hold at any size.  (1) every feature's per-bin counts of a node add up to the node's row count, and sums /
    sums of squares add up to the same totals for every feature; (2) the two children of a split add up to the parent,
    bin by bin, exactly (integer accumulators); (3) a second run gives the same bits (order independence)."""
    n, F, B = 8_000_000, 200, 64
    from shifu_b200.capi import NativeMaster
    runs = []
    for rep in range(2):
        cfg = ConfigArrays([B] * F, max_depth=4, min_instances_per_node=5)
        w = Worker(cfg)
        w.load_synthetic(n, seed=1002)
        m = NativeMaster(cfg, tree_num=1, max_depth=4)
        it = m.build_initial_master_params()
        levels = []
        while it is not None:
            res = w.compute(it)
            levels.append(res[4].reshape(it.n_todo, F, B, 3).copy())
            gains, bits = w.best_splits(it.n_todo)
            it = m.do_compute(gains, bits)
        runs.append(levels)
        w.close()
    root = runs[0][0][0]
    assert np.all(root[:, :, 0].sum(axis=1) == float(n))                       # (1) counts
    assert np.all(root[:, :, 1].sum(axis=1) == root[0, :, 1].sum())            # labels are 0/1: sums are exact integers
    assert np.all(root[:, :, 2].sum(axis=1) == root[0, :, 2].sum())
    lvl2 = runs[0][1]
    assert lvl2.shape[0] == 2 and np.array_equal(lvl2[0] + lvl2[1], root)      # (2) children add up to the parent
    lvl3 = runs[0][2]
    assert np.array_equal(lvl3[0] + lvl3[1], lvl2[0]) and np.array_equal(lvl3[2] + lvl3[3], lvl2[1])
    for a, b in zip(runs[0], runs[1]):                                         # (3) reproducible bits
        assert np.array_equal(a, b)


def test_device_binning_matches_get_bin_index(oracle_lib):
    """f1: sb_bin_values against DTWorker.getBinIndex restated in the oracle, incl. +-inf, NaN, duplicate boundaries,
    values on a boundary, a single-boundary column, and the reference's own cancer-judgement boundaries."""
    import ctypes as C
    import os
    import torch
    lib = capi.load_library()
    rng = np.random.default_rng(61)
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cancer_judgement.npz"))
    cases = [np.array([-np.inf, 1.0, 1.0, 2.0, 5.5]), np.array([-np.inf]), np.array([0.0, 0.25, 0.5]),
             z["boundaries"][0], np.concatenate([[-np.inf], np.sort(rng.normal(size=300))])]
    for bb in cases:
        bb = np.ascontiguousarray(bb, dtype=np.float64)
        v = rng.normal(size=5000).astype(np.float32) * 3
        v[:8] = [np.inf, -np.inf, np.nan, 1.0, 2.0, 0.25, -1e30, 1e30]
        v[8:8 + min(len(bb), 100)] = bb[:100].astype(np.float32)
        width = 2 if len(bb) > 256 else 1
        d_v = torch.from_numpy(v).cuda()
        d_out = torch.zeros(len(v), dtype=torch.int16 if width == 2 else torch.uint8, device="cuda")
        capi.check(lib, lib.sb_bin_values(d_v.data_ptr(), len(v), bb.ctypes.data, len(bb), d_out.data_ptr(), width, None))
        torch.cuda.synchronize()
        got = d_out.cpu().numpy().astype(np.int64)
        vv = np.where(np.isnan(v), np.float32(0), v)
        want = np.array([oracle_lib.orc_get_bin_index(C.c_float(float(x)), bb.ctypes.data, len(bb)) for x in vv])
        want = np.where(want < 0, -1 if width == 2 else 255, want)
        assert np.array_equal(got, want)


def test_device_bagging_mask_is_java_random():
    """f4: subsampleWeights drawn on the device are `new java.util.Random(seed).nextDouble() <= rate`, row by row."""
    from java_random import JavaRandom
    n = 70001
    cfg = ConfigArrays([8] * 3)
    w = Worker(cfg)
    w.load_synthetic(n, seed=5)
    for seed, rate, first in ((42, 0.7, 0), (-123456789012, 0.25, 1000003)):
        w.sample_weights(0, seed, rate, first)
        _, _, _, got = w.read_rows(n)
        r = JavaRandom(seed)
        for _ in range(first):
            r.next_double()
        want = np.array([1.0 if r.next_double() <= rate else 0.0 for _ in range(n)], dtype=np.float32)
        assert np.array_equal(got, want)
    w.close()


def test_device_stratified_mask_is_one_java_random_per_class():
    """f4, isStratifiedSampling: one Random(seed) per class value, so a row gets the k-th nextDouble(), k = earlier rows
    of its class (dt/DTWorker.java:1526-1532)."""
    from java_random import JavaRandom
    n, n_bins = 50021, [8] * 3
    rng = np.random.default_rng(17)
    cols = [rng.integers(0, 8, size=n).astype(np.uint8) for _ in n_bins]
    y = rng.choice([0.0, 1.0, 2.0], size=n, p=[0.6, 0.3, 0.1]).astype(np.float32)
    cfg = ConfigArrays(n_bins)
    w = Worker(cfg)
    w.load_rows(cols, y, None, None)
    seed, rate = 20240517, 0.35
    w.sample_weights_stratified(0, seed, rate, 3)
    _, _, _, got = w.read_rows(n)
    r = JavaRandom(seed)
    draws = np.array([r.next_double() for _ in range(n)])
    want = np.zeros(n, dtype=np.float32)
    for c in range(3):
        idx = np.flatnonzero((y + np.float32(0.01)).astype(np.int32) == c)
        want[idx] = (draws[:len(idx)] <= rate).astype(np.float32)
    assert np.array_equal(got, want)
    with pytest.raises(RuntimeError):
        w.sample_weights_stratified(0, seed, rate, 2)   # class 2 exists
    w.close()


def test_device_poisson_bags_match_the_host_twin_and_the_distribution():
    """f4, RF bag counts: counter-based, so two shards draw what one handle draws; Poisson(rate) in distribution."""
    from shifu_b200 import synthetic
    n = 400003
    cfg = ConfigArrays([8] * 3, is_gbdt=False, tree_num=3, weights_per_row=3, impurity=capi.SB_IMPURITY_GINI, num_classes=2)
    w = Worker(cfg)
    w.load_synthetic(n, seed=5, label_kind=1)
    for t, rate in ((0, 1.0), (2, 0.5)):
        w.sample_poisson(t, 77, rate, row0=1000)
        _, _, _, got = w.read_rows(n, w_index=t)
        assert np.array_equal(got, synthetic.poisson_bags(77, t, rate, 1000, n))
        assert abs(got.mean() - rate) < 0.01 and abs(got.var() - rate) < 0.02
        assert abs((got == 0).mean() - np.exp(-rate)) < 0.005
    half = Worker(cfg)
    half.load_synthetic(n // 2, seed=5, label_kind=1)
    half.sample_poisson(0, 77, 1.0, row0=1000 + (n - n // 2))
    _, _, _, tail = half.read_rows(n // 2, w_index=0)
    _, _, _, full = w.read_rows(n, w_index=0)
    assert np.array_equal(tail, full[n - n // 2:])
    half.close()
    w.close()


def test_c0_cancer_judgement_full_train_and_model_file(tmp_path):
    """Config C0: the reference's own example (429 rows x 30 numeric x 10 bins, binned with its ColumnConfig.json), GBT
    10 trees depth 6 squared loss lr 0.05, all features.  Every iteration's statistics, scalars and chosen splits are
    checked against the oracle; the grown trees are then written in the reference's model format, read back, and the
    model's scores on the RAW feature values must equal the device's Data.predict."""
    import os
    from oracle.pyoracle import OracleWorker
    from shifu_b200 import model_io
    from shifu_b200.capi import NativeMaster
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cancer_judgement.npz"))
    bins, label, raw = z["bins"], z["label"], z["raw"]
    n = len(label)
    cfg = ConfigArrays([10] * 30, column_num=z["column_num"], bin_boundary=z["boundaries"].ravel(), max_depth=6,
                       tree_num=10, learning_rate=0.05, min_instances_per_node=1)
    dev, orc = Worker(cfg), OracleWorker(cfg)
    dev.load_rows(list(bins), label)
    orc.load_rows(list(bins), label)
    m = NativeMaster(cfg, tree_num=10, max_depth=6, learning_rate=0.05)
    it = m.build_initial_master_params()
    n_iter, errs = 0, []
    while it is not None:
        d = dev.compute(it)
        o = orc.compute(it, n_threads=4)
        assert np.array_equal(d[4].reshape(-1, 3)[:, 0], o[4].reshape(-1, 3)[:, 0])
        _, dout, _, _ = dev.read_rows(n)
        # sums of residuals cancel: the yardstick is the sum of |contributions| (two fp64 summation orders of the same
        # numbers already differ by ~1e-16 of it), see helpers.rel_close
        assert rel_close(d[4].reshape(-1, 3)[:, 1], o[4].reshape(-1, 3)[:, 1], scale=float(np.abs(dout).sum())).all()
        assert rel_close(d[4].reshape(-1, 3)[:, 2], o[4].reshape(-1, 3)[:, 2], scale=float((dout * dout).sum())).all()
        for a, b in zip(d[:4], o[:4]):
            assert abs(a - b) <= 1e-9 * abs(b) + 1e-12
        gd, _ = dev.best_splits(it.n_todo)
        go, _ = orc.best_splits(it, d[4])
        go2, _ = orc.best_splits(it, o[4])
        for k in range(it.n_todo):
            assert_gain_equal(gd[k], go[k])
            assert_same_split(gd[k], go2[k])
        errs.append(d[2] / d[0])
        it = m.do_compute(gd, None)
        n_iter += 1
    assert m.n_trees == 10 and n_iter >= 30
    assert errs[-1] < errs[5] < 0.25  # the training error goes down tree after tree
    # the tenth tree is grown but, as in the reference, never applied to Data.predict (training halts, dt/DTMaster.java:396)
    trees = []
    for t in range(10):
        nodes, pool, extras = m.export_tree(t)
        trees.append(model_io.tree_from_export(nodes, pool, extras, z["column_num"], t, 1.0 if t == 0 else 0.05))
    model = {"version": 4, "algorithm": "GBT", "loss": "squared", "is_classification": False, "is_one_vs_all": False,
             "input_count": 30, "numerical_mean": [], "column_names": [(int(c), "column_%d" % c) for c in z["column_num"]],
             "categories": [], "column_mapping": [(int(c), i) for i, c in enumerate(z["column_num"])], "bags": [trees]}
    path = tmp_path / "model0.gbt"
    model_io.write_model(model, path)
    back = model_io.read_model(path)
    assert back["trailing"] == 0 and len(back["bags"][0]) == 10
    back["bags"][0] = back["bags"][0][:9]
    predict, _, _, _ = dev.read_rows(n)
    scores = np.array([model_io.score_gbt(back, raw[:, r]) for r in range(n)])
    assert np.allclose(scores, predict, rtol=2e-5, atol=2e-6)
    dev.close()
    orc.close()


def _export_all(master):
    out = []
    for t in range(master.n_trees):
        nodes, pool, extras = master.export_tree(t)
        out.append((b"".join(bytes(x) for x in nodes), pool.tobytes(), extras.tobytes()))
    return out


def test_native_tree_loop_and_single_rank_communicator():
    """sbh_grow_tree (the iteration loop of a tree in native code) grows the trees of the call-by-call loop, and a
    one-rank NCCL communicator inside the library (sb_dist_init) changes nothing."""
    from shifu_b200.capi import NativeMaster, dist_unique_id
    n_bins = [32] * 24 + [7] * 3
    is_cat = [False] * 24 + [True] * 3
    grown = []
    for mode in ("calls", "native", "native+comm"):
        cfg = ConfigArrays(n_bins, is_categorical=is_cat, max_depth=5, tree_num=3, min_instances_per_node=2, learning_rate=0.1)
        w = Worker(cfg)
        w.load_synthetic(60000, seed=21, row0=0, weighted=True)
        if mode == "native+comm":
            w.dist_init(dist_unique_id(), 0, 1)
        m = NativeMaster(cfg, tree_num=3, max_depth=5, learning_rate=0.1)
        it = m.build_initial_master_params()
        passes = 0
        while it is not None:
            if mode == "calls":
                w.compute_begin(it)
                w.compute_stats()
                _, gains, bits = w.compute_finish_split(it.n_todo, with_cat=True)
                it = m.do_compute(gains, bits)
                passes += 1
            else:
                it, p, scal = m.grow_tree(w, it)
                passes += p
                assert scal[0] > 0
        grown.append((passes, _export_all(m)))
        m.close()
        w.close()
    assert grown[0][0] == grown[1][0] == grown[2][0] >= 9
    assert grown[0][1] == grown[1][1] == grown[2][1]


def test_feature_counts_cover_every_tile_shape():
    """The last feature tile holds 1..4 features per lane (template parameter KACT of hist_rows_kernel): 27 -> 1,
    40 -> 2, 80 -> 3, 128 -> 4 exactly full, 130 -> a full tile plus a 2-feature tile, 200 -> 4 + 3."""
    for F, seed in ((27, 51), (80, 52), (128, 53), (130, 54), (200, 55)):
        run_gbt(6000, [16] * F, seed=seed, n_trees=2, max_depth=3, weighted=(F % 2 == 0), with_validation=False)
    run_gbt(6000, [16] * 80, seed=56, n_trees=1, max_depth=4, feature_subsets=True, with_validation=False)  # per-cell masks


def test_worker_that_joins_a_tree_midway_routes_from_the_root():
    """A handle that first sees a tree at its third level (the message carries the whole tree, as always) must walk its
    rows from the root: the statistics equal those of the handle that saw every pass."""
    n, n_bins = 30000, [32] * 24
    cols, y, sig, _ = make_data(n, n_bins, seed=91, weighted=True)
    cfg = ConfigArrays(n_bins, max_depth=5, min_instances_per_node=1, tree_num=2)
    a = Worker(cfg)
    a.load_rows(cols, y, sig, None)
    root = Node(1)
    todo_nodes, switch, level = [root], True, 1
    while todo_nodes:
        todo = [(k, 0, nd.id, []) for k, nd in enumerate(todo_nodes)]
        it = IterIn(cfg, [(0, root)], todo, switch_to_next_tree=switch, is_first_tree=True)
        ra = a.compute(it)
        if level >= 3:  # a fresh handle sees only this message
            b = Worker(cfg)
            b.load_rows(cols, y, sig, None)
            rb = b.compute(it)
            assert np.array_equal(ra[4], rb[4]) and ra[0] == rb[0]
            b.close()
        gains, bits = a.best_splits(len(todo))
        nxt = []
        for k, nd in enumerate(todo_nodes):
            g = gains[k]
            if not g.found or g.gain <= 0 or index_to_level(nd.id) == 5:
                continue
            nd.predict = g.predict
            left, right = apply_gain(cfg, nd, g, bits, k)
            if index_to_level(nd.id) + 1 < 5:
                nxt += [left, right]
        todo_nodes, switch, level = nxt, False, level + 1
    assert level >= 4
    a.close()
