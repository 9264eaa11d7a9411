"""The C++ DTMaster mirror (csrc/host_master.cpp, sbh_* entry points) against the Python mirror (dt_master.py) on the
same stream of gain records -- no GPU needed: both are pure host logic (queue, level-/leaf-wise growth, feature
sub-sampling with java.util.Random, GBT tree switch, halt rules; dt/DTMaster.java:274-537, 543-605, 759-893)."""
import ctypes as C

import numpy as np
import pytest

from shifu_b200 import capi
from shifu_b200.capi import ConfigArrays, NativeMaster, SbGain
from shifu_b200.dt_master import DTMaster


@pytest.fixture(scope="module", autouse=True)
def built():
    import __graft_entry__ as g
    g.build()


def fake_gains(rng, cfg, n_todo, feats_per_node, p_found=0.9, p_zero_imp=0.1):
    gains = (SbGain * max(n_todo, 1))()
    bits = np.zeros(max(n_todo, 1) * capi.SB_MAX_CAT_BYTES, dtype=np.uint8)
    for i in range(n_todo):
        g = gains[i]
        if rng.random() > p_found:
            continue
        feats = feats_per_node[i]
        f = int(feats[rng.integers(len(feats))])
        g.found, g.feature = 1, f
        nb = int(cfg.n_bins[f])
        if cfg.is_categorical[f]:
            g.split_type = capi.SB_SPLIT_CATEGORICAL
            g.is_left = int(rng.integers(2))
            g.cat_bytes = nb // 8 + 1
            bits[i * capi.SB_MAX_CAT_BYTES: i * capi.SB_MAX_CAT_BYTES + g.cat_bytes] = rng.integers(0, 256, g.cat_bytes)
            g.split_bin = int(rng.integers(1, nb))
        else:
            g.split_type = capi.SB_SPLIT_CONTINUOUS
            g.split_bin = int(rng.integers(0, nb - 1))
        g.gain = float(rng.random()) if rng.random() > 0.05 else 0.0
        g.impurity, g.predict = float(rng.random()), float(rng.normal())
        g.left_impurity = 0.0 if rng.random() < p_zero_imp else float(rng.random())
        g.right_impurity = 0.0 if rng.random() < p_zero_imp else float(rng.random())
        g.left_predict, g.right_predict = float(rng.normal()), float(rng.normal())
        g.wgt_cnt = float(rng.integers(10, 1000))
    return gains, bits


def msg_signature(m):
    """Everything the worker sees in a message."""
    s = m.struct
    sig = [s.n_trees, s.n_todo, s.switch_to_next_tree, s.is_first_tree]
    for t in range(s.n_trees):
        tr = s.trees[t]
        sig.append((tr.tree_id, tr.n_nodes, tr.mask_pool_bytes))
        for i in range(tr.n_nodes):
            n = tr.nodes[i]
            mask = bytes(tr.mask_pool[n.left_mask_offset + k] for k in range(n.left_mask_bytes))
            cats = bytes(tr.mask_pool[n.cat_offset + k] for k in range(n.cat_bytes))
            sig.append((n.id, n.left, n.right, n.feature, n.split_type if n.feature >= 0 else 0, mask, n.has_predict,
                        n.predict, n.threshold if n.feature >= 0 else 0.0, n.is_left, cats))
    for i in range(s.n_todo):
        td = s.todo[i]
        sig.append((td.key, td.tree_id, td.node_id, tuple(td.features[k] for k in range(td.n_features))))
    return sig


@pytest.mark.parametrize("is_gbdt,strategy,max_leaves", [(True, "ALL", -1), (True, "HALF", -1), (False, "TWOTHIRDS", -1),
                                                         (True, None, -1), (True, "ALL", 12), (False, "SQRT", -1)])
def test_native_master_matches_python_mirror(is_gbdt, strategy, max_leaves):
    rng = np.random.default_rng(11 if max_leaves > 0 else 7)
    n_bins = [16] * 9 + [1, 2] + [6, 21]
    is_cat = [0] * 11 + [1, 1]
    cfg = ConfigArrays(n_bins, is_categorical=is_cat, is_gbdt=is_gbdt, tree_num=4, max_depth=5,
                       weights_per_row=1 if is_gbdt else 4)
    kw = dict(tree_num=4, max_depth=5, is_gbdt=is_gbdt, learning_rate=0.1, feature_subset_strategy=strategy,
              feature_subset_rate=0.4, feature_sampling_seed=99, max_leaves=max_leaves, max_stats_memory=40000,
              worker_number=2)
    py, nat = DTMaster(cfg, **kw), NativeMaster(cfg, **kw)
    a, b = py.build_initial_master_params(), nat.build_initial_master_params()
    n_iter = 0
    while a is not None:
        assert b is not None
        assert msg_signature(a) == msg_signature(b)
        feats = []
        for i in range(a.n_todo):
            td = a.struct.todo[i]
            fl = [td.features[k] for k in range(td.n_features)]
            feats.append(fl if fl else py.all_features)
        gains, bits = fake_gains(rng, cfg, a.n_todo, feats, p_zero_imp=0.1 if n_iter > 1 else 0.0, p_found=0.9 if n_iter > 1 else 1.0)
        a = py.do_compute(list(gains)[:a.n_todo], bits)
        b = nat.do_compute(gains, bits)
        n_iter += 1
        assert n_iter < 500
    assert b is None and nat.n_trees == len(py.trees) and n_iter >= 4
    nodes, pool, extras = nat.export_tree(0)
    py_nodes = list(py.trees[0].node.walk())
    assert [n.id for n in nodes] == [n.id for n in py_nodes]
    assert np.allclose(extras[:, 0], [n.gain for n in py_nodes])
    nat.close()
