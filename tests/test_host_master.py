"""The product's DTMaster (csrc/host_master.cpp, sbh_* entry points) against the oracle's restatement of
dt/DTMaster.java (oracle/oracle_master.cpp: doCompute, the two split routines, buildInitialMasterParams,
getSubsamplingFeatures, java.util.PriorityQueue / LinkedList / HashMap order) on the same stream of gain records.
No GPU needed: both are host logic.  Every message (trees, masks, todo nodes, feature lists, flags) and the final
trees must agree."""
import numpy as np
import pytest

from shifu_b200 import capi
from shifu_b200.capi import ConfigArrays, NativeMaster, SbGain

from java_random import JavaRandom


@pytest.fixture(scope="module", autouse=True)
def built():
    import __graft_entry__ as g
    g.build()


def fake_gains(rng, cfg, n_todo, feats_per_node, p_found=0.9, p_zero_imp=0.1, tie_gains=False):
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
        if tie_gains:  # equal priorities: the order then comes from the heap, not from the values
            g.gain = float(rng.integers(1, 3))
        g.impurity, g.predict = float(rng.random()), float(rng.normal())
        g.left_impurity = 0.0 if rng.random() < p_zero_imp else float(rng.random())
        g.right_impurity = 0.0 if rng.random() < p_zero_imp else float(rng.random())
        g.left_predict, g.right_predict = float(rng.normal()), float(rng.normal())
        g.wgt_cnt = float(rng.integers(10, 1000)) if not tie_gains else 100.0
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


CASES = [(True, "ALL", -1, False), (True, "HALF", -1, False), (False, "TWOTHIRDS", -1, False), (True, None, -1, False),
         (True, "ALL", 12, False), (False, "SQRT", -1, False), (False, "LOG2", 9, False), (False, "ALL", 10, True),
         (True, "AUTO", 7, True), (False, "ONETHIRD", -1, False)]


@pytest.mark.parametrize("is_gbdt,strategy,max_leaves,ties", CASES)
def test_native_master_matches_oracle_master(is_gbdt, strategy, max_leaves, ties):
    from oracle.pyoracle import OracleMaster
    rng = np.random.default_rng(11 if max_leaves > 0 else 7)
    n_bins = [16] * 9 + [1, 2] + [6, 21]
    is_cat = [0] * 11 + [1, 1]
    cfg = ConfigArrays(n_bins, is_categorical=is_cat, is_gbdt=is_gbdt, tree_num=4, max_depth=5,
                       weights_per_row=1 if is_gbdt else 4)
    all_features = [f for f in range(len(n_bins)) if (not is_cat[f] and n_bins[f] > 1) or (is_cat[f] and n_bins[f] - 1 > 0)]
    kw = dict(tree_num=4, max_depth=5, is_gbdt=is_gbdt, learning_rate=0.1, feature_subset_strategy=strategy,
              feature_subset_rate=0.4, feature_sampling_seed=99, max_leaves=max_leaves, max_stats_memory=40000,
              worker_number=2)
    orc, nat = OracleMaster(cfg, **kw), NativeMaster(cfg, **kw)
    a, b = orc.build_initial_master_params(), nat.build_initial_master_params()
    n_iter = 0
    while a is not None:
        assert b is not None
        assert msg_signature(a) == msg_signature(b)
        feats = []
        for i in range(a.n_todo):
            td = a.struct.todo[i]
            fl = [td.features[k] for k in range(td.n_features)]
            feats.append(fl if fl else all_features)
        gains, bits = fake_gains(rng, cfg, a.n_todo, feats, p_zero_imp=0.1 if n_iter > 1 else 0.0,
                                 p_found=0.9 if n_iter > 1 else 1.0, tie_gains=ties)
        a = orc.do_compute(gains, bits)
        b = nat.do_compute(gains, bits)
        n_iter += 1
        assert n_iter < 500
    assert b is None and nat.n_trees == orc.n_trees and n_iter >= 3
    for t in range(nat.n_trees):
        nodes, pool, extras = nat.export_tree(t)
        dump = orc.dump_tree(t)
        assert [n.id for n in nodes] == [int(x) for x in dump[:, 0]]
        # gain, impurity, leftImpurity, rightImpurity, leftPredict, rightPredict, wgtCnt
        assert np.array_equal(extras, dump[:, 4:11])
    nat.close()
    orc.close()


def test_oracle_feature_sampling_is_the_reservoir_of_the_reference():
    """sampleFeaturesForNodeStats (dt/DTMaster.java:880-893) re-derived here from java.util.Random: the first `sample`
    features, then feature i replaces slot (int)(nextDouble() * i) when that is below `sample`."""
    from oracle.pyoracle import OracleMaster
    n_bins = [8] * 20
    cfg = ConfigArrays(n_bins, is_gbdt=True, tree_num=3, max_depth=3)
    orc = OracleMaster(cfg, tree_num=3, max_depth=3, feature_subset_strategy="HALF", feature_sampling_seed=1234)
    msg = orc.build_initial_master_params()
    got = [msg.struct.todo[0].features[k] for k in range(msg.struct.todo[0].n_features)]
    r = JavaRandom(1234)
    want = list(range(10))
    for i in range(10, 20):
        j = int(r.next_double() * i)
        if 0 <= j < 10:
            want[j] = i
    assert got == want
    orc.close()
