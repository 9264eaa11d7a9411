"""Host-side mirror of dt/DTMaster.java: the tree-growth state machine around the device path.

What stays on the host is what the reference keeps in the master's scalar logic: the todo queue, level-wise /
leaf-wise growth, per-node feature sub-sampling, the GBT tree switch and the halt rules
(dt/DTMaster.java:274-537, 543-605, 710-727, 759-893).  What moved to the GPU is the merge of worker statistics
(now an integer all-reduce) and Impurity.computeImpurity + max-gain per node (sb_best_splits): `do_compute` below
consumes sb_gain records instead of NodeStats maps.

Method and field names follow the Java class so the two can be read side by side.
"""
import math
from collections import deque

from .capi import SB_MAX_CAT_BYTES, SB_SPLIT_CONTINUOUS
from .messages import IterIn, Node

# core/dtrain/FeatureSubsetStrategy.java:23-31
ALL, HALF, ONETHIRD, TWOTHIRDS, AUTO, SQRT, LOG2 = "ALL", "HALF", "ONETHIRD", "TWOTHIRDS", "AUTO", "SQRT", "LOG2"


class JavaRandom:
    """java.util.Random (48-bit LCG): seeded runs of the reference draw exactly these numbers."""

    MASK = (1 << 48) - 1

    def __init__(self, seed):
        self.seed = (seed ^ 0x5DEECE66D) & self.MASK

    def next(self, bits):
        self.seed = (self.seed * 0x5DEECE66D + 0xB) & self.MASK
        r = self.seed >> (48 - bits)
        if r >= 1 << (bits - 1):
            r -= 1 << bits
        return r

    def next_double(self):
        a = self.next(26) & ((1 << 26) - 1)
        b = self.next(27) & ((1 << 27) - 1)
        return ((a << 27) + b) * (1.0 / (1 << 53))


def index_to_level(node_index):  # dt/Node.java:459-461
    return int(node_index).bit_length()


def _is_pos_zero(x):  # Double.compare(x, 0d) == 0
    return x == 0.0 and math.copysign(1.0, x) > 0


class TreeNode:
    """dt/TreeNode.java:41-104: a tree (or a todo node of it) with its per-node feature subset."""

    def __init__(self, tree_id, node, learning_rate=1.0):
        self.tree_id = tree_id
        self.node = node
        self.learning_rate = learning_rate
        self.features = []
        self.node_num = 1
        self.root_wgt_cnt = 0.0


class DTMaster:
    def __init__(self, cfg, tree_num, max_depth, is_gbdt=True, learning_rate=0.05, feature_subset_strategy=ALL,
                 feature_subset_rate=1.0, feature_sampling_seed=0, max_leaves=-1, max_batch_split_size=32,
                 max_stats_memory=256 << 20, worker_number=1, all_features=None):
        self.cfg = cfg
        self.tree_num = tree_num
        self.max_depth = max_depth
        self.is_gbdt, self.is_rf = bool(is_gbdt), not is_gbdt
        self.learning_rate = learning_rate
        self.feature_subset_strategy = feature_subset_strategy
        self.feature_subset_rate = feature_subset_rate
        # the reference uses an unseeded Random (dt/DTMaster.java:246); a seed is injected for reproducibility
        self.feature_sampling_random = JavaRandom(feature_sampling_seed)
        self.max_leaves = max_leaves
        self.is_leaf_wise = max_leaves > 0  # dt/DTMaster.java:980-982
        self.max_batch_split_size = max_batch_split_size
        self.max_stats_memory = max_stats_memory
        self.worker_number = worker_number
        F = len(cfg.n_bins)
        if all_features is None:  # getAllFeatureList (dt/DTMaster.java:852-878), as feature slots
            all_features = [f for f in range(F)
                            if (not cfg.is_categorical[f] and cfg.n_bins[f] > 1) or (cfg.is_categorical[f] and cfg.n_bins[f] - 1 > 0)]
        self.all_features = list(all_features)
        self.input_num = F
        self.to_do_queue = deque()
        self.to_split_queue = []
        self.trees = []
        if self.is_rf:
            for i in range(tree_num):  # :1078-1080
                self.trees.append(TreeNode(i, Node(1), 1.0))
        else:
            self.trees.append(TreeNode(0, Node(1), 1.0))  # learning rate 1 for the first tree (:1108)
        self.todo_nodes = []  # list of (key, TreeNode)
        self.halt = False
        self.switch_to_next_tree = False
        self.is_first_tree = False
        self.iteration = 0
        self.stats_size = cfg.stats_size

    @property
    def n_trees(self):
        return len(self.trees)

    # ---- feature sub-sampling, dt/DTMaster.java:822-850, 880-893
    def get_subsampling_features(self):
        n = len(self.all_features)
        s = self.feature_subset_strategy
        if s is None:
            return self._sample(int(n * self.feature_subset_rate))
        if s == HALF:
            return self._sample(n // 2)
        if s == ONETHIRD:
            return self._sample(n // 3)
        if s == TWOTHIRDS:
            return self._sample(n * 2 // 3)
        if s == SQRT:
            return self._sample(int(n * math.sqrt(self.input_num) / self.input_num))
        if s == LOG2:
            return self._sample(int(n * math.log(self.input_num) / math.log(2) / self.input_num))
        if s == AUTO:
            return self._sample(n // 2) if self.tree_num > 1 else []
        return []  # ALL: empty list == every valid feature

    def _sample(self, sample):
        feats = list(self.all_features[:sample])
        for i in range(sample, len(self.all_features)):
            j = int(self.feature_sampling_random.next_double() * i)
            if 0 <= j < sample:
                feats[j] = self.all_features[i]
        return feats

    def get_stats_mem(self, subset):  # :729-747
        feats = subset if len(subset) else self.all_features
        mem = 0
        for f in feats:
            mem += int(self.cfg.n_bins[f]) * self.stats_size * 8 * 2
        return mem * self.worker_number // 2

    # ---- iteration 0, dt/DTMaster.java:759-820
    def build_initial_master_params(self):
        self.todo_nodes = []
        if self.is_rf:
            curr_mem = 0
            key = 0
            for tn in self.trees:
                tn.features = self.get_subsampling_features()
                curr_mem += self.get_stats_mem(tn.features)
                if curr_mem < self.max_stats_memory:
                    self.todo_nodes.append((key, tn))
                    key += 1
                else:
                    self.to_do_queue.append(tn)
            self.switch_to_next_tree = False
            self.is_first_tree = False
        else:
            tn = self.trees[-1]
            tn.features = self.get_subsampling_features()
            self.todo_nodes.append((0, tn))
            self.switch_to_next_tree = True
            self.is_first_tree = len(self.trees) == 1
        return self._message()

    def _message(self, new_weights=None, dropout=None):
        if self.is_gbdt:
            sent = self.trees[-2:] if self.switch_to_next_tree and len(self.trees) >= 2 else self.trees[-1:]
            trees = [(t.tree_id, t.node) for t in sent]
        else:
            trees = [(t.tree_id, t.node) for t in self.trees]
        todo = [(key, tn.tree_id, tn.node.id, list(tn.features)) for key, tn in self.todo_nodes]
        return IterIn(self.cfg, trees, todo, switch_to_next_tree=self.switch_to_next_tree,
                      is_first_tree=self.is_first_tree, new_weights=new_weights, dropout=dropout)

    # ---- dt/DTMaster.java:710-727
    def _populate_gain_info_to_node(self, tree_id, node, g, cat_bits, idx):
        node.predict = g.predict
        node.feature = g.feature
        node.split_type = g.split_type
        if g.split_type == SB_SPLIT_CONTINUOUS:
            node.threshold = float(self.cfg.boundaries_of(g.feature)[g.split_bin + 1])  # dt/Impurity.java:200
        else:
            node.is_left = bool(g.is_left)
            node.cat_bits = bytes(cat_bits[idx * SB_MAX_CAT_BYTES: idx * SB_MAX_CAT_BYTES + g.cat_bytes])
        node.gain = g.gain
        node.impurity = g.impurity
        node.left_impurity, node.right_impurity = g.left_impurity, g.right_impurity
        node.left_predict, node.right_predict = g.left_predict, g.right_predict
        node.wgt_cnt = g.wgt_cnt
        if node.id == 1:
            self.trees[tree_id].root_wgt_cnt = g.wgt_cnt

    # ---- dt/DTMaster.java:568-605
    def _split_node_for_level_wised_tree(self, is_leaf, tree_id, node):
        if is_leaf:
            return
        level = index_to_level(node.id)
        for side in (0, 1):
            imp = node.left_impurity if side == 0 else node.right_impurity
            pred = node.left_predict if side == 0 else node.right_predict
            child_is_leaf = level + 1 == self.max_depth or _is_pos_zero(imp)
            child = Node(2 * node.id + side, predict=pred)
            child.impurity = imp
            if side == 0:
                node.left = child
            else:
                node.right = child
            self.trees[tree_id].node_num += 1
            if not child_is_leaf:
                self.to_do_queue.append(TreeNode(tree_id, child))

    # ---- dt/DTMaster.java:543-562
    def _split_node_for_leaf_wised_tree(self, tree_id, node):
        is_over_max_leaves = self.trees[tree_id].node_num + 1 > self.max_leaves
        can_split = (not is_over_max_leaves) and not _is_pos_zero(node.left_impurity)
        if can_split:
            for side in (0, 1):
                imp = node.left_impurity if side == 0 else node.right_impurity
                pred = node.left_predict if side == 0 else node.right_predict
                child = Node(2 * node.id + side, predict=pred)
                child.impurity = imp
                if side == 0:
                    node.left = child
                else:
                    node.right = child
                self.trees[tree_id].node_num += 1
                self.to_do_queue.append(TreeNode(tree_id, child))

    # ---- dt/DTMaster.java:316-537, with the merged statistics already reduced to gain records on the device
    def do_compute(self, gains, cat_bits=None, new_weights=None, dropout=None):
        """gains[i] is the sb_gain of todo node i (key order).  Returns the next message, or None on halt."""
        self.iteration += 1
        for idx, (key, tn) in enumerate(self.todo_nodes):
            g = gains[idx]
            tree_id = tn.tree_id
            done = tn.node
            if not g.found:
                done.is_leaf = True  # :335-339
                continue
            self._populate_gain_info_to_node(tree_id, done, g, cat_bits, idx)
            if self.is_leaf_wise:
                if not (g.gain <= 0.0):
                    # priority = wgtCntRatio * gain (:1064-1070)
                    ratio = 1.0 if done.id == 1 else g.wgt_cnt / self.trees[tree_id].root_wgt_cnt
                    self.to_split_queue.append((ratio * g.gain, len(self.to_split_queue), tree_id, done))
            else:
                is_leaf = g.gain <= 0.0 or index_to_level(done.id) == self.max_depth
                done.is_leaf = is_leaf
                self._split_node_for_level_wised_tree(is_leaf, tree_id, done)
        if self.is_leaf_wise:
            # the reference never increments currSplitIndex (:359-363): the whole queue is split, best first
            self.to_split_queue.sort(key=lambda e: (-e[0], e[1]))
            for _, _, tree_id, nd in self.to_split_queue:
                self._split_node_for_leaf_wised_tree(tree_id, nd)
            self.to_split_queue = []

        self.todo_nodes = []
        if not self.to_do_queue:
            if self.is_gbdt:
                tree = self.trees[-1]
                if len(self.trees) >= self.tree_num:
                    self.halt = True  # :396-399
                elif tree.node.left is None and tree.node.right is None:
                    self.halt = True  # root never split (:400-407)
                else:
                    self.is_first_tree = len(self.trees) == 1  # :418
                    tree.features = None
                    new_root = TreeNode(len(self.trees), Node(1), self.learning_rate)
                    self.trees.append(new_root)
                    new_root.features = self.get_subsampling_features()
                    self.todo_nodes.append((0, new_root))
                    self.switch_to_next_tree = True
            else:
                self.halt = True  # :431-435
        else:
            curr_mem = 0
            key = 0
            while self.to_do_queue and curr_mem <= self.max_stats_memory:  # :450-470
                tn = self.to_do_queue.popleft()
                tn.features = self.get_subsampling_features()
                curr_mem += self.get_stats_mem(tn.features)
                self.todo_nodes.append((key, tn))
                key += 1
            self.switch_to_next_tree = False
            self.is_first_tree = len(self.trees) == 1  # :475
        if self.halt:
            return None
        return self._message(new_weights=new_weights, dropout=dropout)
