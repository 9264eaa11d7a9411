"""Builders for the flattened DTMasterParams message (sb_iter_in) from a Python-side tree model.

The tree model mirrors dt/Node.java + dt/Split.java (id, split, children, predict); `IterIn` mirrors
DTMasterParams (trees, todoNodes, switchToNextTree, isFirstTree: dt/DTMasterParams.java:44-203).  The only
computation here is the host-side encoding the C ABI asks for: each split's `left_mask`, built with the
reference's own routing rule (dt/DTWorker.java:1052-1088).
"""
import ctypes as C

import numpy as np

from .capi import (SB_SPLIT_CATEGORICAL, SB_SPLIT_CONTINUOUS, SbIterIn, SbNode, SbTodoNode, SbTree)


class Node:
    """dt/Node.java: heap-numbered node (root id 1, left 2*id, right 2*id+1)."""

    def __init__(self, nid, predict=None):
        self.id = nid
        self.feature = -1  # feature slot of the split, -1 = no split
        self.split_type = SB_SPLIT_CONTINUOUS
        self.threshold = 0.0
        self.is_left = False
        self.cat_bits = b""  # SimpleBitSet bytes of Split.leftOrRightCategories
        self.left = None
        self.right = None
        self.predict = predict  # None == Node.predict is null
        self.gain = 0.0
        self.impurity = 0.0
        self.left_impurity = 0.0
        self.right_impurity = 0.0
        self.left_predict = 0.0
        self.right_predict = 0.0
        self.wgt_cnt = 0.0
        self.is_leaf = True
        self._mask = None  # left_mask bytes, computed once the split is final (a split never changes afterwards)

    def walk(self):
        yield self
        if self.left is not None:
            yield from self.left.walk()
        if self.right is not None:
            yield from self.right.walk()


def left_mask(cfg, node):
    """bit b set <=> a row with bin index b goes left at `node` (SimpleBitSet byte order)."""
    f = node.feature
    nb = int(cfg.n_bins[f])
    mask = bytearray(nb // 8 + 1)
    if not cfg.is_categorical[f]:
        bb = cfg.boundaries_of(f)
        bits = np.zeros((nb // 8 + 1) * 8, dtype=np.uint8)
        bits[:nb] = bb < node.threshold  # dt/DTWorker.java:1057-1062
        return np.packbits(bits, bitorder="little").tobytes()
    else:
        n_cats = nb - 1
        for b in range(nb):
            idx = b if 0 <= b < n_cats else n_cats  # :1066-1072
            wi = idx >> 3
            has = wi < len(node.cat_bits) and (node.cat_bits[wi] >> (idx % 8)) & 1
            goes_left = bool(has) == bool(node.is_left)  # :1074-1087
            if goes_left:
                mask[b >> 3] |= 1 << (b % 8)
    return bytes(mask)


NODE_DTYPE = np.dtype([("id", "<i4"), ("left", "<i4"), ("right", "<i4"), ("feature", "<i4"), ("split_type", "<i4"),
                       ("left_mask_offset", "<i4"), ("left_mask_bytes", "<i4"), ("has_predict", "<i4"),
                       ("predict", "<f8"), ("threshold", "<f8"), ("is_left", "<i4"), ("cat_offset", "<i4"),
                       ("cat_bytes", "<i4"), ("reserved", "<i4")], align=True)
assert NODE_DTYPE.itemsize == C.sizeof(SbNode)


TODO_DTYPE = np.dtype([("key", "<i4"), ("tree_id", "<i4"), ("node_id", "<i4"), ("n_features", "<i4"),
                       ("features", "<u8")], align=True)
assert TODO_DTYPE.itemsize == C.sizeof(SbTodoNode)


class FlatTree:
    """An sb_tree plus the arrays it points at (sb_node records are filled through a numpy structured array)."""

    def __init__(self, cfg, tree_id, root):
        self.tree_id = tree_id
        if root is None:  # placeholder tree (root id INVALID_INDEX, dt/DTMaster.java:500-504)
            self.nodes = np.zeros(1, dtype=NODE_DTYPE)
            self.pool = np.zeros(1, dtype=np.uint8)
            n_nodes = 0
        else:
            lst = list(root.walk())
            index = {id(n): i for i, n in enumerate(lst)}
            pool = bytearray()
            recs = []
            for n in lst:
                mo = mb = co = cb = 0
                if n.feature >= 0:
                    if n._mask is None:
                        n._mask = left_mask(cfg, n)
                    mo, mb = len(pool), len(n._mask)
                    pool += n._mask
                    co, cb = len(pool), len(n.cat_bits)
                    pool += n.cat_bits
                recs.append((n.id, index[id(n.left)] if n.left is not None else -1,
                             index[id(n.right)] if n.right is not None else -1, n.feature, n.split_type, mo, mb,
                             0 if n.predict is None else 1, 0.0 if n.predict is None else n.predict, n.threshold,
                             1 if n.is_left else 0, co, cb, 0))
            self.nodes = np.array(recs, dtype=NODE_DTYPE)
            self.pool = np.frombuffer(bytes(pool) + b"\0", dtype=np.uint8).copy()
            n_nodes = len(lst)
        t = SbTree()
        t.tree_id = tree_id
        t.n_nodes = n_nodes
        t.nodes = C.cast(self.nodes.ctypes.data, C.POINTER(SbNode))
        t.mask_pool = self.pool.ctypes.data_as(C.POINTER(C.c_uint8))
        t.mask_pool_bytes = len(self.pool) - 1
        self.struct = t


class IterIn:
    """DTMasterParams as the worker consumes it.  trees: list of (tree_id, root Node or None);
    todo: list of (key, tree_id, node_id, [feature slots] or [])."""

    def __init__(self, cfg, trees, todo, switch_to_next_tree=False, is_first_tree=False, new_weights=None,
                 dropout=None):
        self.flat = [FlatTree(cfg, tid, root) for tid, root in trees]
        self.trees = (SbTree * max(len(self.flat), 1))(*[f.struct for f in self.flat])
        # sb_todo_node records through a structured array; all feature lists share one int32 pool
        flat_feats = [f for t in todo for f in t[3]]
        self.todo_feats = np.array(flat_feats if flat_feats else [0], dtype=np.int32)
        base = self.todo_feats.ctypes.data
        recs, off = [], 0
        for t in todo:
            nf = len(t[3])
            recs.append((t[0], t[1], t[2], nf, base + 4 * off))
            off += nf
        self.todo = np.array(recs if recs else [(0, 0, 0, 0, 0)], dtype=TODO_DTYPE)
        s = SbIterIn()
        s.n_trees = len(self.flat)
        s.trees = C.cast(self.trees, C.POINTER(SbTree))
        s.n_todo = len(todo)
        s.todo = C.cast(self.todo.ctypes.data, C.POINTER(SbTodoNode))
        s.switch_to_next_tree = int(bool(switch_to_next_tree))
        s.is_first_tree = int(bool(is_first_tree))
        s.continuous_running_start = 0
        self.new_weights = None if new_weights is None else np.ascontiguousarray(new_weights, dtype=np.float32)
        self.dropout = None if dropout is None else np.ascontiguousarray(dropout, dtype=np.uint8)
        if self.new_weights is not None:
            s.new_weights = self.new_weights.ctypes.data_as(C.POINTER(C.c_float))
        if self.dropout is not None:
            s.dropout = self.dropout.ctypes.data_as(C.POINTER(C.c_uint8))
        self.struct = s
        self.n_todo = len(todo)
