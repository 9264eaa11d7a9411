"""The wire format of sb_iter_in (what the JNI layer hands over), pinned from plain C: tests/wire_check.c writes a
message byte by byte from the documented layout, sb_msg_unpack must read it back field by field, sb_msg_pack must
write the same bytes, damaged buffers must be refused.  The gpu-marked variant also runs sb_compute from the unpacked
message.  jni/shifu_b200_jni.c itself is not compiled: no JDK on this box."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def build(tmp_path):
    import __graft_entry__ as g
    g.build()
    exe = str(tmp_path / "wire_check")
    lib_dir = os.path.join(ROOT, "shifu_b200")
    subprocess.check_call(["gcc", "-std=c11", "-O1", "-Wall", "-Wextra", "-o", exe, os.path.join(HERE, "wire_check.c"),
                           "-L" + lib_dir, "-lshifu_b200", "-Wl,-rpath," + lib_dir])
    return exe


def test_wire_format_unpacks_field_by_field(tmp_path):
    out = subprocess.run([build(tmp_path)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "wire_check ok" in out.stdout, out.stdout + out.stderr


@pytest.mark.gpu
def test_sb_compute_from_an_unpacked_message(tmp_path):
    out = subprocess.run([build(tmp_path), "gpu"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "sb_compute from the unpacked message" in out.stdout, out.stdout + out.stderr


def test_python_message_packs_to_the_same_layout():
    """An IterIn built by the ctypes message builder, packed by the library and unpacked again, is the same message."""
    import ctypes as C
    import numpy as np
    from shifu_b200 import capi
    from shifu_b200.capi import ConfigArrays
    from shifu_b200.messages import IterIn, Node
    lib = capi.load_library()
    cfg = ConfigArrays([8, 8, 5], is_categorical=[0, 0, 1])
    root = Node(1, predict=0.5)
    root.feature, root.threshold = 0, 3.0
    root.left, root.right = Node(2, predict=-1.0), Node(3, predict=1.0)
    root.left.feature, root.left.split_type, root.left.is_left, root.left.cat_bits = 2, capi.SB_SPLIT_CATEGORICAL, True, bytes([0b101])
    root.left.left, root.left.right = Node(4, predict=0.0), Node(5, predict=2.0)
    n = 9
    it = IterIn(cfg, [(3, root)], [(0, 3, 4, [2, 0]), (1, 3, 5, []), (2, 3, 3, [1])], switch_to_next_tree=True,
                is_first_tree=False, new_weights=np.arange(n, dtype=np.float32), dropout=np.arange(n) % 2)
    need = lib.sb_msg_pack(C.byref(it.struct), n, None, 0)
    buf = np.zeros(need // 8 + 1, dtype=np.uint64)  # 8-byte aligned storage
    assert lib.sb_msg_pack(C.byref(it.struct), n, buf.ctypes.data, need) == need
    h = C.c_void_p()
    capi.check(lib, lib.sb_msg_unpack(buf.ctypes.data, need, C.byref(h)))
    back = lib.sb_msg_iter_in(h).contents
    a, b = it.struct, back
    assert (a.n_trees, a.n_todo, a.switch_to_next_tree, a.is_first_tree) == (b.n_trees, b.n_todo, b.switch_to_next_tree, b.is_first_tree)
    ta, tb = a.trees[0], b.trees[0]
    assert (ta.tree_id, ta.n_nodes, ta.mask_pool_bytes) == (tb.tree_id, tb.n_nodes, tb.mask_pool_bytes)
    assert C.string_at(ta.nodes, C.sizeof(capi.SbNode) * ta.n_nodes) == C.string_at(tb.nodes, C.sizeof(capi.SbNode) * tb.n_nodes)
    assert C.string_at(ta.mask_pool, ta.mask_pool_bytes) == C.string_at(tb.mask_pool, tb.mask_pool_bytes)
    for i in range(a.n_todo):
        x, y = a.todo[i], b.todo[i]
        assert (x.key, x.tree_id, x.node_id, x.n_features) == (y.key, y.tree_id, y.node_id, y.n_features)
        assert [x.features[k] for k in range(x.n_features)] == [y.features[k] for k in range(y.n_features)]
    assert [b.new_weights[i] for i in range(n)] == list(range(n)) and [b.dropout[i] for i in range(n)] == [i % 2 for i in range(n)]
    lib.sb_msg_free(h)
