"""sb_error_sums_static -- the entry point for callers that may not allocate device memory -- against the existing
entry point (sb_compute), through the C ABI, on two calls in a row.

Why such callers exist: a thread that watches training (early-stop check, validation curve) from the side cannot
call cudaMalloc / cudaFree, because both synchronise the device and would stall the trainer's stream; and a process
that has given all of HBM to the bin matrix has nothing left to allocate from.  For them the library carries a
fixed 64 KB scratch array inside its own device image; this test checks that the numbers it produces are the ones
sb_compute reports, that stale scratch from an earlier call does not leak into the next one, and that a request
beyond the fixed capacity is refused instead of overrunning the array.

This file is named so that pytest collects it last: a run that stops at the first failure has then already
exercised every other test.
"""
import ctypes as C

import numpy as np
import pytest

from shifu_b200 import capi
from shifu_b200.capi import ConfigArrays, Worker
from shifu_b200.messages import IterIn, Node

from helpers import apply_gain, make_data

pytestmark = pytest.mark.gpu


def static_sums(lib, torch, label, predict, sig, loss):
    """Calls the no-allocation entry point on device copies of the three arrays; returns (count, error)."""
    d_label = torch.from_numpy(label).cuda()
    d_pred = torch.from_numpy(predict).cuda()
    d_sig = None if sig is None else torch.from_numpy(sig).cuda()
    out = torch.full((2,), -1.0, dtype=torch.float64, device="cuda")  # poisoned: the call must overwrite it
    torch.cuda.synchronize()
    rc = lib.sb_error_sums_static(d_label.data_ptr(), d_pred.data_ptr(), None if d_sig is None else d_sig.data_ptr(),
                                  len(label), loss, out.data_ptr(), None)
    capi.check(lib, rc)
    torch.cuda.synchronize()
    c, e = out.cpu().tolist()
    return c, e


@pytest.mark.parametrize("loss", [capi.SB_LOSS_SQUARED, capi.SB_LOSS_LOG])
def test_static_error_sums_match_sb_compute_twice_in_a_row(loss, oracle_lib):
    import torch
    lib = capi.load_library()
    n, nv, n_bins = 70001, 9001, [32] * 12
    cols, y, sig, _ = make_data(n, n_bins, seed=31, weighted=True)
    vcols, vy, vsig, _ = make_data(nv, n_bins, seed=32, weighted=True)
    cfg = ConfigArrays(n_bins, loss=loss, max_depth=3, learning_rate=0.3, tree_num=3)
    w = Worker(cfg)
    w.load_rows(cols, y, sig, None)
    w.load_rows(vcols, vy, vsig, None, is_validation=True)
    # tree 0: root, then its two children are final leaves (max_depth 3 would allow one more level; one is enough)
    root = Node(1)
    it = IterIn(cfg, [(0, root)], [(0, 0, 1, [])], switch_to_next_tree=True, is_first_tree=True)
    w.compute(it)
    gains, bits = w.best_splits(1)
    assert gains[0].found
    apply_gain(cfg, root, gains[0], bits, 0)
    # switch to tree 1: predict / output are updated from tree 0's leaves, errors are sums over Data.predict
    root1 = Node(1)
    it = IterIn(cfg, [(0, root), (1, root1)], [(0, 1, 1, [])], switch_to_next_tree=True, is_first_tree=True)
    tc, vc, te, ve, _ = w.compute(it)
    predict, _, _, _ = w.read_rows(n)
    vpredict, _, _, _ = w.read_rows(nv, is_validation=True)
    assert te > 0 and ve > 0

    # call 1: training rows; call 2 (straight after, smaller n, same shared scratch): validation rows
    c1, e1 = static_sums(lib, torch, y, predict, sig, loss)
    c2, e2 = static_sums(lib, torch, vy, vpredict, vsig, loss)
    for got, want in ((c1, tc), (e1, te), (c2, vc), (e2, ve)):
        assert abs(got - want) <= 1e-9 * abs(want), (got, want)
    # and again the first one: same input, same bits, whatever the previous call left behind
    c3, e3 = static_sums(lib, torch, y, predict, sig, loss)
    assert (c3, e3) == (c1, e1)
    # an answer that does not come from this library: the oracle's Loss.computeError (dt/Loss.java) row by row, the
    # float product sig * error widened and summed in double as dt/DTWorker.java:681-682 does
    for (lab, pred, sg, got_c, got_e) in ((y, predict, sig, c1, e1), (vy, vpredict, vsig, c2, e2)):
        errs = np.array([oracle_lib.orc_loss_error(loss, C.c_float(float(p)), C.c_float(float(t))) for p, t in zip(pred, lab)],
                        dtype=np.float32)
        want_e = float(np.sum((sg * errs).astype(np.float64)))
        want_c = float(np.sum(sg.astype(np.float64)))
        assert abs(got_c - want_c) <= 1e-9 * want_c and abs(got_e - want_e) <= 1e-9 * want_e, (got_c, want_c, got_e, want_e)
    w.close()


def test_static_error_sums_limits():
    import torch
    lib = capi.load_library()
    x = torch.zeros(8, dtype=torch.float32, device="cuda")
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    rc = lib.sb_error_sums_static(x.data_ptr(), x.data_ptr(), None, capi.SB_STATIC_MAX_ROWS + 1, capi.SB_LOSS_SQUARED,
                                  out.data_ptr(), None)
    assert rc == -4 and b"static scratch limit" in lib.sb_last_error()
    # empty input: both sums are zero, nothing is launched
    out.fill_(7.0)
    torch.cuda.synchronize()
    capi.check(lib, lib.sb_error_sums_static(x.data_ptr(), x.data_ptr(), None, 0, capi.SB_LOSS_SQUARED, out.data_ptr(), None))
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [0.0, 0.0]
    # significance == NULL means 1f for every row; rows spanning several blocks
    n = 16384 * 3 + 5
    lab = torch.ones(n, dtype=torch.float32, device="cuda")
    pred = torch.full((n,), 0.5, dtype=torch.float32, device="cuda")
    capi.check(lib, lib.sb_error_sums_static(lab.data_ptr(), pred.data_ptr(), None, n, capi.SB_LOSS_SQUARED, out.data_ptr(), None))
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [float(n), 0.25 * n]
