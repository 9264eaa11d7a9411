"""world_size-2 gloo test of the N>1 path's host logic on CPU: contiguous row shards + SUM/MAX collectives give every
rank the statistics (and hence the split decisions) of a single worker over all rows.  The per-rank worker here is
the CPU oracle (no GPU in this container); on the GPU box tools/multi_gpu_check.py runs the same protocol over NCCL
with the CUDA path and checks bit-identity."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rank_main(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from helpers import make_data
    from oracle.pyoracle import OracleWorker
    from shifu_b200.capi import ConfigArrays
    from shifu_b200.messages import IterIn, Node
    from shifu_b200.sharding import reduce_pass, shard_rows

    n, n_bins = 9001, [16] * 7 + [5]
    cols, y, sig, _ = make_data(n, n_bins, seed=3)  # unit weights: every per-bin sum is an exact integer or a short sum
    cfg = ConfigArrays(n_bins, max_depth=4)
    row0, cnt = shard_rows(n, world, rank)
    o = OracleWorker(cfg)
    o.load_rows([c[row0:row0 + cnt] for c in cols], y[row0:row0 + cnt], None, None)
    root = Node(1)
    it = IterIn(cfg, [(0, root)], [(0, 0, 1, [])], switch_to_next_tree=True, is_first_tree=True)
    tc, vc, te, ve, stats = o.compute(it, n_threads=1)
    # the protocol of one pass (scales are a MAX of float bit patterns)
    m = np.array([1.0, float(np.abs(y[row0:row0 + cnt]).max()), float((y[row0:row0 + cnt] ** 2).max())], dtype=np.float32)
    scales = torch.from_numpy(m.view(np.int32).copy())
    raw = torch.from_numpy(np.round(stats * 1024).astype(np.int64))  # exact: labels are 0/1
    acc = torch.tensor([tc, vc, te, ve], dtype=torch.float64)
    reduce_pass(dist, scales, raw, acc, True, True)
    merged = raw.numpy().astype(np.float64) / 1024
    gains, _ = o.best_splits(it, merged)
    np.save(os.path.join(out_dir, "stats_%d.npy" % rank), merged)
    np.save(os.path.join(out_dir, "meta_%d.npy" % rank),
            np.array([acc[0].item(), gains[0].feature, gains[0].split_bin, gains[0].gain, row0, cnt,
                      float(scales.numpy().view(np.float32)[1])]))
    dist.destroy_process_group()


def test_two_ranks_equal_single_worker(tmp_path):
    world, port = 2, 29731
    mp.spawn(_rank_main, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import make_data
    from oracle.pyoracle import OracleWorker
    from shifu_b200.capi import ConfigArrays
    from shifu_b200.messages import IterIn, Node
    n, n_bins = 9001, [16] * 7 + [5]
    cols, y, sig, _ = make_data(n, n_bins, seed=3)
    cfg = ConfigArrays(n_bins, max_depth=4)
    o = OracleWorker(cfg)
    o.load_rows(cols, y, None, None)
    it = IterIn(cfg, [(0, Node(1))], [(0, 0, 1, [])], switch_to_next_tree=True, is_first_tree=True)
    tc, vc, te, ve, stats = o.compute(it, n_threads=1)
    g, _ = o.best_splits(it, stats)
    s0, s1 = np.load(tmp_path / "stats_0.npy"), np.load(tmp_path / "stats_1.npy")
    m0, m1 = np.load(tmp_path / "meta_0.npy"), np.load(tmp_path / "meta_1.npy")
    assert np.array_equal(s0, s1) and np.array_equal(s0, stats)          # merged == single worker, on every rank
    assert m0[0] == m1[0] == tc == 9001.0                                 # counts add up
    assert (m0[1], m0[2]) == (m1[1], m1[2]) == (g[0].feature, g[0].split_bin) and m0[3] == m1[3] == g[0].gain
    assert (m0[4], m0[5], m1[4], m1[5]) == (0, 4501, 4501, 4500)          # contiguous shards, remainder to rank 0
    assert m0[6] == m1[6] == 1.0                                          # MAX collective


def test_shard_rows_partition():
    from shifu_b200.sharding import shard_rows
    for n in (0, 1, 7, 100, 100_000_003):
        for world in (1, 2, 3, 8):
            parts = [shard_rows(n, world, r) for r in range(world)]
            assert parts[0][0] == 0 and sum(p[1] for p in parts) == n
            for a, b in zip(parts, parts[1:]):
                assert a[0] + a[1] == b[0] and a[1] - b[1] in (0, 1)
