"""Multi-GPU consistency check (run under torchrun): the statistics after the integer all-reduce must be identical on
every rank and equal to a single-GPU run over the concatenated rows (rank 0 also holds the full data set)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import DevView  # noqa: E402
from shifu_b200.capi import ConfigArrays, Worker  # noqa: E402
from shifu_b200.capi import NativeMaster as DTMaster  # noqa: E402  (the C++ DTMaster mirror)

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
R, F = 400_000, 70
cfg = ConfigArrays([64] * F, max_depth=5, tree_num=3, device=int(os.environ["LOCAL_RANK"]), min_instances_per_node=5)
w = Worker(cfg)
per = R // world
w.load_synthetic(per, seed=5, row0=rank * per, weighted=True)
full = Worker(cfg) if rank == 0 else None
if full is not None:
    full.load_synthetic(per * world, seed=5, row0=0, weighted=True)
m = DTMaster(cfg, tree_num=3, max_depth=5, worker_number=world)
it = m.build_initial_master_params()
n_iter = 0
while it is not None:
    w.compute_begin(it)
    scales_fresh, acc_fresh = w.pass_info()
    sp, sb, rp, rb, ap, ab = w.device_buffers()
    if scales_fresh:
        dist.all_reduce(torch.as_tensor(DevView(sp, 3, "<i4"), device="cuda"), op=dist.ReduceOp.MAX)
    w.compute_stats()
    sp, sb, rp, rb, ap, ab = w.device_buffers()
    dist.all_reduce(torch.as_tensor(DevView(rp, rb // 8, "<i8"), device="cuda"), op=dist.ReduceOp.SUM)
    if acc_fresh:
        dist.all_reduce(torch.as_tensor(DevView(ap, 4, "<f8"), device="cuda"), op=dist.ReduceOp.SUM)
    n = w.stats_len(it)
    sc = w.compute_finish(n, want_stats=True)
    gains, bits = w.best_splits(it.n_todo)
    if full is not None:
        ref = full.compute(it)
        assert np.array_equal(ref[4], sc[4]), "sharded statistics differ from the single-GPU run"
        assert abs(ref[2] - sc[2]) <= 1e-9 * abs(ref[2]) + 1e-12 and ref[0] == sc[0]
    # identical on all ranks
    t = torch.from_numpy(sc[4].copy()).cuda()
    t0 = t.clone()
    dist.broadcast(t0, 0)
    assert torch.equal(t, t0), "ranks disagree"
    it = m.do_compute(gains, bits)
    n_iter += 1
if rank == 0:
    print("multi-gpu check ok: %d iterations, %d trees, bit-identical to the single-GPU run" % (n_iter, m.n_trees))

# ---- the same through the library's own NCCL communicator (sb_dist_init) and the native tree loop (sbh_grow_tree)
from shifu_b200.capi import NativeMaster, dist_unique_id  # noqa: E402

ids = [dist_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
w2 = Worker(cfg)
w2.load_synthetic(per, seed=5, row0=rank * per, weighted=True)
w2.dist_init(ids[0], rank, world)
full2 = None
if rank == 0:
    full2 = Worker(cfg)
    full2.load_synthetic(per * world, seed=5, row0=0, weighted=True)
m2 = DTMaster(cfg, tree_num=2, max_depth=5, worker_number=world)
it = m2.build_initial_master_params()
n_iter = 0
while it is not None:
    sc = w2.compute(it)  # begin + stats (+ all-reduce inside) + finish
    gains, bits = w2.best_splits(it.n_todo)
    if full2 is not None:
        ref = full2.compute(it)
        assert np.array_equal(ref[4], sc[4]), "in-library merge differs from the single-GPU run"
        assert abs(ref[2] - sc[2]) <= 1e-9 * abs(ref[2]) + 1e-12 and ref[0] == sc[0]
    it = m2.do_compute(gains, bits)
    n_iter += 1
w2.close()
if full2 is not None:
    full2.close()


def grow_all(worker, n_workers):
    nm = NativeMaster(cfg, tree_num=3, max_depth=5, worker_number=n_workers)
    msg, total = nm.build_initial_master_params(), 0
    while msg is not None:
        msg, passes, _ = nm.grow_tree(worker, msg)
        total += passes
    out = []
    for t in range(nm.n_trees):
        nodes, pool, extras = nm.export_tree(t)
        out.append((bytes(b"".join(bytes(x) for x in nodes)), pool.tobytes(), extras.tobytes()))
    return out, total


ids = [dist_unique_id() if rank == 0 else None]  # a communicator per handle
dist.broadcast_object_list(ids, src=0)
w3 = Worker(cfg)
w3.load_synthetic(per, seed=5, row0=rank * per, weighted=True)
w3.dist_init(ids[0], rank, world)
trees_sharded, passes = grow_all(w3, world)
w3.close()
if rank == 0:
    full3 = Worker(cfg)
    full3.load_synthetic(per * world, seed=5, row0=0, weighted=True)
    trees_full, passes_full = grow_all(full3, 1)
    full3.close()
    assert passes == passes_full and trees_sharded == trees_full, "native tree loop: sharded trees differ from the single-GPU trees"
    print("in-library NCCL merge ok: %d iterations bit-identical; native tree loop: %d trees, %d passes, identical to single GPU"
          % (n_iter, len(trees_full), passes))

# ---- a shard without a single active row (all its sample weights are 0): its row lists are empty, its kernels must do
# nothing, and the merged result must still equal the single-GPU run over the concatenated rows
ids = [dist_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
w4 = Worker(cfg)
w4.load_synthetic(per, seed=5, row0=rank * per, weighted=False)
w4.dist_init(ids[0], rank, world)
if rank == world - 1:
    w4.set_weights(0, np.zeros(per, np.float32))
full4 = None
if rank == 0:
    full4 = Worker(cfg)
    full4.load_synthetic(per * world, seed=5, row0=0, weighted=False)
    wz = np.ones(per * world, np.float32)
    wz[(world - 1) * per:] = 0.0
    full4.set_weights(0, wz)
m4 = DTMaster(cfg, tree_num=1, max_depth=5, worker_number=world)
it = m4.build_initial_master_params()
n_iter4 = 0
while it is not None:
    sc = w4.compute(it)
    gains, bits = w4.best_splits(it.n_todo)
    if full4 is not None:
        ref = full4.compute(it)
        assert np.array_equal(ref[4], sc[4]), "empty shard: merged statistics differ from the single-GPU run"
    it = m4.do_compute(gains, bits)
    n_iter4 += 1
w4.close()
if full4 is not None:
    full4.close()
    print("empty-shard check ok: %d iterations bit-identical" % n_iter4)

# ---- a random forest: all trees advance together, every pass is one forest pass (one list for all trees) on every shard
from shifu_b200 import capi  # noqa: E402

T_RF = 5
cfg_rf = ConfigArrays([32] * 40, impurity=capi.SB_IMPURITY_GINI, num_classes=2, loss=capi.SB_LOSS_SQUARED, is_gbdt=False,
                      tree_num=T_RF, weights_per_row=T_RF, max_depth=6, min_instances_per_node=5,
                      device=int(os.environ["LOCAL_RANK"]))


def grow_forest(worker, n_workers):
    nm = NativeMaster(cfg_rf, tree_num=T_RF, max_depth=6, is_gbdt=False, feature_subset_strategy="TWOTHIRDS",
                      feature_sampling_seed=11, max_stats_memory=(1 << 40), worker_number=n_workers)
    msg, total = nm.build_initial_master_params(), 0
    while msg is not None:
        msg, passes, _ = nm.grow_tree(worker, msg)
        total += passes
    out = []
    for t in range(nm.n_trees):
        nodes, pool, extras = nm.export_tree(t)
        out.append((bytes(b"".join(bytes(x) for x in nodes)), pool.tobytes(), extras.tobytes()))
    nm.close()
    return out, total


ids = [dist_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
w5 = Worker(cfg_rf)
w5.load_synthetic(per, seed=9, row0=rank * per, label_kind=1)
for t in range(T_RF):
    w5.sample_poisson(t, 9, 1.0, row0=rank * per)
w5.dist_init(ids[0], rank, world)
forest_sharded, passes_rf = grow_forest(w5, world)
w5.close()
if rank == 0:
    full5 = Worker(cfg_rf)
    full5.load_synthetic(per * world, seed=9, row0=0, label_kind=1)
    for t in range(T_RF):
        full5.sample_poisson(t, 9, 1.0, row0=0)
    forest_full, passes_rf_full = grow_forest(full5, 1)
    full5.close()
    assert passes_rf == passes_rf_full and forest_sharded == forest_full, "random forest: sharded trees differ from the single-GPU trees"
    print("forest check ok: %d trees, %d passes, identical to single GPU" % (len(forest_full), passes_rf))
dist.barrier()
dist.destroy_process_group()
