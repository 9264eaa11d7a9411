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
from shifu_b200.dt_master import DTMaster  # noqa: E402

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
    print("multi-gpu check ok: %d iterations, %d trees, bit-identical to the single-GPU run" % (n_iter, len(m.trees)))
dist.destroy_process_group()
