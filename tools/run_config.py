"""Runs one of BASELINE.json's configs through the device path and prints one JSON line (developer tool; bench.py is
the contract benchmark for C2).  Under torchrun the rows are sharded as in bench.py.

  c1   10M x 100 numeric x 32 bins, RF 50 trees depth 10, Gini, Poisson(1) bag counts, TWOTHIRDS features per node
  c3   50M x (120 numeric x 64 bins + 30 categorical with 10/50/200/1000 categories), GBT log loss, variance, depth 8
  c3rf the same columns, RF 20 trees depth 8, entropy (the reference rejects GBT + entropy: ModelInspector.java:898-904)
  c4   1B x 100 x 64 bins, GBT depth 8, 0.5 row bagging per tree (injected mask), 0.5 features per node
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from shifu_b200 import capi  # noqa: E402
from shifu_b200.capi import ConfigArrays, NativeMaster, Worker  # noqa: E402
from shifu_b200.sharding import shard_rows  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", required=True, choices=["c1", "c3", "c3rf", "c4"])
ap.add_argument("--rows", type=int, default=0)
ap.add_argument("--trees", type=int, default=0)
a = ap.parse_args()

import torch  # noqa: E402

world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

cat_bins = [11, 51, 201, 1001]
C = {
    "c1": dict(rows=10_000_000, n_bins=[32] * 100, is_cat=None, gbdt=False, trees=50, depth=10, imp=capi.SB_IMPURITY_GINI,
               loss=capi.SB_LOSS_SQUARED, fss="TWOTHIRDS", rate=1.0, label_kind=1, seed=1001),
    "c3": dict(rows=50_000_000, n_bins=[64] * 120 + [cat_bins[i % 4] for i in range(30)], is_cat=[0] * 120 + [1] * 30, gbdt=True,
               trees=3, depth=8, imp=capi.SB_IMPURITY_VARIANCE, loss=capi.SB_LOSS_LOG, fss="ALL", rate=1.0, label_kind=0, seed=1003),
    "c3rf": dict(rows=50_000_000, n_bins=[64] * 120 + [cat_bins[i % 4] for i in range(30)], is_cat=[0] * 120 + [1] * 30, gbdt=False,
                 trees=20, depth=8, imp=capi.SB_IMPURITY_ENTROPY, loss=capi.SB_LOSS_SQUARED, fss="TWOTHIRDS", rate=1.0, label_kind=0, seed=1003),
    "c4": dict(rows=1_000_000_000, n_bins=[64] * 100, is_cat=None, gbdt=True, trees=3, depth=8, imp=capi.SB_IMPURITY_VARIANCE,
               loss=capi.SB_LOSS_SQUARED, fss=None, rate=0.5, label_kind=0, seed=1004),
}[a.config]
if a.rows:
    C["rows"] = a.rows
if a.trees:
    C["trees"] = a.trees
F = len(C["n_bins"])
row0, n_local = shard_rows(C["rows"], world, rank)
T = C["trees"]
cfg = ConfigArrays(C["n_bins"], is_categorical=C["is_cat"], impurity=C["imp"], num_classes=2, loss=C["loss"], is_gbdt=C["gbdt"],
                   tree_num=T + (1 if C["gbdt"] else 0), weights_per_row=1 if C["gbdt"] else T, max_depth=C["depth"],
                   min_instances_per_node=5, learning_rate=0.05, device=local)
w = Worker(cfg)
t0 = time.time()
w.load_synthetic(n_local, seed=C["seed"], row0=row0, label_kind=C["label_kind"])
rng = np.random.default_rng(C["seed"] + rank)
if not C["gbdt"]:
    for t in range(T):  # Poisson(1) bag counts per (row, tree) (dt/DTWorker.java:1547-1571), drawn on the device
        w.sample_poisson(t, C["seed"], 1.0, row0=row0)
load_s = time.time() - t0
m = NativeMaster(cfg, tree_num=T + (1 if C["gbdt"] else 0), max_depth=C["depth"], is_gbdt=C["gbdt"], learning_rate=0.05,
                 feature_subset_strategy=C["fss"], feature_subset_rate=C["rate"], feature_sampling_seed=7,
                 max_stats_memory=(1 << 40), worker_number=world)
it = m.build_initial_master_params()
w.set_profiling(True)
torch.cuda.synchronize()
if dist is not None:
    dist.barrier()
t0 = time.time()
passes, cells, per_pass = 0, 0.0, []
trees_done = 0
while it is not None:
    tp = time.time()
    n_todo = it.n_todo
    todo_trees = len({it.struct.todo[i].tree_id for i in range(n_todo)})
    feats = sum((it.struct.todo[i].n_features or F) for i in range(n_todo)) / max(n_todo, 1)
    if C["gbdt"] and C["rate"] < 1.0 and it.struct.switch_to_next_tree and m.n_trees >= 2:
        # bagging mask redrawn per tree (dt/DTWorker.java:643-668) on the device, java.util.Random semantics
        w.sample_weights(0, 1000 + m.n_trees, 0.5, 0)
    nxt, p, sc = None, 0, None
    # one pass (bench.grow_tree grows a whole GBT tree; here passes are timed one by one)
    w.compute_begin(it)
    if dist is not None:
        from shifu_b200.sharding import reduce_pass
        sf, af = w.pass_info()
        sp, sb, rp, rb, ap_, ab = w.device_buffers()
        reduce_pass(dist, torch.as_tensor(bench.DevView(sp, 3, "<i4"), device="cuda") if sf else None, None, None, sf, False)
    w.compute_stats()
    if dist is not None:
        sp, sb, rp, rb, ap_, ab = w.device_buffers()
        reduce_pass(dist, None, torch.as_tensor(bench.DevView(rp, rb // 8, "<i8"), device="cuda") if rb > 0 else None,
                    torch.as_tensor(bench.DevView(ap_, 4, "<f8"), device="cuda") if af else None, False, af)
    sc = w.compute_finish()
    has_cat = C["is_cat"] is not None
    gains, bits = w.best_splits(n_todo, with_cat=has_cat)
    it = m.do_compute(gains, bits)
    dt = time.time() - tp
    passes += 1
    cells += float(C["rows"]) * feats * todo_trees
    per_pass.append((n_todo, todo_trees, round(dt * 1e3, 2)))
    if C["gbdt"] and m.n_trees > T:
        break
torch.cuda.synchronize()
wall = time.time() - t0
hist_ms, hp, hl = w.hist_time()
if rank == 0:
    print(json.dumps({"config": a.config, "rows": C["rows"], "features": F, "n_gpus": world, "trees": T, "depth": C["depth"],
                      "passes": passes, "wall_s": wall, "sec_per_tree": wall / T, "hist_ms_total": hist_ms,
                      "rows_features_trees_per_s": cells / wall, "load_s": load_s,
                      "train_error": (sc[2] / sc[0]) if sc and sc[0] else None,
                      "per_pass_(todo,trees,ms)": per_pass[:12] + (["..."] if len(per_pass) > 12 else [])}))
w.close()
if dist is not None:
    dist.destroy_process_group()
