"""Per-level timing of the histogram kernel on a synthetic GBT workload (developer tool, not the bench)."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shifu_b200.capi import ConfigArrays, Worker  # noqa: E402
from shifu_b200.capi import NativeMaster  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=10_000_000)
ap.add_argument("--features", type=int, default=200)
ap.add_argument("--bins", type=int, default=64)
ap.add_argument("--depth", type=int, default=8)
ap.add_argument("--trees", type=int, default=2)
ap.add_argument("--weighted", type=int, default=0)
ap.add_argument("--flags", type=int, default=0, help="sb_config.flags (SB_FLAG_*)")
ap.add_argument("--feature-rate", type=float, default=1.0, help="featureSubsetRate per node (C4: 0.5)")
ap.add_argument("--lib", default=None, help="alternative libshifu_b200.so (kernel experiments)")
a = ap.parse_args()
if a.lib:
    from shifu_b200 import capi
    capi.LIB_PATH = os.path.abspath(a.lib)

cfg = ConfigArrays([a.bins] * a.features, max_depth=a.depth, tree_num=a.trees, learning_rate=0.05, min_instances_per_node=5, flags=a.flags)
w = Worker(cfg)
t0 = time.time()
w.load_synthetic(a.rows, seed=1002, weighted=bool(a.weighted))
print("load_synthetic %.2fs" % (time.time() - t0))
w.set_profiling(True)
m = NativeMaster(cfg, tree_num=a.trees, max_depth=a.depth, learning_rate=0.05,
                 feature_subset_strategy=(None if a.feature_rate < 1.0 else "ALL"), feature_subset_rate=a.feature_rate, feature_sampling_seed=7)
it = m.build_initial_master_params()
algo_bytes = a.rows * (a.features + 16)
tm = time.time()
while it is not None:
    t0 = time.time()
    w.compute_begin(it)
    t1 = time.time()
    w.compute_stats()
    t2 = time.time()
    sc = w.compute_finish()
    t3 = time.time()
    gains, bits = w.best_splits(it.n_todo)
    t4 = time.time()
    wall = (t4 - t0) * 1e3
    host_master = (t0 - tm) * 1e3
    print("   host: master+message %.2f  begin %.2f  stats %.2f  finish(sync) %.2f  best_splits %.2f ms" % (
        host_master, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3))
    ms, npass, nl = w.hist_time()
    lvl = int(max(it.struct.todo[i].node_id for i in range(it.n_todo))).bit_length()
    print("tree %d level %d todo %3d  hist %.3f ms  wall %.3f ms  %.1f GB/s (algorithmic)  %.3e cells/s  err=%.6g" % (
        m.n_trees - 1, lvl, it.n_todo, ms, wall, algo_bytes / ms / 1e6, a.rows * a.features / ms * 1e3,
        sc[2] / max(sc[0], 1)))
    tm = time.time()
    it = m.do_compute(gains, bits)
w.close()
