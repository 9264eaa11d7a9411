#!/usr/bin/env python
"""bench.py -- throughput of the GBT statistics-build hot path on B200 (BASELINE.json metric).

Workload (BASELINE.json configs[2], the one the north-star target is quoted on): synthetic 100M rows x 200
numeric features x 64 bins, GBT, squared loss, depth-8 level-wise trees.  One "step" = one GBT tree =
the D-1 statistics passes of its levels + routing + residual update + error sums + all-reduce + on-device best
split, driven by the host DTMaster mirror.  `value` = rows x features x passes / time over all GPUs, with the data
resident in HBM; `e2e` = the same through sb_compute with HOST in/out buffers (message in, statistics out).
Rows are sharded contiguously over the GPUs (strong scaling: the 100M rows are fixed); the DTMaster merge is an
NCCL all-reduce of the integer accumulators.

`--impl reference` times the CPU restatement of the reference path (oracle/, all host threads) on a bounded
sample of the same workload; the Java original cannot run here (no JVM).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 1002


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=100_000_000)
    ap.add_argument("--features", type=int, default=200)
    ap.add_argument("--bins", type=int, default=64)
    ap.add_argument("--depth", type=int, default=8)
    ap.add_argument("--cpu-sample-rows", type=int, default=2_000_000)
    ap.add_argument("--no-other-configs", action="store_true", help="skip the short runs of the other BASELINE.json configs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--flags", type=int, default=0, help="sb_config.flags (SB_FLAG_*), for A/B runs")
    return ap.parse_args()


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"  # /opt/skills/guides/B200_PROFILING.md


class ClockSampler:
    """nvidia-smi samples of one GPU during the timed region (clocks line of the profiling recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.samples = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            parts = [p.strip() for p in s.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class DevView:
    """A raw device allocation exposed through __cuda_array_interface__ so torch can wrap it without a copy."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def grow_tree(worker, master, it, e2e_buffers=None):
    """One GBT tree through the device path; returns (next message, passes, last scalars).

    Device-resident measurement: the native iteration loop (sbh_grow_tree).  End-to-end measurement: the call a
    user of the reference makes, DTWorker.doCompute with HOST buffers (message in, statistics out: sb_compute), then
    sb_best_splits and the master step.  With several GPUs the merge is the library's own NCCL all-reduce in both, and
    only rank 0 asks for the merged statistics on its host."""
    from shifu_b200 import capi
    if e2e_buffers is None:
        return master.grow_tree(worker, it)
    passes = 0
    tree_index = master.n_trees
    scalars = None
    while it is not None:
        n = worker.stats_len(it)
        out = capi.SbIterOut()
        if e2e_buffers["want_stats"]:
            out.stats = C.cast(e2e_buffers["stats_ptr"], C.POINTER(C.c_double))
            out.stats_len = e2e_buffers["stats_cap"]
            assert n <= e2e_buffers["stats_cap"]
        capi.check(worker.lib, worker.lib.sb_compute(worker.h, C.byref(it.struct), C.byref(out)))
        e2e_buffers["d2h"] += n * 8 + 32
        scalars = (out.train_count, out.validation_count, out.train_error, out.validation_error)
        gains, bits = worker.best_splits(it.n_todo)
        passes += 1
        it = master.do_compute(gains, bits)
        if it is None or master.n_trees != tree_index:
            break
    return it, passes, scalars


def measured_traffic(a, rows_local):
    """dram__bytes_read + dram__bytes_write per launch of the histogram kernel.  NOT measured in this run: it is the
    mean over the 7 passes of a tree in the committed ncu --set full capture of this very workload at 100M rows on one
    GPU (profiles/r02_hist_rows_traffic.json), scaled to this rank's share of the rows; the bench line says so in
    roofline.traffic_source.  None for any other shape."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r02_hist_rows_traffic.json")
    try:
        t = json.load(open(path))
    except Exception:
        return None
    if (a.rows, a.features, a.bins, a.depth) != (t.get("rows"), 200, 64, 8):
        return None
    return t["mean_dram_bytes_per_launch"] * rows_local / t["rows"]


def trees_digest(master, n_trees):
    """sha256 over the exported trees (node records, masks, gains): the same at every GPU count, because the merged
    integer statistics are."""
    import hashlib
    h = hashlib.sha256()
    for t in range(n_trees):
        nodes, pool, extras = master.export_tree(t)
        for nd in nodes:
            h.update(bytes(nd))
        h.update(pool.tobytes())
        h.update(extras.tobytes())
    return h.hexdigest()[:16]


CAT_BINS = [11, 51, 201, 1001]
OTHER_CONFIGS = {
    # BASELINE.json configs[1], [3], [4]; seeds as in SURVEY.md 8(d).  `gpus`: the GPU counts at which the config is run.
    "C1": dict(rows=10_000_000, n_bins=[32] * 100, is_cat=None, gbdt=False, trees=50, depth=10, impurity="gini", loss="squared",
               fss="TWOTHIRDS", rate=1.0, bag=None, label_kind=1, seed=1001, gpus=(1,),
               what="synthetic 10M rows x 100 numeric features x 32 bins, RF 50 trees depth 10, Gini, Poisson(1) bags, TWOTHIRDS features per node"),
    "C3": dict(rows=50_000_000, n_bins=[64] * 120 + [CAT_BINS[i % 4] for i in range(30)], is_cat=[0] * 120 + [1] * 30, gbdt=True,
               trees=3, depth=8, impurity="variance", loss="log", fss="ALL", rate=1.0, bag=None, label_kind=0, seed=1003, gpus=(1, 8),
               what="synthetic 50M rows x 150 features (120 numeric x 64 bins + 30 categorical with 10/50/200/1000 categories), GBT log loss depth 8 "
                    "(variance impurity: the reference rejects GBT + entropy, ModelInspector.java:898-904)"),
    # not a BASELINE.json config: C3's columns as a random forest (the probe VERDICT.md round 1 used); run by tools/run_other.py only
    "C3RF": dict(rows=50_000_000, n_bins=[64] * 120 + [CAT_BINS[i % 4] for i in range(30)], is_cat=[0] * 120 + [1] * 30, gbdt=False,
                 trees=20, depth=8, impurity="entropy", loss="squared", fss="TWOTHIRDS", rate=1.0, bag=None, label_kind=0, seed=1003, gpus=(),
                 what="C3's 50M rows x 150 features as RF 20 trees depth 8, Entropy, Poisson(1) bags, TWOTHIRDS features per node"),
    "C4": dict(rows=1_000_000_000, n_bins=[64] * 100, is_cat=None, gbdt=True, trees=3, depth=8, impurity="variance", loss="squared",
               fss=None, rate=0.5, bag=0.5, label_kind=0, seed=1004, gpus=(8,),
               what="synthetic 1B rows x 100 features x 64 bins, GBT depth 8, 0.5 row bagging per tree, 0.5 features per node"),
}


def measure_config(name, world, rank, local_rank, dist_ctx, barrier):
    """A short run of one of the other BASELINE.json configs through the same device path: GBT = `trees` trees after one
    warm-up tree; RF = the whole forest (all trees advance together) after a small warm-up forest on other rows."""
    import torch
    from shifu_b200 import capi
    from shifu_b200.capi import ConfigArrays, NativeMaster, Worker, dist_unique_id
    from shifu_b200.sharding import shard_rows
    c = OTHER_CONFIGS[name]
    imp = {"gini": capi.SB_IMPURITY_GINI, "entropy": capi.SB_IMPURITY_ENTROPY, "variance": capi.SB_IMPURITY_VARIANCE}[c["impurity"]]
    loss = {"squared": capi.SB_LOSS_SQUARED, "log": capi.SB_LOSS_LOG}[c["loss"]]
    F, T, gbdt = len(c["n_bins"]), c["trees"], c["gbdt"]
    row0, n_local = shard_rows(c["rows"], world, rank)
    has_cat = c["is_cat"] is not None

    def make(rows_local, first_row, n_trees):
        cfg = ConfigArrays(c["n_bins"], is_categorical=c["is_cat"], impurity=imp, num_classes=2, loss=loss, is_gbdt=gbdt,
                           tree_num=n_trees + (2 if gbdt else 0), weights_per_row=1 if gbdt else n_trees, max_depth=c["depth"],
                           min_instances_per_node=5, learning_rate=0.05, device=local_rank)
        w = Worker(cfg)
        w.load_synthetic(rows_local, seed=c["seed"], row0=first_row, label_kind=c["label_kind"])
        if not gbdt:
            for t in range(n_trees):  # Poisson(1) bag counts per (row, tree), dt/DTWorker.java:1547-1571
                w.sample_poisson(t, c["seed"], 1.0, row0=first_row)
        if dist_ctx is not None:
            ids = [dist_unique_id() if rank == 0 else None]
            dist_ctx.broadcast_object_list(ids, src=0)
            w.dist_init(ids[0], rank, world)
        m = NativeMaster(cfg, tree_num=n_trees + (2 if gbdt else 0), max_depth=c["depth"], is_gbdt=gbdt, learning_rate=0.05,
                         feature_subset_strategy=c["fss"], feature_subset_rate=c["rate"], feature_sampling_seed=7,
                         max_stats_memory=(1 << 40), worker_number=world)
        return cfg, w, m

    def grow(w, m, it):
        if c["bag"] is not None and m.n_trees >= 2:  # bagging mask redrawn per tree (dt/DTWorker.java:643-668), on the device
            w.sample_weights(0, 1000 + m.n_trees, c["bag"], 0)
        return m.grow_tree(w, it)

    if not gbdt:  # warm the kernels of this configuration on a small forest
        _, ww, wm = make(min(n_local, 500_000), row0, 4)
        wit = wm.build_initial_master_params()
        while wit is not None:
            wit, _, _ = wm.grow_tree(ww, wit)
        ww.close()
        wm.close()
    cfg, w, m = make(n_local, row0, T)
    it = m.build_initial_master_params()
    if gbdt:
        it, _, _ = grow(w, m, it)
    w.hist_time()
    w.set_profiling(True)
    w.rows_streamed(reset=True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    passes, scal, done = 0, None, 0
    while it is not None and done < T:
        it, p, scal = grow(w, m, it)
        passes += p
        done += 1 if gbdt else T
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    hist_ms, hist_passes, _ = w.hist_time()
    streamed = w.rows_streamed()
    if dist_ctx is not None:
        t = torch.tensor([ms, hist_ms], dtype=torch.float64, device="cuda")
        dist_ctx.all_reduce(t, op=dist_ctx.ReduceOp.MAX)
        ms, hist_ms = [float(x) for x in t.tolist()]
    digest = trees_digest(m, T + (1 if gbdt else 0)) if rank == 0 else None
    w.close()
    m.close()
    peak, _ = measured_peak_gbs()
    out = {"workload": c["what"], "n_gpus": world, "trees": T, "passes": passes, "seed": c["seed"],
           "ms_total": ms, "hist_ms_per_pass": hist_ms / max(hist_passes, 1),
           "frac_streamed": float(streamed) * (F + 14) / max(hist_ms * 1e-3, 1e-12) / 1e9 / peak,
           "train_error_last": (scal[2] / scal[0]) if scal and scal[0] else None, "trees_sha256_16": digest}
    if gbdt:
        out["sec_per_tree"] = ms / 1e3 / T
        out["rows_features_per_s"] = float(c["rows"]) * F * passes / (ms * 1e-3)
    else:
        out["sec_per_forest"] = ms / 1e3
        out["rows_features_trees_per_s"] = float(c["rows"]) * F * T * passes / (ms * 1e-3)
        out["note"] = "all %d trees advance together: a pass builds one level of every tree" % T
    return out


def run_b200(a):
    import torch
    from shifu_b200.capi import ConfigArrays, NativeMaster, Worker

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dist_ctx = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist_ctx = dist
    n_gpus = world
    from shifu_b200.sharding import shard_rows
    row0, rows_local = shard_rows(a.rows, world, rank)

    n_trees_total = a.warmup + a.steps + a.steps + 2
    cfg = ConfigArrays([a.bins] * a.features, max_depth=a.depth, tree_num=n_trees_total, learning_rate=0.05,
                       min_instances_per_node=5, device=local_rank, flags=a.flags)
    worker = Worker(cfg)
    worker.load_synthetic(rows_local, seed=SEED, row0=row0, label_kind=0)
    if dist_ctx is not None:
        # the library's own communicator (sb_dist_init): torch.distributed only carries the 128-byte id and the barriers
        from shifu_b200.capi import dist_unique_id
        ids = [dist_unique_id() if rank == 0 else None]
        dist_ctx.broadcast_object_list(ids, src=0)
        worker.dist_init(ids[0], rank, world)
    master = NativeMaster(cfg, tree_num=n_trees_total, max_depth=a.depth, learning_rate=0.05, worker_number=world)
    it = master.build_initial_master_params()

    def barrier():
        torch.cuda.synchronize()
        if dist_ctx is not None:
            dist_ctx.barrier()
            torch.cuda.synchronize()

    # ---- warm-up trees
    for _ in range(a.warmup):
        it, _, _ = grow_tree(worker, master, it)
    worker.hist_time()
    worker.set_profiling(True)
    worker.rows_streamed(reset=True)
    l0, h2d0, d2h0 = worker.counters()

    # ---- timed region 1: device-resident
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    passes = 0
    for _ in range(a.steps):
        it, p, scalars = grow_tree(worker, master, it)
        passes += p
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    l1, h2d1, d2h1 = worker.counters()
    hist_ms, hist_passes, hist_launches = worker.hist_time()
    rows_streamed = worker.rows_streamed(reset=True)
    worker.set_profiling(False)

    # ---- timed region 2: end to end through sb_compute with pinned host buffers
    stats_cap = (1 << (a.depth - 2)) * a.features * a.bins * 3 + 16
    pinned = torch.empty(stats_cap, dtype=torch.float64, pin_memory=True)
    # After the in-library merge every rank holds the same statistics: ONE copy reaches a host (rank 0, where a Guagua
    # master would consume it); the other ranks leave theirs on the device (sb_iter_out.stats = NULL) and only read the
    # four scalars and the gain records.
    e2e_buf = {"stats_ptr": pinned.data_ptr(), "stats_cap": stats_cap, "d2h": 0, "want_stats": rank == 0}
    _, h2d_a, d2h_a = worker.counters()
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    e2e_passes = 0
    for _ in range(a.steps):
        it, p, _ = grow_tree(worker, master, it, e2e_buffers=e2e_buf)
        e2e_passes += p
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1)
    _, h2d_b, d2h_b = worker.counters()

    if dist_ctx is not None:
        t = torch.tensor([ms, ms_e2e, hist_ms], dtype=torch.float64, device="cuda")
        dist_ctx.all_reduce(t, op=dist_ctx.ReduceOp.MAX)
        ms, ms_e2e, hist_ms = [float(x) for x in t.tolist()]

    digest = trees_digest(master, master.n_trees - 1) if rank == 0 else None  # the last tree is the next step's bare root
    n_trees_hashed = master.n_trees - 1
    worker.close()
    master.close()
    # short runs of the other BASELINE.json configs (every rank takes part: the merges are collectives)
    others = {}
    if not a.no_other_configs and (a.rows, a.features, a.bins, a.depth) == (100_000_000, 200, 64, 8):
        for name, c in OTHER_CONFIGS.items():
            if world in c["gpus"]:
                others[name] = measure_config(name, world, rank, local_rank, dist_ctx, barrier)
    if rank != 0:
        if dist_ctx is not None:
            dist_ctx.destroy_process_group()
        return

    cells = float(a.rows) * a.features
    value = cells * passes / (ms * 1e-3)
    e2e_value = cells * e2e_passes / (ms_e2e * 1e-3)
    peak, peak_kind = measured_peak_gbs()
    # Algorithmic bytes of one pass on one GPU: every bin index once + 14 B of per-row
    # scalars (output, significance, sample weight as fp32 and a 2-byte node id; this build keeps a 4-byte id, which
    # is its own cost, not the algorithm's).
    bytes_per_row = a.features * 1 + 14
    algo_bytes = float(rows_local) * bytes_per_row
    avg_hist_ms = hist_ms / max(hist_passes, 1)
    # effective: credits a pass with the bytes of ALL its rows although histogram subtraction lets it read only the
    # rows of the smaller child of every sibling pair (exact, a real saving -- but not bandwidth);
    # streamed: the rows the statistics kernels actually read (sb_rows_streamed) times the same bytes per row.
    achieved_effective = algo_bytes / (avg_hist_ms * 1e-3) / 1e9
    achieved = float(rows_streamed) * bytes_per_row / (hist_ms * 1e-3) / 1e9
    # second ceiling, the one that binds: shared-memory atomic word updates.  A cell costs 5 words with integer
    # counts (4 on a tree root whose bin counts are reused); one conflict-free 32-lane ATOMS per cycle and SM is what
    # tools/ubench_smem.cu measured as the upper limit (profiles/r01_ubench_smem_atomics.txt: 31.97 words/cycle/SM).
    root_share = float(a.steps) * rows_local / max(float(rows_streamed), 1.0)
    words_per_cell = 5.0 - min(max(root_share, 0.0), 1.0)
    sm_clock_hz = (clocks or {}).get("sm_mhz") or 1965.0
    atom_peak = 148 * 32 * sm_clock_hz * 1e6 / 1e9            # G word-updates/s
    atom_achieved = float(rows_streamed) * a.features * words_per_cell / (hist_ms * 1e-3) / 1e9
    line = {
        "metric": "binned rows*features/sec histogram build (GBT tree levels, incl. routing, residual update, best split)",
        "value": value,
        "unit": "rows*features/s",
        "n_gpus": n_gpus,
        "steps": a.steps,
        "warmup": a.warmup,
        "ms_per_step": ms / a.steps,
        "sec_per_tree": ms / a.steps / 1e3,
        "higher_is_better": True,
        "scaling": "strong",
        "vs_baseline": None,
        "dtype": "u8 bins -> i64 fixed-point accumulators -> f64 statistics",
        "data": "synthetic",
        "config": {"workload": "C2: synthetic %d rows x %d features x %d bins, GBT squared loss depth %d (level-wise, %d stats passes/tree)"
                               % (a.rows, a.features, a.bins, a.depth, passes // max(a.steps, 1)),
                   "parallelism": "rows sharded over %d GPU(s); integer all-reduce of node stats" % n_gpus,
                   "cache": "inputs (%.1f GB/GPU) far larger than L2; no flush needed" % (rows_local * (a.features + 24) / 1e9),
                   "seed": SEED},
        "passes_per_step": passes / max(a.steps, 1),
        "hist_ms_per_pass": avg_hist_ms,
        "hist_share_of_step": hist_ms / ms,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "frac_streamed": achieved / peak, "frac_effective": achieved_effective / peak,
                     "achieved_effective": achieved_effective,
                     "rows_streamed_per_pass": float(rows_streamed) / max(hist_passes, 1),
                     "traffic": measured_traffic(a, rows_local),
                     "traffic_source": "committed ncu capture (profiles/r02_hist_rows_traffic.json), not measured in this run",
                     "peak_source": peak_kind,
                     "algorithmic_bytes_per_row": bytes_per_row, "algorithmic_bytes_per_launch": algo_bytes,
                     "second_bound": {"bound": "smem_atomics", "achieved": atom_achieved, "peak": atom_peak,
                                      "unit": "G word-updates/s", "frac": atom_achieved / atom_peak,
                                      "words_per_cell": words_per_cell,
                                      "peak_source": "one conflict-free 32-lane shared-memory atomic per cycle and SM at the sampled SM clock (tools/ubench_smem.cu)"},
                     "note": "frac = frac_streamed: bytes of the rows the statistics kernels read / their time (row-list build included) / measured HBM peak; "
                             "frac_effective credits every pass with all rows (histogram subtraction skips the larger child). "
                             "The kernel is bound by shared-memory atomics, not HBM: see second_bound and DESIGN.md 4.1"},
        "e2e": {"value": e2e_value, "unit": "rows*features/s",
                "h2d_bytes_per_step": (h2d_b - h2d_a) / max(a.steps, 1),
                "d2h_bytes_per_step": (d2h_b - d2h_a) / max(a.steps, 1),
                "ms_per_step": ms_e2e / a.steps},
        "gpu_launches": int(l1 - l0),
        "clocks": clocks,
        "train_error_last": (scalars[2] / scalars[0]) if scalars and scalars[0] else None,
        "trees_sha256_16": digest,
        "trees_hashed": n_trees_hashed,
        "other_configs": others,
    }
    if n_gpus == 1 and not a.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(a, a.cpu_sample_rows, steps=5, warmup=1)  # about 10 s with all cores + one tree with 4 threads
    if dist_ctx is not None:
        dist_ctx.destroy_process_group()
    OUT.emit(line) if OUT is not None else print(json.dumps(line))


def oracle_tree(orc, cfg, max_depth, tree_index, prev_root, threads):
    """One level-wise GBT tree grown with oracle calls only (DTWorker.doCompute + DTMaster best split restated)."""
    from shifu_b200.messages import IterIn, Node
    root = Node(1)
    todo_nodes = [root]
    switch = True
    passes = 0
    while todo_nodes:
        todo = [(k, tree_index, nd.id, []) for k, nd in enumerate(todo_nodes)]
        sent = [(tree_index - 1, prev_root), (tree_index, root)] if (switch and prev_root is not None) else [(tree_index, root)]
        it = IterIn(cfg, sent, todo, switch_to_next_tree=switch,
                    is_first_tree=(tree_index == 0) or (switch and tree_index == 1))
        res = orc.compute(it, n_threads=threads)
        gains, _ = orc.best_splits(it, res[4])
        passes += 1
        nxt = []
        for k, nd in enumerate(todo_nodes):
            g = gains[k]
            if not g.found:
                continue
            level = int(nd.id).bit_length()
            nd.predict = g.predict
            if g.gain <= 0 or level == max_depth:
                continue
            nd.feature, nd.split_type = g.feature, g.split_type
            nd.threshold = float(cfg.boundaries_of(g.feature)[g.split_bin + 1])
            nd.left, nd.right = Node(2 * nd.id, g.left_predict), Node(2 * nd.id + 1, g.right_predict)
            if level + 1 != max_depth:
                if not (g.left_impurity == 0.0 and not np.signbit(g.left_impurity)):
                    nxt.append(nd.left)
                if not (g.right_impurity == 0.0 and not np.signbit(g.right_impurity)):
                    nxt.append(nd.right)
        todo_nodes = nxt
        switch = False
    return root, passes


def cpu_baseline(a, sample_rows, steps=1, warmup=0, default_threads_steps=1):
    """The reference path restated on the CPU (oracle/) on a bounded sample of the workload: with all host threads, and
    with the reference's default workerThreadCount = 4 (container/obj/ModelTrainConf.java:143).  The sample is smaller
    than BASELINE.md 3 planned (<= 20M rows) so that both figures fit in the default run: 4 threads need about 8 s per
    tree of 2M rows."""
    from oracle.pyoracle import OracleWorker
    from shifu_b200 import synthetic
    from shifu_b200.capi import ConfigArrays
    cores = os.cpu_count() or 1
    cfg = ConfigArrays([a.bins] * a.features, max_depth=a.depth, tree_num=steps + warmup + default_threads_steps + 1,
                       learning_rate=0.05, min_instances_per_node=5)
    cols, y, sig = synthetic.make(SEED, [a.bins] * a.features, 0, sample_rows)
    orc = OracleWorker(cfg)
    orc.load_rows(cols, y, sig, None)
    prev = None
    t_idx = 0
    for _ in range(warmup):
        prev, _ = oracle_tree(orc, cfg, a.depth, t_idx, prev, cores)
        t_idx += 1

    def timed(n_trees, threads):
        nonlocal prev, t_idx
        t0 = time.perf_counter()
        passes = 0
        for _ in range(n_trees):
            prev, p = oracle_tree(orc, cfg, a.depth, t_idx, prev, threads)
            passes += p
            t_idx += 1
        dt = time.perf_counter() - t0
        return {"value": float(sample_rows) * a.features * passes / max(dt, 1e-9), "unit": "rows*features/s", "cores": threads,
                "seconds": dt, "ms_per_step": dt * 1e3 / max(n_trees, 1), "trees": n_trees, "passes": passes}

    full = timed(steps, cores)
    four = timed(default_threads_steps, min(4, cores)) if default_threads_steps > 0 else None
    orc.close()
    full.update({
        "kind": "port",
        "sample": "%d rows x %d features x %d bins of the same generator, %d GBT tree(s) of depth %d (%d passes), "
                  "C++ restatement of DTWorker.doCompute with workerThreadCount=%d (phase A single-threaded as in the reference); "
                  "the Java original cannot run here (no JVM)" % (sample_rows, a.features, a.bins, steps, a.depth, full["passes"], cores),
        "reference_default_threads": four,
    })
    return full


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    base = cpu_baseline(a, a.cpu_sample_rows, steps=a.steps, warmup=min(a.warmup, 1), default_threads_steps=0)
    line = {
        "impl": "reference",
        "metric": "binned rows*features/sec histogram build (GBT tree levels, incl. routing, residual update, best split)",
        "value": base["value"],
        "unit": "rows*features/s",
        "n_gpus": a.gpus,
        "steps": a.steps,
        "warmup": min(a.warmup, 1),
        "ms_per_step": base["ms_per_step"],
        "higher_is_better": True,
        "scaling": "strong",
        "vs_baseline": None,
        "dtype": "i16 bins -> f64 statistics (CPU)",
        "data": "synthetic",
        "config": {"workload": "C2 sample: %d of %d rows x %d features x %d bins, GBT squared loss depth %d"
                               % (a.cpu_sample_rows, a.rows, a.features, a.bins, a.depth), "seed": SEED},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": "rows*features/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    OUT.emit(line) if OUT is not None else print(json.dumps(line))


class OneLineStdout:
    """stdout carries ONE JSON line.  Libraries write there too (NCCL prints its version banner with printf when a
    communicator is created), so file descriptor 1 points at stderr until the line is ready."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def emit(self, line):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)
        return False


OUT = None

if __name__ == "__main__":
    args = parse()
    with OneLineStdout() as OUT:
        if args.impl == "reference":
            run_reference(args)
        else:
            run_b200(args)
