"""Throughput of the device scorer (f2) on the reference's own 100-tree GBT fixture: rows/s with the rows resident in
HBM (kernel time from CUDA events) and end to end from host memory, next to the CPU oracle on a sample."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from shifu_b200 import model_io  # noqa: E402
from shifu_b200.capi import Scorer  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
m = model_io.read_model(os.path.join(ROOT, "tests", "golden", "readablespec_model0.gbt"))
sm = model_io.to_score_model(m)
rng = np.random.default_rng(1)
x = rng.normal(size=(n, sm.n_columns)) * 30.0 + 10.0
sc = Scorer(sm)
sc.score(x[:1000])
t0 = time.perf_counter()
got = sc.score(x)
t1 = time.perf_counter()
k_ms = sc.last_kernel_ms()
sample = min(n, 200_000)
t2 = time.perf_counter()
ref = pyoracle.score_rows(sm, x[:sample])
t3 = time.perf_counter()
assert np.array_equal(got[:sample], ref)
bytes_alg = n * (sm.n_columns * 8 + 8)
print("scoring %d rows x %d columns, %d trees (%d nodes): kernel %.3f ms = %.3e rows/s, %.1f GB/s algorithmic; "
      "end to end from pageable host memory %.3f s = %.3e rows/s; CPU oracle (1 thread) %.3e rows/s"
      % (n, sm.n_columns, sm.n_trees, len(sm.nodes), k_ms, n / (k_ms * 1e-3), bytes_alg / (k_ms * 1e-3) / 1e9,
         t1 - t0, n / (t1 - t0), sample / (t3 - t2)))
