"""Host-side twin of the device generator of SURVEY.md 8(d) (rows.cu synthetic_*_kernel): every value is a pure
function of (seed, global row, feature), so any shard of the data set can be produced anywhere.

Bins are bit-identical to the device generator.  Labels use float32 exp and can differ from the device in a
handful of rows per million (different exp implementations); code that compares against the GPU reads the
labels back with sb_read_labels instead.
"""
import numpy as np

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M
    x = ((x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
    x = ((x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
    return x ^ (x >> np.uint64(31))


def bins_column(seed, feature, n_features, n_bins, row0, n_rows):
    with np.errstate(over="ignore"):
        r = np.arange(row0, row0 + n_rows, dtype=np.uint64)
        h = splitmix64(np.uint64(seed) ^ (r * np.uint64(n_features) + np.uint64(feature)))
    return ((h >> np.uint64(33)).astype(np.uint32) % np.uint32(n_bins)).astype(np.uint16 if n_bins > 256 else np.uint8)


def labels(seed, cols, n_bins, row0, n_rows, label_kind=0, weighted=False):
    with np.errstate(over="ignore"):
        r = np.arange(row0, row0 + n_rows, dtype=np.uint64)
        h = splitmix64(np.uint64(seed) ^ np.uint64(0xA5A5A5A5DEADBEEF) ^ r)
    u = (h >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    u2 = ((h >> np.uint64(16)) & np.uint64(0xFFFFFF)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    F = len(cols)
    if label_kind == 0:
        z = np.zeros(n_rows, dtype=np.float32)
        for f in range(min(F, 5)):
            # the device label kernel reads the uint8 matrix, where a wide (> 256 bins) feature is stored as 0
            c = cols[f] if n_bins[f] <= 256 else np.zeros(n_rows, dtype=np.uint8)
            x = c.astype(np.float32) / np.float32(n_bins[f]) - np.float32(0.5)
            z += np.float32(-2.0 if f & 1 else 3.0) * x
        p = np.float32(1.0) / (np.float32(1.0) + np.exp(-z, dtype=np.float32))
        y = (u < p).astype(np.float32)
        y = np.where(u2 < np.float32(0.1), np.float32(1.0) - y, y).astype(np.float32)
    else:
        b0 = cols[0].astype(np.float32)
        b1 = cols[1].astype(np.float32) if F > 1 else 0.0
        b2 = cols[2].astype(np.float32) if F > 2 else 0.0
        nz = np.float32(2.0) * u - np.float32(1.0)
        y = (((b0 + b1 + np.float32(2.0) * b2) * np.float32(0.25) + np.float32(4.0) * nz)
             > np.float32(0.5) * np.float32(n_bins[0])).astype(np.float32)
    sig = (np.float32(0.5) + np.float32(1.5) * u2).astype(np.float32) if weighted else np.ones(n_rows, np.float32)
    return y, sig


def make(seed, n_bins, row0, n_rows, label_kind=0, weighted=False):
    F = len(n_bins)
    if n_rows * F >= 50_000_000:  # numpy releases the GIL inside the hash arithmetic: one column per thread
        import os
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(32, os.cpu_count() or 1)) as ex:
            cols = list(ex.map(lambda f: bins_column(seed, f, F, int(n_bins[f]), row0, n_rows), range(F)))
    else:
        cols = [bins_column(seed, f, F, int(n_bins[f]), row0, n_rows) for f in range(F)]
    y, sig = labels(seed, cols, n_bins, row0, n_rows, label_kind, weighted)
    return cols, y, sig


def poisson_bags(seed, tree, rate, row0, n_rows):
    """Host twin of sb_sample_poisson (prep.cu poisson_bags_kernel): inverse Poisson CDF of a counter-based uniform."""
    import math
    with np.errstate(over="ignore"):
        r = np.arange(row0, row0 + n_rows, dtype=np.uint64)
        key = (r * np.uint64(0xD1342543DE82EF95) + np.uint64(tree) * np.uint64(0x9E3779B97F4A7C15)) & _M
        u = (splitmix64(np.uint64(seed) ^ key) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    p0 = math.exp(-rate)
    k = np.zeros(n_rows, dtype=np.int64)
    pk = np.full(n_rows, p0)
    cdf = np.full(n_rows, p0)
    for _ in range(1000):
        more = u > cdf
        if not more.any():
            break
        k[more] += 1
        pk[more] = pk[more] * rate / k[more]
        cdf[more] = cdf[more] + pk[more]
    return k.astype(np.float32)
