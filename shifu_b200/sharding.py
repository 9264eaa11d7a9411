"""Row sharding and the per-pass collectives of the multi-GPU path (one process per GPU, torch.distributed).

The reference shards rows over Guagua workers and the master sums their NodeStats (dt/DTMaster.java:298-315).  Here
every rank holds a contiguous shard of the rows; per statistics pass the ranks exchange
  * the three scale maxima (MAX)            -- only when a pass recomputed them (tree switch),
  * the integer accumulators of the pass (SUM)  -- the DTMaster merge,
  * the four error/count scalars (SUM)      -- only when a pass recomputed them,
after which every rank holds identical numbers and takes identical split decisions, so nothing is broadcast.
"""


def shard_rows(n_rows, world, rank):
    """Contiguous shard [row0, row0 + n) of rank `rank`; the first n_rows % world ranks get one extra row."""
    base, extra = divmod(n_rows, world)
    n = base + (1 if rank < extra else 0)
    row0 = rank * base + min(rank, extra)
    return row0, n


def reduce_pass(dist, scales, raw, acc, scales_fresh, acc_fresh):
    """The collectives of one pass on already-wrapped tensors (any backend: NCCL on the GPUs, gloo in the CPU tests).
    `scales` int32[3] bit patterns of non-negative floats (integer MAX == float MAX), `raw` int64, `acc` float64[4]."""
    if scales_fresh and scales is not None:
        dist.all_reduce(scales, op=dist.ReduceOp.MAX)
    if raw is not None and raw.numel() > 0:
        dist.all_reduce(raw, op=dist.ReduceOp.SUM)
    if acc_fresh and acc is not None:
        dist.all_reduce(acc, op=dist.ReduceOp.SUM)
