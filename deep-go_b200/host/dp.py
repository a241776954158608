"""Host side of synchronous data parallelism (replaces makeDataParallel / nn.DataParallelTable(1),
experiments.lua:155-168): one process per GPU, the batch is split on dim 0 exactly as DataParallelTable(1)
does, every rank scales its loss gradient by 1/global_batch and the library sums gradients over ranks."""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous dim-0 slice of rank `rank` (DataParallelTable splits the batch into `world` near-equal chunks)."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_batch(batch, rank, world):
    """batch: dict of arrays with a leading batch dim -> this rank's slice of every array."""
    n = len(next(iter(batch.values())))
    lo, hi = shard_range(n, rank, world)
    return {k: v[lo:hi] for k, v in batch.items()}


def exchange_unique_id(dist, rank, make_id):
    """Rank 0 creates the 128-byte communicator id (engine.dp_unique_id) and ships it over the existing
    torch.distributed group (gloo or nccl); every rank returns the same bytes."""
    obj = [make_id() if rank == 0 else None]
    dist.broadcast_object_list(obj, src=0)
    assert isinstance(obj[0], (bytes, bytearray)) and len(obj[0]) == 128
    return bytes(obj[0])


def init_data_parallel(ctx, model, dist, rank, world, global_batch):
    """Bind the library's NCCL communicator for this rank and set the 1/global_batch loss scale."""
    from . import engine
    if world > 1:
        ctx.init_dp(rank, world, exchange_unique_id(dist, rank, engine.dp_unique_id))
    model.set_global_batch(global_batch)
