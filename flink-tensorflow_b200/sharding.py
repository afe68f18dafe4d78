"""How the path shards across GPUs: by record, with no collective on the data path.

One process (or Flink subtask) per GPU.  Records are dealt round-robin -- Flink's rebalance partitioner, which is what
sits in front of a parallel mapWithModel (FT/streaming/package.scala:23-34) -- subtask i serves GPU i % n_gpus, weights are
replicated.  torch.distributed is used only around the timed region: a barrier before and after, and a MAX reduction of the
per-rank device time.  Nothing in this module touches CUDA, so it is testable with the gloo backend on CPU.  torch is
imported only inside the two reductions (bench.py and the gloo test call them); the partitioning helpers are plain Python.
"""


def shard_indices(n_records, rank, world):
    """Indices of the records rank `rank` of `world` processes handles (round-robin: i % world == rank)."""
    if not (0 <= rank < world):
        raise ValueError("rank must be in [0, world)")
    return range(rank, n_records, world)


def device_for_subtask(subtask_index, n_gpus):
    """GPU ordinal of an operator subtask (getRuntimeContext.getIndexOfThisSubtask % visible GPUs)."""
    if n_gpus <= 0:
        raise ValueError("no GPU visible")
    return subtask_index % n_gpus


def merge_sharded(parts, n_records):
    """Inverse of shard_indices: parts[r] holds the results of rank r in its own order; returns them in record order."""
    world = len(parts)
    out = [None] * n_records
    for r, part in enumerate(parts):
        idx = shard_indices(n_records, r, world)
        if len(part) != len(idx):
            raise ValueError(f"rank {r} returned {len(part)} results for {len(idx)} records")
        for i, v in zip(idx, part):
            out[i] = v
    return out


def max_over_ranks(value, device="cpu"):
    """The slowest rank's time: what a multi-GPU step takes."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_throughput(units_per_rank, ms_per_step_this_rank, device="cpu"):
    """Whole-job throughput: every rank's units per step over the slowest rank's step time."""
    import torch.distributed as dist
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    ms = max_over_ranks(ms_per_step_this_rank, device)
    return units_per_rank * world / (ms / 1e3), ms
