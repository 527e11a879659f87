"""Host-side multi-process plumbing (one process per GPU, torch.distributed): the two scalar exchanges of the path and
the benchmark's sharding / timing reductions.  Device-agnostic so the logic is testable with gloo on CPU."""
import torch
import torch.distributed as dist


def is_dist():
    return dist.is_available() and dist.is_initialized()


def global_num_boxes(local_count, distributed=True):
    """set_criterion.py:185-193: num_boxes = clamp(sum over ranks of the local target count / world_size, min=1).
    `local_count` is a 1-element float tensor; the result stays on its device (no .item())."""
    n = local_count.clone()
    if distributed and is_dist():
        dist.all_reduce(n)
        n = n / dist.get_world_size()
    return n.clamp(min=1.0)


def max_over_ranks(value, device="cpu"):
    """Timing reduction of bench.py: every rank reports its device-side elapsed time, the job's time is the max."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if is_dist():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def shard_seed(rank, index):
    """Synthetic-clip seed of the `index`-th batch of rank `rank`: disjoint streams per rank (weak scaling: every rank
    processes its own clips, as DistributedSampler shards a dataset, runs/_loader.py:99-101)."""
    return 1000 * rank + index


def reduce_stats(stats):
    """utils/distributed.py:63-75 reduce_distrib_loss: average the 5 logged scalars over ranks in ONE all-reduce."""
    keys = sorted(stats)
    t = torch.stack([stats[k].detach().float() for k in keys])
    if is_dist():
        dist.all_reduce(t)
        t = t / dist.get_world_size()
    return {k: t[i] for i, k in enumerate(keys)}
