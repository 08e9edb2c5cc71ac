"""Multi-GPU sharding of the step: worlds are independent (no cross-world reads anywhere in
src/sim/sim.cpp), so rank r of k simply owns a contiguous range of GLOBAL world ids and runs its own
simulator on its own GPU. RNG streams are keyed on the global id (src/sim/sim.cpp:1238-1239), which
makes the result independent of k. The only exchange is a sum all-reduce of the 16-scalar
episode-statistics vector (NCCL over NVLink on GPUs; gloo in the CPU tests).
"""
from typing import List, Tuple

STAT_NAMES = ("agent_steps", "steps", "count_s1", "count_s2", "count_s3", "count_s4",
              "health_s1", "health_s2", "health_s3", "health_s4",
              "births", "deaths", "respawns", "overflow", "food_eaten", "shots_hit")


def shard_range(num_worlds: int, rank: int, world_size: int) -> Tuple[int, int]:
    """[begin, end) of the global world ids rank `rank` owns; the first `num_worlds % world_size` ranks
    own one extra world."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world_size")
    q, r = divmod(num_worlds, world_size)
    begin = rank * q + min(rank, r)
    return begin, begin + q + (1 if rank < r else 0)


def make_shard(sim_cls, gpu_id: int, total_worlds: int, rand_seed: int, init_agents: int, rank: int,
               world_size: int, **kw):
    """Construct this rank's simulator over its world range. `sim_cls` has SimManager's signature."""
    b, e = shard_range(total_worlds, rank, world_size)
    return sim_cls(gpu_id, e - b, rand_seed, init_agents, world_id_offset=b, **kw), (b, e)


def allreduce_stats(stats: List[int], device=None, group=None) -> List[int]:
    """Sum the episode-statistics vector over all ranks. `steps` (index 1) is identical on every rank
    and is therefore divided back; everything else is additive across shards."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return list(stats)
    t = torch.tensor(list(stats), dtype=torch.int64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    out = t.tolist()
    out[1] //= dist.get_world_size(group)
    return out


def allreduce_max(value: float, device=None, group=None) -> float:
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
