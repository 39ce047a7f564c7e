"""Env sharding over ranks and the single collective of the path.

Environments never interact: rank g of G owns the global env ids
[g*B/G, (g+1)*B/G); the in-kernel RNG is keyed by GLOBAL env id, so trajectories - and the
merged statistics - do not depend on G.  Nothing is exchanged on the step path; the only
collective is one all_reduce(SUM) of the episode-length histogram and counters (NCCL over
NVLink on GPUs; gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Tuple

import numpy as np
import torch
import torch.distributed as dist


def shard(num_envs_total: int, rank: int, world: int) -> Tuple[int, int]:
    """(env_id_offset, num_envs) of `rank`: contiguous blocks, remainder to the low ranks."""
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    base, rem = divmod(int(num_envs_total), int(world))
    n = base + (1 if rank < rem else 0)
    off = rank * base + min(rank, rem)
    return off, n


def allreduce_stats_arrays(hist: np.ndarray, header: np.ndarray, reward_sum: float, device=None):
    """One all_reduce(SUM) over [hist | header | reward_sum]; returns the merged triple.
    Works without an initialised process group (world of 1)."""
    hist = np.asarray(hist, dtype=np.int64)
    header = np.asarray(header, dtype=np.int64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return hist.copy(), header.copy(), float(reward_sum)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    # the counts travel as int64 (exact); the reward sum rides along as its float64 bit pattern
    # is not summable, so it goes in a second, 1-element float64 reduce fused by the coalescing
    # manager when available - two tiny messages at worst
    packed = torch.from_numpy(np.concatenate([hist, header])).to(device)
    rs = torch.tensor([reward_sum], dtype=torch.float64, device=device)
    dist.all_reduce(packed, op=dist.ReduceOp.SUM)
    dist.all_reduce(rs, op=dist.ReduceOp.SUM)
    packed = packed.cpu().numpy()
    return packed[:hist.size].copy(), packed[hist.size:].copy(), float(rs.item())


def allreduce_stats(env):
    """Merged (hist, header dict, reward_sum) of a sharded BatchedBulletCartpole."""
    hist, header, rsum = env.stats()
    names = list(header.keys())
    h, hd, rs = allreduce_stats_arrays(hist, np.array([header[k] for k in names], dtype=np.int64), rsum,
                                       device=getattr(env, "device", None))
    return h, dict(zip(names, (int(v) for v in hd))), rs
