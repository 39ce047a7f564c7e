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
    # ONE message: counts and the reward sum travel together as float64 (the counts stay exact below 2^53)
    if max(int(hist.sum()), int(np.abs(header).max(initial=0))) >= 2 ** 53:
        raise OverflowError("statistics beyond 2^53 do not travel exactly as float64")
    packed = torch.from_numpy(np.concatenate([hist.astype(np.float64), header.astype(np.float64), [float(reward_sum)]])).to(device)
    dist.all_reduce(packed, op=dist.ReduceOp.SUM)
    packed = packed.cpu().numpy()
    nh, nd = hist.size, header.size
    return np.rint(packed[:nh]).astype(np.int64), np.rint(packed[nh:nh + nd]).astype(np.int64), float(packed[nh + nd])


def allreduce_stats(env):
    """Merged (hist, header dict, reward_sum) of a sharded BatchedBulletCartpole."""
    hist, header, rsum = env.stats()
    names = list(header.keys())
    h, hd, rs = allreduce_stats_arrays(hist, np.array([header[k] for k in names], dtype=np.int64), rsum,
                                       device=getattr(env, "device", None))
    return h, dict(zip(names, (int(v) for v in hd))), rs
