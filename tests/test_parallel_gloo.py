"""CPU tier, world_size 2 over gloo: the N>1 host logic of the path - shard by global env id,
run each shard, one all_reduce of the statistics - must reproduce the single-rank histogram
exactly (the RNG is keyed by GLOBAL env id, so trajectories do not depend on the sharding).
The shards run the oracle here (no GPU in this tier); -m gpu repeats it on the CUDA path."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rollout(offset, n, total_cfg_kwargs):
    from cartpoleplusplus_b200 import default_opts, make_config, config as C
    from oracle import Oracle
    cfg = make_config(default_opts(**total_cfg_kwargs), True, num_envs=n, dtype="fp64", seed=7, env_id_offset=offset)
    ora = Oracle(cfg, dtype="fp64")
    ora.state[:, C.S_DONE] = 1
    ora.rollout_random(400, action_mask=0x1F, max_episodes=3)
    return ora.hist.copy(), ora.header.copy(), float(ora.reward_sum[0])


def _worker(rank, world, port, total, kwargs, out_path):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from cartpoleplusplus_b200.parallel import allreduce_stats_arrays, shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    off, n = shard(total, rank, world)
    hist, header, rsum = _rollout(off, n, kwargs)
    h, hd, rs = allreduce_stats_arrays(hist, header, rsum)
    if rank == 0:
        np.savez(out_path, hist=h, header=hd, rsum=rs)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_histogram_equals_single_rank(tmp_path, built):
    total, kwargs = 48, dict(initial_force=55.0, max_episode_len=40)
    out = str(tmp_path / "merged.npz")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, total, kwargs, out), nprocs=2, join=True)
    got = np.load(out)
    hist1, header1, rsum1 = _rollout(0, total, kwargs)
    assert np.array_equal(got["hist"], hist1)
    assert np.array_equal(got["header"], header1)
    assert float(got["rsum"]) == rsum1
    assert hist1.sum() == 3 * total  # every env delivered its 3 episodes
