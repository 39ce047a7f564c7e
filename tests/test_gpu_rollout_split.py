"""-m gpu: bcp_rollout_random through the split pipeline (fp32, >= 4096 envs: k_draw_actions + the pipeline's kernels per
call) against the one-launch k_rollout (BCP_ROLLOUT_SPLIT=0): the same machine and the same Philox action stream, so every
counter identity holds on both and the episode statistics agree (not bit for bit: the two fp32 paths round differently)."""
import numpy as np
import pytest
import torch

from cartpoleplusplus_b200 import default_opts

pytestmark = pytest.mark.gpu


def _run(n, discrete, split, monkeypatch, **ov):
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    monkeypatch.setenv("BCP_ROLLOUT_SPLIT", "1" if split else "0")
    env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined", **ov), discrete_actions=discrete,
                                num_envs=n, device=0, dtype="fp32", seed=11)
    calls = 0
    header = {"episodes": 0}
    while header["episodes"] < n and calls < 12:
        env.rollout_random(64, max_episodes=1)
        calls += 1
        hist, header, rsum = env.stats()
    launches = env.launch_count
    env.close()
    return hist, header, rsum, launches


@pytest.mark.parametrize("discrete,ov", [(False, dict(initial_force=55.0)), (True, dict(initial_force=55.0))])
def test_one_episode_per_env_on_both_paths(built, monkeypatch, discrete, ov):
    n = 8192
    hs, cs, rs, ls = _run(n, discrete, True, monkeypatch, **ov)
    ho, co, ro, lo = _run(n, discrete, False, monkeypatch, **ov)
    assert ls > lo * 10  # the split path launched its kernels (one k_rollout per call otherwise)
    for hist, c, r in ((hs, cs, rs), (ho, co, ro)):
        assert c["episodes"] == n == int(hist.sum())                      # exactly one episode per env, then parked
        assert c["env_steps"] == int((hist * np.arange(len(hist))).sum())  # every live step is in some episode's length
        assert r == float(c["env_steps"])                                 # reward 1 per live step
        assert c["resets"] == 0 and c["substeps"] == 2 * c["env_steps"]    # the constructor reset the envs; R * S = 2 substeps per live step
    # same statistics: deciles of the episode lengths within 2 steps, means within 2 %
    def deciles(h):
        lens = np.repeat(np.arange(len(h)), h)
        return np.percentile(lens, np.arange(0, 101, 10))
    ds, do = deciles(hs), deciles(ho)
    assert np.abs(ds - do).max() <= 2, (ds, do)
    assert abs(cs["env_steps"] - co["env_steps"]) <= 0.02 * co["env_steps"]
