"""-m gpu: the split pipeline (k_prep / k_solve, bcp_split.cuh) against the one-launch fast
kernel and the oracle, and the pipelined auto-reset machine against a host-driven oracle."""
import os

import numpy as np
import pytest

from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200 import config as C

pytestmark = pytest.mark.gpu


def _env(n, split, dtype="fp32", discrete=False, **ov):
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    os.environ["BCP_SPLIT"] = "1" if split else "0"
    try:
        return BatchedBulletCartpole(default_opts(**ov), discrete_actions=discrete, num_envs=n, device=0, dtype=dtype, seed=11)
    finally:
        os.environ.pop("BCP_SPLIT", None)


@pytest.mark.parametrize("ov", [dict(initial_force=55.0, action_force=100.0),
                                dict(initial_force=55.0, action_force=100.0, cart_friction=1.0, steps_per_repeat=3),
                                dict(initial_force=200.0, action_force=100.0, auto_reset="pipelined")])
def test_split_equals_one_launch_kernel(built, ov):
    """Same per-env functions in the same order: the two fast paths agree to rounding (the
    compiler may contract FMAs differently in the two kernels) on every env and every output."""
    import torch
    n = 3000  # not a multiple of the block size: tail lanes
    a, b = _env(n, False, **ov), _env(n, True, **ov)
    oa, ob = a.reset(), b.reset()
    assert torch.equal(oa, ob)
    rng = np.random.default_rng(2)
    worst = 0.0
    for t in range(12):
        act = torch.from_numpy(rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)).cuda()
        # identical starting states every step, so that differences do not compound chaotically
        b.set_state(a.get_state())
        ra, rb = a.step(act), b.step(act)
        assert torch.equal(ra[2], rb[2]) or (ra[2] != rb[2]).float().mean() < 2e-3
        assert torch.equal(ra[1], rb[1])
        d = (ra[0] - rb[0]).abs().reshape(n, -1).max(dim=1).values
        worst = max(worst, float(d.median()))
        assert float(d.median()) <= 1e-6 and float(torch.quantile(d, 0.99)) <= 1e-4, (t, float(d.max()))
        sa, sb = a.get_state(), b.get_state()
        assert np.array_equal(sa[:, C.S_STEPS:C.S_RESET_LEFT + 1], sb[:, C.S_STEPS:C.S_RESET_LEFT + 1]) or \
            (sa[:, C.S_DONE] != sb[:, C.S_DONE]).mean() < 2e-3
    a.close(); b.close()


def test_pipelined_auto_reset_matches_host_driven_oracle(built):
    """fp64 (bit-exact path): an env that finishes runs its reset R*S substeps per call and comes
    back FRESH with the observation the oracle's synchronous reset() returns; live envs match the
    oracle step for step."""
    import torch
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    from oracle import Oracle
    n = 96
    opts = default_opts(initial_force=55.0, action_force=100.0, max_episode_len=7, steps_per_repeat=3,
                        auto_reset="pipelined")
    env = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=n, device=0, dtype="fp64", seed=5)
    ocfg = C.copy_config(env.cfg)
    ocfg.auto_reset = C.AUTO_RESET_OFF
    ora = Oracle(ocfg, dtype="fp64")
    assert np.array_equal(env.reset().cpu().numpy(), ora.reset())
    rng = np.random.default_rng(9)
    n_fresh = 0
    calls_resetting = np.zeros(n, dtype=int)
    for t in range(80):
        a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
        obs, rew, done, info = env.step(torch.from_numpy(a).cuda())
        reason = info["done_reason"].cpu().numpy()
        obs, rew, done = obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy()
        fresh = reason == C.REASON_FRESH
        resetting = reason == C.REASON_RESETTING
        live = ~(fresh | resetting)
        obs_o, rew_o, done_o, reason_o = ora.step(a)  # done envs: old obs, reward 0 (:179-181)
        assert np.array_equal(obs[live], obs_o[live]) and np.array_equal(done[live], done_o[live])
        assert np.array_equal(reason[live], reason_o[live]) and (rew[live] == 1).all()
        assert (rew[~live] == 0).all() and done[resetting].all() and not done[fresh].any()
        calls_resetting[resetting] += 1
        if fresh.any():
            o1 = ora.reset(fresh.astype(np.uint8))
            assert np.array_equal(obs[fresh], o1[fresh])
            # 130 substeps at 6 per call: 21 RESETTING calls, FRESH on the 22nd
            assert (calls_resetting[fresh] == 21).all()
            calls_resetting[fresh] = 0
            n_fresh += int(fresh.sum())
    assert n_fresh > n  # every env went round at least once
    st = env.get_state()
    ok = st[:, C.S_RESET_LEFT] == 0
    assert np.array_equal(st[ok][:, :C.S_RESET_LEFT], ora.state[ok][:, :C.S_RESET_LEFT])
    env.close()
