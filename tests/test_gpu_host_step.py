"""GPU tier: step_host (bcp_step_host: host buffers, observation slabs copied repeat by repeat beside the
kernels) returns exactly what step() returns, on the split pipeline and on the one-launch kernel."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.mark.parametrize("n,dtype,kw", [(8192, "fp32", dict(auto_reset="pipelined", max_episode_len=7)),
                                        (8192, "fp32", dict(auto_reset="pipelined", max_episode_len=7, action_repeats=3, steps_per_repeat=2)),
                                        (300, "fp64", dict(max_episode_len=9)),
                                        (4096, "fp32", dict(auto_reset="pipelined", max_episode_len=5, action_repeats=1))])
def test_step_host_equals_step(built, n, dtype, kw):
    from cartpoleplusplus_b200 import default_opts
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    opts = default_opts(action_force=100.0, initial_force=55.0, **kw)
    a = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=n, device=0, dtype=dtype, seed=4)
    b = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=n, device=0, dtype=dtype, seed=4)
    assert torch.equal(a.reset(), b.reset())
    rng = np.random.default_rng(0)
    fresh = 0
    for t in range(90 if kw.get("auto_reset") else 12):
        act = torch.from_numpy(rng.uniform(-1, 1, size=(n, 2)).astype(np.float32))
        obs, rew, done, info = a.step(act.cuda())
        h_obs, h_rew, h_done, h_info = b.step_host(act.pin_memory() if t % 2 else act)  # pinned and pageable callers
        assert h_obs.shape == obs.shape
        assert torch.equal(h_obs, obs.cpu()), t
        assert torch.equal(h_rew, rew.cpu()) and torch.equal(h_done.bool(), done.cpu())
        assert torch.equal(h_info["done_reason"], info["done_reason"].cpu())
        fresh += int((info["done_reason"] == 5).sum())
    if kw.get("auto_reset") and a.repeats > 1:
        assert fresh > 0  # the reset observations that arrive after the early slabs were exercised
    a.close()
    b.close()
