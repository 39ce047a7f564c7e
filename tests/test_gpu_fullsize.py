"""-m gpu: BASELINE.json's full batch (65 536 envs, configs[2]) through properties that do not
need the oracle at that size: bookkeeping identities between the per-call outputs and the
episode statistics, unit quaternions / finite poses, and independence of the sharding (bit
for bit).  (The scene's half-turn symmetry is NOT such a property: dBoxBox2's clipper and the
4-point manifold reduction depend on the order in which points are emitted, which is not
symmetric - measured, turned start states drift apart by ~5e-4 within three env.steps.)"""
import numpy as np
import pytest
import torch

from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200 import config as C

pytestmark = pytest.mark.gpu
N = 65536


def _env(n=N, offset=0, dtype="fp32", **ov):
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    opts = default_opts(action_force=100.0, **ov)
    return BatchedBulletCartpole(opts, discrete_actions=False, num_envs=n, device=0, dtype=dtype, seed=5, env_id_offset=offset)


def test_bookkeeping_identities_at_full_size(built):
    env = _env(auto_reset="pipelined", initial_force=200.0)
    env.reset(copy=False)
    env.rollout_random(150)            # into the stationary mix of episode phases
    h0, c0, _ = env.stats()
    g = torch.Generator(device="cuda")
    g.manual_seed(1)
    reward_sum, finished = 0.0, 0
    for _ in range(40):
        a = torch.rand((N, 2), device="cuda", generator=g) * 2 - 1
        obs, reward, done, info = env.step(a, copy=False)
        reason = info["done_reason"]
        reward_sum += float(reward.sum())
        live = reward > 0
        # reward 1 exactly on live steps; an episode ends only on a live step, with a reason 1..3
        assert bool(((reward == 0) | (reward == 1)).all())
        ended = live & done
        assert bool(((reason[ended] >= 1) & (reason[ended] <= 3)).all())
        assert bool((reason[live & ~done] == 0).all())
        assert bool(torch.isin(reason[~live], torch.tensor([0, 4, 5], device="cuda", dtype=reason.dtype)).all())
        finished += int(ended.sum())
        assert bool(torch.isfinite(obs).all())
        q = obs[..., 3:7]
        assert float((q.norm(dim=-1) - 1).abs().max()) < 1e-4
    h1, c1, _ = env.stats()
    assert c1["env_steps"] - c0["env_steps"] == int(round(reward_sum))
    assert c1["episodes"] - c0["episodes"] == finished
    assert int((h1 - h0).sum()) == finished


@pytest.mark.parametrize("N", [65536, 131072])
def test_trajectories_do_not_depend_on_the_sharding(built, N):
    """Whole batch vs two half-size handles with env_id_offset: bit for bit - whole and halves run the same per-env
    programs whatever the number of cohorts a handle is split into (bcp_abi.cu)."""
    exact = True  # every size runs the impulse-space solve phase since the handle is run as cohorts (round 2)
    whole = _env(N, initial_force=55.0)
    halves = [_env(N // 2, offset=0, initial_force=55.0), _env(N // 2, offset=N // 2, initial_force=55.0)]
    def same(x, y):
        if exact:
            return torch.equal(x, y)
        d = (x.float() - y.float()).abs().reshape(x.shape[0], -1).amax(dim=1)
        return float(d.median()) <= 1e-6 and float(torch.quantile(d[:60000], 0.99)) <= 1e-3

    o = whole.reset()
    oh = torch.cat([h.reset() for h in halves])
    assert torch.equal(o, oh)  # reset runs the one-launch kernel on every size
    g = torch.Generator(device="cuda")
    g.manual_seed(2)
    for _ in range(5):
        a = torch.rand((N, 2), device="cuda", generator=g) * 2 - 1
        o, r, d, _ = whole.step(a)
        parts = [h.step(a[i * (N // 2):(i + 1) * (N // 2)]) for i, h in enumerate(halves)]
        assert same(o, torch.cat([p[0] for p in parts]))
        if exact:
            assert torch.equal(d, torch.cat([p[2] for p in parts]))
        else:  # start every step from identical states, so that rounding does not compound chaotically
            st = whole.get_state()
            for i, h in enumerate(halves):
                h.set_state(st[i * (N // 2):(i + 1) * (N // 2)])


@pytest.mark.parametrize("n,cohorts", [(40037, 3), (5000, 4), (70001, 8)])
def test_cohort_split_of_an_odd_batch_is_invisible(built, monkeypatch, n, cohorts):
    """A batch whose size is no multiple of anything, run as one pipeline and as `cohorts` (BCP_COHORTS; the cohort bases
    are rounded to whole blocks, the last cohort takes the rest): same outputs bit for bit over env.steps with resets, and
    through bcp_step_host."""
    monkeypatch.setenv("BCP_COHORTS", "1")
    one = _env(n, initial_force=55.0, auto_reset="pipelined")
    monkeypatch.setenv("BCP_COHORTS", str(cohorts))
    many = _env(n, initial_force=55.0, auto_reset="pipelined")
    assert torch.equal(one.reset(), many.reset())
    for e in (one, many):
        e.rollout_random(30)  # into the mix of live / resetting envs
    assert np.array_equal(one.get_state(), many.get_state())
    g = torch.Generator(device="cuda")
    g.manual_seed(11)
    for t in range(6):
        a = torch.rand((n, 2), device="cuda", generator=g) * 2 - 1
        if t < 4:
            ra, rb = one.step(a), many.step(a)
        else:
            ha = a.cpu().pin_memory()
            ra, rb = one.step_host(ha), many.step_host(ha)
        for x, y in zip(ra[:3], rb[:3]):
            assert torch.equal(torch.as_tensor(x), torch.as_tensor(y)), t
    assert np.array_equal(one.get_state(), many.get_state())
