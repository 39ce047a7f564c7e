"""-m gpu: --reset-settle-cache (with --no-carry-over-reset): the 100 settle substeps of reset() computed once per handle and
replayed.  In that mode every env settles from the same state (teleport, pending force and contact caches cleared,
bullet_cartpole.py:319-327 minus Bullet's carry-over), so the replay is the same computation: the fp64 kernel must stay bit
for bit on the oracle, which runs the settle loop every time."""
import numpy as np
import pytest

from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200 import config as C

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _mk(n, dtype, **ov):
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    return BatchedBulletCartpole(default_opts(**ov), discrete_actions=False, num_envs=n, device=0, dtype=dtype, seed=5)


def test_flag_needs_the_no_carry_over_mode(built):
    with pytest.raises(ValueError, match="no-carry-over-reset"):
        _mk(8, "fp64", reset_settle_cache=True)


def test_fp64_synchronous_resets_bit_exact_with_the_oracle(built):
    from oracle import Oracle
    n = 200
    env = _mk(n, "fp64", initial_force=55.0, action_force=100.0, no_carry_over_reset=True, reset_settle_cache=True, max_episode_len=9)
    ora = Oracle(env.cfg, dtype="fp64")   # runs all 100 settle substeps on every reset
    assert np.array_equal(env.reset().cpu().numpy(), ora.reset())
    assert np.array_equal(env.get_state(), ora.state)
    rng = np.random.default_rng(1)
    for t in range(30):
        a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
        obs, rew, done, info = env.step(torch.from_numpy(a).cuda())
        o = ora.step(a)
        assert np.array_equal(obs.cpu().numpy(), o[0]) and np.array_equal(done.cpu().numpy(), o[2]), t
        if t % 7 == 6 and o[2].any():
            m = o[2].astype(np.uint8)
            assert np.array_equal(env.reset(torch.from_numpy(m).cuda()).cpu().numpy(), ora.reset(m))
    assert np.array_equal(env.get_state(), ora.state)
    env.close()


def test_fp64_pipelined_reset_takes_the_push_substeps_only(built):
    from oracle import Oracle
    n = 96
    env = _mk(n, "fp64", initial_force=55.0, action_force=100.0, no_carry_over_reset=True, reset_settle_cache=True, max_episode_len=6,
              steps_per_repeat=3, auto_reset="pipelined")
    ocfg = C.copy_config(env.cfg)
    ocfg.auto_reset = C.AUTO_RESET_OFF
    ora = Oracle(ocfg, dtype="fp64")
    assert np.array_equal(env.reset().cpu().numpy(), ora.reset())
    rng = np.random.default_rng(2)
    waited = np.zeros(n, dtype=int)
    n_fresh = 0
    for t in range(60):
        a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
        obs, rew, done, info = env.step(torch.from_numpy(a).cuda())
        reason = info["done_reason"].cpu().numpy()
        obs, done = obs.cpu().numpy(), done.cpu().numpy()
        fresh, resetting = reason == C.REASON_FRESH, reason == C.REASON_RESETTING
        live = ~(fresh | resetting)
        o = ora.step(a)
        assert np.array_equal(obs[live], o[0][live]) and np.array_equal(done[live], o[2][live])
        waited[resetting] += 1
        if fresh.any():
            o1 = ora.reset(fresh.astype(np.uint8))      # the oracle settles for 100 substeps; the kernel replayed them
            assert np.array_equal(obs[fresh], o1[fresh])
            assert (waited[fresh] == 4).all()            # 30 push substeps at 6 per call: FRESH on the 5th call (130 would take 22)
            waited[fresh] = 0
            n_fresh += int(fresh.sum())
    assert n_fresh > n
    env.close()


def test_fp32_split_pipeline_cache_on_and_off_agree(built):
    """fp32, the pipeline the bench times: the cached settle (computed by the one-launch kernel) against settling in place
    (split pipeline) - the same state up to fp32 rounding, and the cache removes 100 of the 130 reset substeps."""
    n = 8192
    kw = dict(initial_force=55.0, action_force=100.0, no_carry_over_reset=True, auto_reset="pipelined", max_episode_len=5)
    on, off = _mk(n, "fp32", reset_settle_cache=True, **kw), _mk(n, "fp32", **kw)
    assert on.solver_kind == off.solver_kind == "split/impulse-space"
    d = (on.reset() - off.reset()).abs().reshape(n, -1).amax(dim=1)
    assert float(d.median()) <= 1e-6 and float(d.max()) <= 1e-4, (float(d.median()), float(d.max()))
    g = torch.Generator(device="cuda").manual_seed(0)
    fresh_on = fresh_off = 0
    for t in range(40):
        a = torch.rand((n, 2), device="cuda", generator=g) * 2 - 1
        r_on, r_off = on.step(a), off.step(a)
        fresh_on += int((r_on[3]["done_reason"] == 5).sum())
        fresh_off += int((r_off[3]["done_reason"] == 5).sum())
        q = r_on[0][..., 3:7]
        assert bool(torch.isfinite(r_on[0]).all()) and float((q.norm(dim=-1) - 1).abs().max()) < 1e-4
    assert fresh_on >= 2 * n and fresh_off == 0   # 5 live + 15 resetting calls per cycle against 5 + 65
    on.close(); off.close()
