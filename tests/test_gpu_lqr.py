"""GPU tier: bcp_lqr_control (the fork's 8-state observation, done rule and LQR feedback, one launch per
batch) against oracle/lqr_oracle.py on the env's own canonical state records."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _env(n, dtype, **kw):
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    from cartpoleplusplus_b200.lqr import lqr_opts
    return BatchedBulletCartpole(lqr_opts(**kw), discrete_actions=False, num_envs=n, device=0, dtype=dtype, seed=3)


@pytest.mark.parametrize("n,dtype", [(257, "fp64"), (4096, "fp32")])
def test_state_done_and_forces_match_the_oracle(built, n, dtype):
    from cartpoleplusplus_b200.lqr import BatchedLqrCartpole
    from oracle import lqr_oracle as L
    env = _env(n, dtype)
    rng = np.random.default_rng(0)
    x0 = rng.normal(scale=0.001, size=n).astype(np.float32)
    ips = np.zeros((n, 8), np.float32)
    ips[:, 0] = x0
    zero = rng.normal(scale=0.05, size=8).astype(np.float32)
    PT, AT = 0.004, 0.006  # tight thresholds: both rules fire within a second of simulated time
    lq = BatchedLqrCartpole(env, PT, AT, initial_pole_state=ips, lqr_zero_point=zero)
    lq.reset()
    K = torch.from_numpy(rng.normal(scale=3.0, size=(n, 2, 8)).astype(np.float32)).cuda()
    seen = set()
    bias = rng.uniform(-30, 30, size=(n, 2))
    for t in range(260):
        # a steady push per env plus noise: poles topple (angle rule) and, later, carts leave the 3 m box (position rule)
        dist = torch.from_numpy((bias + rng.uniform(-5, 5, size=(n, 2))).astype(np.float32)).cuda()
        u = lq.lqr_control_forces(K if t % 2 else K[0], dist).clone()
        rec = env.get_state()
        st = L.lqr_state(rec, x0)
        got = lq.pole_state.cpu().numpy()
        assert np.allclose(got, st, rtol=0, atol=2e-6), (t, np.abs(got - st).max())
        # thresholds and the control law are exact functions of the float32 state the kernel produced
        assert np.array_equal(lq.done.cpu().numpy(), L.check_done(got, PT, AT))
        want_u = L.actions(got, (K if t % 2 else K[0]).cpu().numpy(), zero, dist.cpu().numpy())
        assert np.array_equal(u.cpu().numpy(), want_u), t
        seen |= set(np.unique(lq.done.cpu().numpy()).tolist())
        state, reward, done, info = lq.step(dist)  # open loop: the disturbance alone
        assert state.shape == (n, 8) and float(reward.abs().sum()) == 0.0
        assert torch.equal(done, info["done_reason"] != 0)
    assert seen == {0, 1, 2}
    env.close()


def test_shared_inputs_and_null_outputs(built):
    import ctypes
    from cartpoleplusplus_b200 import _lib
    from cartpoleplusplus_b200.lqr import BatchedLqrCartpole
    env = _env(64, "fp32")
    lq = BatchedLqrCartpole(env)
    lq.reset()
    s1 = lq.observe_state().clone()
    u = lq.lqr_control_forces(np.zeros((2, 8), np.float32), disturbance=np.array([1.5, -2.0], np.float32))
    assert torch.equal(u, torch.tensor([1.5, -2.0], device="cuda").expand(64, 2))
    assert torch.equal(lq.pole_state, s1)
    with pytest.raises(ValueError):
        lq.lqr_control_forces(np.zeros((3, 8), np.float32))
    lib = _lib.load()
    args = _lib.BcpLqrArgs()  # no outputs at all
    assert lib.bcp_lqr_control(env._h, ctypes.byref(args), None) != 0
    assert b"no output" in lib.bcp_last_error()
    env.close()


def test_population_rollout_bookkeeping(built):
    """rollout(): a gain matrix per env, rolled out side by side; survival counts and first done reasons equal a
    step-by-step replay of the fork's loop body (control from the state, then step) on a twin env."""
    from cartpoleplusplus_b200.lqr import BatchedLqrCartpole
    from oracle import lqr_oracle as L
    n, T = 1024, 150
    envs = [_env(n, "fp32", cart_friction=1.0, action_force=1.0) for _ in range(2)]
    a, b = (BatchedLqrCartpole(e, position_threshold=0.01, angle_threshold=0.5) for e in envs)
    a.reset()
    b.reset()
    rng = np.random.default_rng(5)
    K = rng.normal(scale=20.0, size=(n, 2, 8)).astype(np.float32)
    push = torch.from_numpy(rng.uniform(-150, 150, size=(n, 2)).astype(np.float32)).cuda()
    fn = lambda t: push if t < 40 else None  # noqa: E731
    out = a.rollout(K, T, disturbance_fn=fn, stop_when_all_done=False)
    alive = np.ones(n, bool)
    survived = np.zeros(n, np.int32)
    reason = np.zeros(n, np.uint8)
    for t in range(T):
        st = b.observe_state().cpu().numpy()
        u = L.actions(st, K, None, None if t >= 40 else push.cpu().numpy())
        state, _, done, info = b.step(torch.from_numpy(u).cuda())
        d = L.check_done(state.cpu().numpy(), 0.01, 0.5)
        assert np.array_equal(d, info["done_reason"].cpu().numpy())
        reason = np.where(alive & (d != 0), d, reason)
        alive &= d == 0
        survived += alive
    assert np.array_equal(out["steps_survived"].cpu().numpy(), survived)
    assert np.array_equal(out["first_done_reason"].cpu().numpy(), reason)
    assert torch.equal(out["final_state"], b.pole_state)
    assert survived.max() > survived.min() and (reason != 0).any()  # the scenario did end episodes, at different times
    for e in envs:
        e.close()
