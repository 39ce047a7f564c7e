"""-m gpu: the CUDA path, called through the C-ABI, against the CPU oracle.

FP64 parity mode must be BIT-EXACT (-fmad=false kernel vs -ffp-contract=off oracle, same
IEEE operation sequence); FP32 fast mode is held to 1e-4 absolute on single substeps from
identical states (BASELINE.json north_star tolerance)."""
import numpy as np
import pytest

from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200 import config as C

pytestmark = pytest.mark.gpu


def _mk(dtype, n, discrete, **ov):
    import torch
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    from oracle import Oracle
    assert torch.cuda.is_available()
    opts = default_opts(**ov)
    env = BatchedBulletCartpole(opts, discrete_actions=discrete, num_envs=n, device=0, dtype=dtype, seed=11)
    return env, Oracle(env.cfg, dtype=dtype)


@pytest.mark.parametrize("ov", [dict(), dict(cart_friction=1.0, initial_force=55.0, steps_per_repeat=3),
                                dict(initial_force=55.0, no_termination=True)])
def test_fp64_rollout_bit_exact(built, ov):
    import torch
    n = 512
    env, ora = _mk("fp64", n, False, **ov)
    obs = env.reset().cpu().numpy()
    assert np.array_equal(obs, ora.reset())
    assert np.array_equal(env.get_state(), ora.state)
    rng = np.random.default_rng(5)
    for t in range(40):
        a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
        obs, rew, done, info = env.step(torch.from_numpy(a).cuda())
        obs_o, rew_o, done_o, reason_o = ora.step(a)
        assert np.array_equal(obs.cpu().numpy(), obs_o), t
        assert np.array_equal(rew.cpu().numpy(), rew_o)
        assert np.array_equal(done.cpu().numpy(), done_o)
        assert np.array_equal(info["done_reason"].cpu().numpy(), reason_o)
        if t % 10 == 9:  # masked reset of finished envs
            m = done_o.astype(np.uint8)
            if m.any():
                o1 = env.reset(torch.from_numpy(m).cuda()).cpu().numpy()
                assert np.array_equal(o1, ora.reset(m))
    assert np.array_equal(env.get_state(), ora.state)
    env.close()


def test_fp64_discrete_bit_exact(built):
    import torch
    n = 256
    env, ora = _mk("fp64", n, True, initial_force=55.0)
    env.reset(); ora.reset()
    rng = np.random.default_rng(6)
    for t in range(30):
        a = rng.integers(0, 5, size=n).astype(np.int32)
        obs, rew, done, info = env.step(torch.from_numpy(a).cuda())
        obs_o, rew_o, done_o, reason_o = ora.step(a)
        assert np.array_equal(obs.cpu().numpy(), obs_o), t
        assert np.array_equal(done.cpu().numpy(), done_o)
    assert np.array_equal(env.get_state(), ora.state)
    env.close()


def test_fp32_single_substep_within_1e4(built):
    """north_star: fp32 fast mode within 1e-4 of the fp64 poses after a single substep from identical states - every
    env, except those where a discrete decision provably differs (tests/parity_util.py states the rule; none may be
    unexplained).  action_repeats = steps_per_repeat = 1, so one env.step is one p.stepSimulation() (:199) plus the force
    application (:201-203), which moves no pose.  1024 envs: the one-launch fast kernel."""
    import parity_util as PU
    n = 1024
    env, ora32 = _mk("fp32", n, False, initial_force=55.0, action_repeats=1)
    assert env.solver_kind == "one-launch"
    from oracle import Oracle
    ora64 = Oracle(env.cfg, dtype="fp64")
    ora64.reset()
    rng = np.random.default_rng(7)
    for _ in range(20):
        ora64.step(rng.uniform(-1, 1, size=(n, 2)).astype(np.float32))
    st = ora64.state.copy()
    st[:, C.S_DONE] = 0
    st32 = st.astype(np.float32).astype(np.float64)  # identical start representable in fp32
    rep = PU.classify_step(env, ora64, st32, rng.uniform(-1, 1, size=(n, 2)).astype(np.float32), rng=rng)
    rep.pop("dev"), rep.pop("spread")
    print(rep)
    assert rep["median"] <= 1e-6 and rep["q99"] <= 1e-4, rep
    assert not rep["unexplained"] and not rep["bookkeeping_unexplained"], rep
    assert rep["above_tol_frac"] <= 5e-3, rep
    env.close()


def test_step_after_done_and_errors(built):
    import torch
    env, ora = _mk("fp64", 8, True, initial_force=0.0, max_episode_len=3)
    env.reset()
    for t in range(3):
        obs, rew, done, info = env.step(torch.zeros(8, dtype=torch.int64).cuda())
    assert done.all() and (info["done_reason"] == 3).all() and (rew == 1).all()
    obs2, rew2, done2, info2 = env.step(torch.zeros(8, dtype=torch.int64).cuda())
    assert torch.equal(obs2, obs) and (rew2 == 0).all() and done2.all()
    env.close()


def test_compound_cart_fp64_bit_exact_and_fp32_classified(built):
    """the fork's compound cart (models/double_cart_1.urdf:15-80; --cart-model double_cart): fp64 kernel == oracle bit for
    bit over a rollout with resets; one env.step of the fp32 split pipeline (row space: the cart-pole rows have friction)
    within 1e-4 of the oracle or a classified decision flip."""
    import torch
    import parity_util as PU
    n = 256
    env, ora = _mk("fp64", n, False, initial_force=55.0, action_force=100.0, cart_model="double_cart", steps_per_repeat=2)
    assert np.array_equal(env.reset().cpu().numpy(), ora.reset())
    rng = np.random.default_rng(3)
    for t in range(20):
        a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
        obs, rew, done, info = env.step(torch.from_numpy(a).cuda())
        o = ora.step(a)
        assert np.array_equal(obs.cpu().numpy(), o[0]) and np.array_equal(done.cpu().numpy(), o[2]), t
    assert np.array_equal(env.get_state(), ora.state)
    env.close()
    n = 8192
    env32, _ = _mk("fp32", n, False, initial_force=55.0, action_force=100.0, cart_model="double_cart", action_repeats=1, auto_reset="pipelined")
    assert env32.solver_kind == "split/row-space"
    from oracle import Oracle
    ora64 = Oracle(env32.cfg, dtype="fp64")
    state = PU.stationary_mix(env32, 600)
    rep = PU.classify_step(env32, ora64, state, rng.uniform(-1, 1, size=(n, 2)).astype(np.float32), rng=rng)
    rep.pop("dev"), rep.pop("spread")
    print(rep)
    assert rep["median"] <= 1e-6 and rep["q99"] <= 1e-4 and not rep["unexplained"] and not rep["bookkeeping_unexplained"], rep
    env32.close()
