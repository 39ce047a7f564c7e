"""CPU tier: the oracle's own statement of the pipelined auto-reset machine (bcp.h BCP_AUTO_RESET_PIPELINED, no reference
counterpart) against the reference-shaped synchronous form - reset() of bullet_cartpole.py:313-346 called by the host
when an env is done - and against the kernels' per-env program compiled for the host."""
import numpy as np

from cartpoleplusplus_b200 import default_opts, make_config
from cartpoleplusplus_b200 import config as C
from hostharness import HostKernels
from oracle import Oracle


def _pair(n, **kw):
    cfg_p = make_config(default_opts(auto_reset="pipelined", **kw), False, num_envs=n, dtype="fp64", seed=5)
    cfg_o = make_config(default_opts(**kw), False, num_envs=n, dtype="fp64", seed=5)
    return cfg_p, cfg_o


def test_pipelined_oracle_equals_synchronous_oracle():
    n = 40
    for kw in (dict(initial_force=55.0, action_force=100.0, max_episode_len=5, steps_per_repeat=3),
               dict(initial_force=200.0, action_force=100.0, max_episode_len=9, theta_per_reset=True, world_frame_force=True,
                    cart_friction=1.0)):
        cfg_p, cfg_o = _pair(n, **kw)
        pip, syn = Oracle(cfg_p, dtype="fp64"), Oracle(cfg_o, dtype="fp64")
        assert np.array_equal(pip.reset(), syn.reset())
        rs = cfg_p.action_repeats * cfg_p.steps_per_repeat
        calls = -(-130 // rs)  # FRESH on the call in which the 130th reset substep runs
        rng = np.random.default_rng(9)
        waited = np.zeros(n, dtype=int)
        n_fresh = 0
        for t in range(3 * calls + 40):
            a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
            obs, rew, done, reason, info = pip.step(a, with_info=True)
            fresh, resetting = reason == C.REASON_FRESH, reason == C.REASON_RESETTING
            live = ~(fresh | resetting)
            obs_o, rew_o, done_o, reason_o = syn.step(a)
            assert np.array_equal(obs[live], obs_o[live]) and np.array_equal(done[live], done_o[live])
            assert np.array_equal(reason[live], reason_o[live]) and (rew[live] == 1).all() and (rew[~live] == 0).all()
            assert done[resetting].all() and not done[fresh].any()
            assert (info[live][:, :, 0] > 0).all() and (info[resetting][:, :, 0] > 0).all()
            waited[resetting] += 1
            if fresh.any():
                o1 = syn.reset(fresh.astype(np.uint8))
                assert np.array_equal(obs[fresh], o1[fresh])
                assert (waited[fresh] == calls - 1).all()
                ran = (info[fresh][:, :, 0] >= 0).sum(axis=1)
                assert (ran == 130 - (calls - 1) * rs).all()  # the substeps after the reset's last one do not run
                waited[fresh] = 0
                n_fresh += int(fresh.sum())
            ok = pip.state[:, C.S_RESET_LEFT] == 0
            assert np.array_equal(pip.state[ok][:, :C.S_RESET_LEFT], syn.state[ok][:, :C.S_RESET_LEFT])
        assert n_fresh >= n


def test_pipelined_oracle_equals_the_kernels_host_program_bit_for_bit():
    n = 48
    cfg_p, _ = _pair(n, initial_force=55.0, action_force=100.0, max_episode_len=5, steps_per_repeat=3)
    env, ora = HostKernels(cfg_p, "fp64", fast=0), Oracle(cfg_p, dtype="fp64")
    assert np.array_equal(env.reset(), ora.reset())
    rng = np.random.default_rng(3)
    seen = set()
    for t in range(60):
        a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
        re, ro = env.step(a), ora.step(a)
        live = (ro[3] != C.REASON_RESETTING)  # a resetting env's observation block is left untouched by both
        assert np.array_equal(re[0][live], ro[0][live])
        assert all(np.array_equal(x, y) for x, y in zip(re[1:], ro[1:])), t
        assert np.array_equal(env.state, ora.state), t
        seen |= set(ro[3].tolist())
    assert {0, C.REASON_RESETTING, C.REASON_FRESH} <= seen
