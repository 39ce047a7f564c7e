"""CPU tier: the per-env programs of the CUDA kernels (csrc/*.cuh compiled for the host by
tests/hostharness.py) against the C oracle - no GPU needed.
  parity path (substep<>)        : bit-exact in fp64
  fast path   (substep_fast<>)   : same algorithm re-laid; fp64 instantiation <= 1e-10 per substep
  split pipeline pieces          : bitwise equal to the fast path
  pipelined auto-reset machine   : equals the oracle driven with synchronous resets"""
import numpy as np
import pytest

from cartpoleplusplus_b200 import default_opts, make_config
from cartpoleplusplus_b200 import config as C
from hostharness import HostKernels
from oracle import Oracle

CONFIGS = [dict(initial_force=55.0, action_force=100.0),
           dict(initial_force=55.0, action_force=100.0, cart_friction=1.0, steps_per_repeat=3),
           dict(initial_force=200.0, action_force=100.0, theta_per_reset=True, world_frame_force=True)]


def _cfg(n, dtype="fp64", **kw):
    return make_config(default_opts(**kw), False, num_envs=n, dtype=dtype, seed=3)


@pytest.mark.parametrize("kw", CONFIGS)
def test_parity_path_bit_exact_and_fast_path_to_rounding(kw):
    n = 96
    cfg = _cfg(n, **kw)
    ora, par = Oracle(cfg, dtype="fp64"), HostKernels(cfg, "fp64", fast=0)
    assert np.array_equal(ora.reset(), par.reset()) and np.array_equal(ora.state, par.state)
    rng = np.random.default_rng(1)
    for t in range(25):
        a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
        ro, rp = ora.step(a), par.step(a)
        assert all(np.array_equal(x, y) for x, y in zip(ro, rp)), t
    assert np.array_equal(ora.state, par.state)
    # single substeps from identical states: fast path, fp64 instantiation and fp32
    st = ora.state.copy()
    st[:, C.S_DONE] = 0
    fst = HostKernels(cfg, "fp64", fast=1)
    fst.state[:] = st
    ora.state[:] = st
    i_o, i_f = ora.substep(with_info=True), fst.substep(with_info=True)
    assert np.array_equal(i_o, i_f)  # same contacts, same rows, same iteration counts
    assert np.abs(ora.state - fst.state)[:, :C.S_MANI].max() <= 1e-10
    st32 = st.astype(np.float32).astype(np.float64)
    f32 = HostKernels(cfg, "fp32", fast=1)
    f32.state[:] = st32
    ora.state[:] = st32
    ora.substep(); f32.substep()
    d = np.abs(ora.state - f32.state)
    pose = np.concatenate([d[:, 0:7], d[:, 13:20]], axis=1).max(axis=1)
    assert np.median(pose) <= 1e-6 and np.quantile(pose, 0.99) <= 1e-4, (np.median(pose), pose.max())


@pytest.mark.parametrize("kw", CONFIGS[:2])
def test_split_pipeline_pieces_against_oracle(kw):
    """split_setup -> solve_iteration (16-byte rows in velocity units, zero rows in unused
    slots) -> split_finish: fp64 instantiation vs the oracle per substep (same contacts, same
    iteration counts, state to rounding), and whole env.steps in fp32 stay on the fast path's."""
    n = 96
    cfg = _cfg(n, **kw)
    ora = Oracle(cfg, dtype="fp64")
    ora.reset()
    rng = np.random.default_rng(2)
    for t in range(20):
        ora.step(rng.uniform(-1, 1, size=(n, 2)).astype(np.float32))
    st = ora.state.copy()
    st[:, C.S_DONE] = 0
    spl = HostKernels(cfg, "fp64", fast=2)
    spl.state[:] = st
    ora.state[:] = st
    i_o, i_s = ora.substep(with_info=True), spl.substep(with_info=True)
    assert np.array_equal(i_o, i_s)
    assert np.abs(ora.state - spl.state)[:, :C.S_MANI_CNT].max() <= 1e-10
    cfg32 = _cfg(n, dtype="fp32", **kw)
    a, b = HostKernels(cfg32, "fp32", fast=1), HostKernels(cfg32, "fp32", fast=2)
    st32 = st.astype(np.float32).astype(np.float64)
    a.state[:] = st32
    b.state[:] = st32
    act = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
    ra, rb = a.step(act), b.step(act)
    d = np.abs(ra[0].astype(np.float64) - rb[0]).reshape(n, -1).max(axis=1)
    assert np.median(d) <= 1e-6 and np.quantile(d, 0.99) <= 1e-4, (np.median(d), d.max())
    assert np.array_equal(ra[1], rb[1]) and (ra[2] != rb[2]).mean() <= 0.02


def test_pipelined_machine_equals_oracle_with_synchronous_resets():
    n = 48
    kw = dict(initial_force=55.0, action_force=100.0, max_episode_len=5, steps_per_repeat=3)
    cfg_p = make_config(default_opts(auto_reset="pipelined", **kw), False, num_envs=n, dtype="fp64", seed=5)
    cfg_o = make_config(default_opts(**kw), False, num_envs=n, dtype="fp64", seed=5)
    env, ora = HostKernels(cfg_p, "fp64", fast=0), Oracle(cfg_o, dtype="fp64")
    assert np.array_equal(env.reset(), ora.reset())
    rng = np.random.default_rng(9)
    waited = np.zeros(n, dtype=int)
    n_fresh = 0
    for t in range(70):
        a = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
        obs, rew, done, reason = env.step(a)
        fresh, resetting = reason == C.REASON_FRESH, reason == C.REASON_RESETTING
        live = ~(fresh | resetting)
        obs_o, rew_o, done_o, reason_o = ora.step(a)
        assert np.array_equal(obs[live], obs_o[live]) and np.array_equal(done[live], done_o[live])
        assert np.array_equal(reason[live], reason_o[live]) and (rew[live] == 1).all() and (rew[~live] == 0).all()
        waited[resetting] += 1
        if fresh.any():
            o1 = ora.reset(fresh.astype(np.uint8))
            assert np.array_equal(obs[fresh], o1[fresh])
            assert (waited[fresh] == 21).all()  # 130 substeps at 6 per call: FRESH on the 22nd call
            waited[fresh] = 0
            n_fresh += int(fresh.sum())
    assert n_fresh >= n


def test_generic_regime_in_impulse_space_against_oracle():
    """States in which the pole touches the cart AND the ground (the regime neither STACK nor SPLIT covers, LamGen in
    bcp_lambda.cuh, k_solve_gen_once on the GPU), collected from an oracle rollout of the bench workload: one substep of the
    split pipeline's host composition in fp64 vs the oracle - same contacts, same iteration counts, state within 1e-10 -
    and the fp32 instantiation against the row-space sweep it replaces (BCP_NO_LAMGEN)."""
    import os
    n = 1024
    cfg = _cfg(n, initial_force=200.0, action_force=100.0)
    ora = Oracle(cfg, dtype="fp64")
    ora.reset()
    rng = np.random.default_rng(7)
    found = []
    for t in range(120):
        ora.step(rng.uniform(-1, 1, size=(n, 2)).astype(np.float32))
        probe = Oracle(cfg, dtype="fp64")
        probe.state[:] = ora.state
        probe.state[:, C.S_DONE] = 0
        before = probe.state.copy()
        info = probe.substep(with_info=True)
        gen = (info[:, 3 + 1] > 0) & (info[:, 3 + 2] > 0)  # contacts of cart-pole and ground-pole
        found.append(before[gen])
        if sum(len(f) for f in found) >= 64:
            break
    st = np.concatenate(found)
    assert len(st) >= 16, "the rollout met too few states of the generic regime"
    m = len(st)
    cfg_m = _cfg(m, initial_force=200.0, action_force=100.0)
    ora_m, spl = Oracle(cfg_m, dtype="fp64"), HostKernels(cfg_m, "fp64", fast=2)
    ora_m.state[:] = st
    spl.state[:] = st
    i_o, i_s = ora_m.substep(with_info=True), spl.substep(with_info=True)
    assert (i_o[:, 4] > 0).all() and (i_o[:, 5] > 0).all()
    assert np.array_equal(i_o, i_s)
    assert np.abs(ora_m.state - spl.state)[:, :C.S_MANI_CNT].max() <= 1e-10
    assert (i_o[:, 0] >= 50).any(), "no solve of the sample ran all the iterations"
    # fp32: impulse space vs the row-space sweep from the same states
    cfg32 = _cfg(m, dtype="fp32", initial_force=200.0, action_force=100.0)
    st32 = st.astype(np.float32).astype(np.float64)
    a = HostKernels(cfg32, "fp32", fast=2)
    a.state[:] = st32
    ia = a.substep(with_info=True)
    os.environ["BCP_NO_LAMGEN"] = "1"
    try:
        b = HostKernels(cfg32, "fp32", fast=2)
        b.state[:] = st32
        ib = b.substep(with_info=True)
    finally:
        del os.environ["BCP_NO_LAMGEN"]
    d = np.abs(a.state - b.state)
    pose = np.concatenate([d[:, 0:7], d[:, 13:20]], axis=1).max(axis=1)  # positions and quaternions (north_star's 1e-4)
    same_it = ia[:, 0] == ib[:, 0]
    assert np.median(pose) <= 1e-6 and (pose[same_it] <= 1e-4).all(), (np.median(pose), pose.max())
    assert same_it.mean() >= 0.8
