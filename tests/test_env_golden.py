"""CPU tier: the oracle's reset/step bookkeeping against traces produced by EXECUTING the reference's own
BulletCartpole source (tests/golden/gen_env_golden.py; bullet_cartpole.py:50-167,178-359) - call order, the one-substep
force lag, the force pending across reset(), 100 + 30 reset substeps, observation bootstrap, pose capture per repeat,
steps / done / reward / info, step after done, the push RNG's keying.  Bit for bit where the reference's own arithmetic is
exact (--no-random-theta), 1e-12 where it calls libm's cos / sin."""
import numpy as np
import pytest

import env_golden_util as EG
from cartpoleplusplus_b200 import config as C
from oracle import Oracle


class _World:
    def __init__(self, cfg):
        self.o = Oracle(cfg, dtype="fp64")
    def reset(self):
        return self.o.reset()
    def step(self, a):
        return self.o.step(a)


@pytest.mark.parametrize("case", EG.CASES)
def test_oracle_reproduces_the_reference_class_traces(case):
    g = EG.load(case)
    worlds = [_World(c) for c in EG.configs(g)]
    exact = case != "random_theta"
    EG.replay(g, worlds, exact)
    final = np.stack([w.o.state[0] for w in worlds])
    want = g["final_state"]
    keep = np.r_[0:C.S_STEPS, C.S_EPISODE]   # steps / done are the reference class's own fields, not the shim's
    if exact:
        assert np.array_equal(final[:, keep], want[:, keep])
    else:
        assert np.abs(final[:, keep] - want[:, keep]).max() <= 1e-10
    # the second cart-pole was driven by action[1]: its cart moved differently from the first
    assert not np.allclose(final[0, :3] + [1, 0, 0], final[1, :3], atol=1e-6)


@pytest.mark.parametrize("case", EG.CASES)
def test_reference_call_order_is_what_the_oracle_assumes(case):
    g = EG.load(case)
    got = bytes(g["call_order"]).decode()
    assert got == EG.expected_call_order(g)
    # the pushes of reset(): 30 per cart, magnitude --initial-force, direction theta (:354-359); actions: action * action_force
    f = g["forces"]
    R = 3 if case == "naf_recipe" else 2
    S = 4 if case == "naf_recipe" else 1
    per_reset, per_step = 60, 2 * R * S
    n_steps = int((g["kind"] == 1).sum()) // 2
    pushes = np.concatenate([f[k * (per_reset + n_steps * per_step):][:per_reset] for k in range(2)])
    mag = np.hypot(pushes[:, 1], pushes[:, 2])
    init = 200.0 if case == "random_theta" else 55.0
    assert np.allclose(mag, init, rtol=1e-12)
    if case != "random_theta":
        assert (pushes[:, 2] == 0).all() and (pushes[:, 1] == init).all()
    else:
        assert len(np.unique(np.round(pushes[:, 1], 9))) == 120  # a new theta on every push substep, per cart, per reset
