"""-m gpu: the pipeline bench.py times - k_prep / k_solve_lam / k_solve_tail (+ side-stream k_solve), pipelined
auto-reset - against the fp64 oracle ON THE GPU, at the bench's own configuration and batch size, from the
stationary mix of episode phases (live, resetting and fresh envs all present), through the C-ABI.

north_star: "the FP32 fast mode stays within 1e-4" of single-substep / single-call poses from identical states.
Rule enforced (tests/parity_util.py): every env within 1e-4 of the oracle unless a discrete decision provably
differs or the fp64 oracle itself is discontinuous at fp32 resolution at that state; anything else fails.
Reference lines under test: bullet_cartpole.py:197-203 (substeps + force), :237 (pose capture), :239-260
(bookkeeping), :313-346 (reset, run pipelined)."""
import json
import os

import numpy as np
import pytest

from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200 import config as C

pytestmark = pytest.mark.gpu

CASES = {
    # BASELINE.json configs[2]: what bench.py times (split / impulse-space solve phase)
    "configs2_65536": dict(n=65536, discrete=False, ov=dict(initial_force=200.0, action_force=100.0, auto_reset="pipelined")),
    # the same workload on the row-space solve phase (what configurations outside the impulse-space regimes run), two cohorts of 65 536
    "configs2_131072_rowspace": dict(n=131072, discrete=False, env={"BCP_LAMBDA": "0"},
                                     ov=dict(initial_force=200.0, action_force=100.0, auto_reset="pipelined")),
    # configs[1]-sized, discrete random actions with a push (the README's 4th protocol, README.md:90-92)
    "configs1_4096_discrete": dict(n=4096, discrete=True, ov=dict(initial_force=55.0, auto_reset="pipelined")),
    # upstream-style physics (pole stays on the cart, two-body friction rows live), one substep per call: north_star's
    # tolerance is stated for single-substep poses, and with the pole balanced on the cart a flipped decision grows
    # through the R*S = 10 substeps of that workload's call (measured there: median 3.5e-7, 2.9 % above 1e-4)
    "friction_single_substep_8192": dict(n=8192, discrete=False, mix=700,
                                         ov=dict(initial_force=55.0, action_force=100.0, cart_friction=1.0, action_repeats=1,
                                                 theta_per_reset=True, world_frame_force=True, auto_reset="pipelined")),
}


@pytest.mark.parametrize("name", list(CASES))
def test_timed_pipeline_against_fp64_oracle(built, name):
    import parity_util as PU
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    from oracle import Oracle
    case = CASES[name]
    n = case["n"]
    os.environ.update(case.get("env", {}))
    try:
        env = BatchedBulletCartpole(default_opts(**case["ov"]), discrete_actions=case["discrete"], num_envs=n, device=0, dtype="fp32", seed=17)
    finally:
        for k in case.get("env", {}):
            os.environ.pop(k, None)
    assert env.solver_kind == ("split/row-space" if case.get("env") or case["ov"].get("cart_friction") else "split/impulse-space"), env.solver_kind
    ora = Oracle(env.cfg, dtype="fp64")
    state = PU.stationary_mix(env, case.get("mix", 400))
    rng = np.random.default_rng(23)
    acts = rng.integers(0, 5, size=n).astype(np.int32) if case["discrete"] else rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
    rep = PU.classify_step(env, ora, state, acts, rng=rng)
    dev = rep.pop("dev")
    rep.pop("spread")
    print(name, json.dumps(rep))
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", f"pipeline_parity_{name}.json"), "w") as fh:
        json.dump(rep, fh, indent=1)
    ph = rep["phases"]
    assert ph["live"] > 0.3 * n and ph["resetting"] > 0.02 * n, ph   # the mix the bench times, not a fresh batch
    assert rep["median"] <= 1e-6 and rep["q99"] <= 1e-4, rep
    assert rep["above_tol_frac"] <= 5e-3, rep                           # decision flips are rare
    assert not rep["unexplained"], rep
    assert not rep["bookkeeping_unexplained"], rep
    env.close()
