"""Replays tests/golden/env_golden.npz - traces RETURNED BY THE REFERENCE'S OWN BulletCartpole class, whose source
tests/golden/gen_env_golden.py executed over a pybullet shim - through anything with the env surface
(reset() -> obs, step(a) -> obs, reward, done, reason).  Used by the CPU tier (oracle) and the GPU tier (fp64 kernel)."""
import argparse
import os

import numpy as np

from cartpoleplusplus_b200 import add_opts, make_config

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "env_golden.npz")
CASES = ("no_theta", "random_theta", "naf_recipe")


def load(case):
    z = np.load(GOLDEN)
    g = {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(case + "/")}
    g["seed"] = int(z["seed"])
    return g


def configs(g, dtype="fp64"):
    """the two cart-poles of the fork as two single-env configs: the reference's own CLI flags parsed by this package's
    add_opts (same names), fork semantics (no termination rule; theta per push substep and LINK frame are the defaults)"""
    parser = argparse.ArgumentParser()
    add_opts(parser)
    opts = parser.parse_args([str(a) for a in g["argv"]] + ["--no-termination"])
    return [make_config(opts, False, num_envs=1, dtype=dtype, seed=g["seed"], env_id_offset=i, init_offset=(float(i), 0.0)) for i in range(2)]


def expected_call_order(g):
    """what bullet_cartpole.py:313-346 and :197-207 must have asked of pybullet, as a string over
    t = resetBasePositionAndOrientation, s = stepSimulation, f = applyExternalForce"""
    R, = [int(a.split("=")[1]) for a in g["argv"] if str(a).startswith("--action-repeats")] or [2]
    S, = [int(a.split("=")[1]) for a in g["argv"] if str(a).startswith("--steps-per-repeat")] or [1]
    out = ""
    for kind in g["kind"]:
        if kind == 0:
            out += "tttt" + "s" * 100 + "sff" * 30
        elif kind == 1:
            out += "sff" * (R * S)
        # kind 2, step after done: no pybullet call at all (:179-181)
    return out


def replay(g, worlds, exact):
    """worlds: two env-like objects (cart-pole at x = 0 observed by the reference, and the one at x = 1)"""
    tol = 0.0 if exact else 1e-12
    def same(a, b):
        return np.array_equal(a, b) if exact else float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max()) <= tol
    for i, kind in enumerate(g["kind"]):
        if kind == 0:
            obs = [w.reset() for w in worlds]
            assert same(obs[0][0], g["obs"][i]), ("reset", i)
            done = False
        else:
            res = [w.step(g["actions"][i][k][None]) for k, w in enumerate(worlds)]
            obs, rew, dn, reason = res[0][0][0], float(res[0][1][0]), bool(res[0][2][0]), int(res[0][3][0])
            assert same(obs, g["obs"][i]), ("step", i)
            assert rew == g["reward"][i] and dn == bool(g["done"][i]), ("step", i, rew, dn)
            if kind == 1:
                assert (reason == 3) == bool(g["info"][i]), ("done_reason", i)
            else:  # step after done: old observation, reward 0, done (:179-181)
                assert rew == 0.0 and dn
