"""Regenerates tests/golden/env_golden.npz by EXECUTING THE REFERENCE'S OWN ENV SOURCE
(run in the build container only; /root/reference does not exist on the GPU box).

bullet_cartpole.py cannot be imported as shipped: it imports gym and pybullet, both absent.  This script takes the
SOURCE TEXT of /root/reference/bullet_cartpole.py at run time (nothing is copied into the repo) and executes it as a
module against
  * a stub `gym` (Env base class; spaces from cartpoleplusplus_b200.spaces), and
  * a `pybullet` shim `p` whose world is two cart-poles - the fork loads cart/pole at x = 0 and cart2/pole2 at x = 1
    (:155-160) - each backed by one single-env instance of the CPU oracle: p.stepSimulation() = one oracle substep of
    both, p.applyExternalForce = the oracle's force accumulator, p.resetBasePositionAndOrientation = pose overwrite with
    the velocities zeroed (caches and pending force survive, [BULLET-RECALL]), p.getBasePositionAndOrientation = the pose,
and with `np.random.random` served from the Philox stream the kernels use (key: seed, cart-pole index, reset count, push
index; the reference draws cart then cart2 on every push substep, :329-332).

What this pins: everything bullet_cartpole.py itself decides - call order of stepSimulation / applyExternalForce, the
one-substep force lag, the force that is still pending when reset() teleports the bodies, the 100 + 30 substeps of reset,
the observation bootstrap, pose capture per repeat, steps / done / reward / info, step-after-done.  What it does NOT pin
is the shim's physics (the oracle's substep): that remains a restatement of Bullet.

The stored traces are what the REFERENCE CLASS returned; tests replay the actions through oracle.Oracle.reset/step (CPU
tier) and through the fp64 CUDA kernel (GPU tier) and must reproduce them - bit for bit with --no-random-theta (theta = 0:
cos and sin are exact), to 1e-12 with random theta (np.cos / np.sin of libm vs the oracle's fixed-sequence sin / cos).
"""
import argparse
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
REF = "/root/reference/bullet_cartpole.py"
HERE = os.path.dirname(os.path.abspath(__file__))

from cartpoleplusplus_b200 import config as C  # noqa: E402
from cartpoleplusplus_b200 import default_opts, make_config, spaces  # noqa: E402
from oracle import Oracle  # noqa: E402
from oracle import oracle as O  # noqa: E402

SEED = 1234


class BulletShim:
    """the ~10 pybullet entry points bullet_cartpole.py calls, over two single-env oracles"""
    GUI, DIRECT, LINK_FRAME, WORLD_FRAME = 1, 2, 1, 2

    def __init__(self, ref_opts):
        # fork semantics for the oracle: no termination rule, theta per push substep, LINK frame, the shipped URDFs
        kw = dict(action_force=ref_opts.action_force, initial_force=ref_opts.initial_force, no_random_theta=ref_opts.no_random_theta,
                  action_repeats=ref_opts.action_repeats, steps_per_repeat=ref_opts.steps_per_repeat,
                  max_episode_len=ref_opts.max_episode_len, no_termination=True)
        self.cfgs = [make_config(default_opts(**kw), False, num_envs=1, dtype="fp64", seed=SEED, env_id_offset=i, init_offset=(float(i), 0.0))
                     for i in range(2)]
        self.worlds = [Oracle(c, dtype="fp64") for c in self.cfgs]
        self.bodies = {}      # body id -> (world, 'cart' | 'pole' | 'ground')
        self.log = []         # (what, body, fx, fy) in call order
        self.n_loaded = 0
        self.resets = 0       # finished resets = the RNG's episode word
        self.draws = 0        # np.random.random calls since the last new_episode()

    # -- world construction, :151-166
    def connect(self, mode):
        assert mode == self.DIRECT
    def setGravity(self, x, y, z):
        assert (x, y, z) == (0, 0, -9.81)
    def loadURDF(self, path, x, y, z, qx, qy, qz, qw):
        name = os.path.basename(path)
        bid = self.n_loaded
        self.n_loaded += 1
        kind = {"ground.urdf": "ground", "cart.urdf": "cart", "pole.urdf": "pole", "cart2.urdf": "cart", "pole2.urdf": "pole"}[name]
        world = 1 if name in ("cart2.urdf", "pole2.urdf") else 0
        self.bodies[bid] = (world, kind)
        if kind != "ground":
            self._set_pose(world, kind, (x, y, z), (qx, qy, qz, qw))
        return bid
    def resetDebugVisualizerCamera(self, **kw):
        pass

    def _base(self, kind):
        return C.S_CART if kind == "cart" else C.S_POLE
    def _set_pose(self, world, kind, pos, orn):
        s = self.worlds[world].state[0]
        b = self._base(kind)
        s[b:b + 3] = pos
        s[b + 3:b + 7] = orn
        s[b + 7:b + 13] = 0.0

    # -- the hot path
    def stepSimulation(self):
        self.log.append(("step", -1, 0.0, 0.0))
        for w in self.worlds:
            w.substep()
    def applyExternalForce(self, body, link, force, pos, frame):
        world, kind = self.bodies[body]
        assert kind == "cart" and link == -1 and tuple(pos) == (0, 0, 0) and frame == self.LINK_FRAME and force[2] == 0
        self.log.append(("force", body, float(force[0]), float(force[1])))
        self.worlds[world].apply_force([float(force[0]), float(force[1])])
    def resetBasePositionAndOrientation(self, body, pos, orn):
        world, kind = self.bodies[body]
        self.log.append(("teleport", body, 0.0, 0.0))
        self._set_pose(world, kind, pos, orn)
    def getBasePositionAndOrientation(self, body):
        world, kind = self.bodies[body]
        s = self.worlds[world].state[0]
        b = self._base(kind)
        return tuple(float(v) for v in s[b:b + 3]), tuple(float(v) for v in s[b + 3:b + 7])
    def getBaseVelocity(self, body):
        world, kind = self.bodies[body]
        s = self.worlds[world].state[0]
        b = self._base(kind)
        return tuple(float(v) for v in s[b + 7:b + 10]), tuple(float(v) for v in s[b + 10:b + 13])
    def getEulerFromQuaternion(self, q):  # feeds the fork's debug side channel (monkey_*) only
        x, y, z, w = q
        return (float(np.arctan2(2 * (y * z + w * x), w * w - x * x - y * y + z * z)), float(np.arcsin(np.clip(-2 * (x * z - w * y), -1, 1))),
                float(np.arctan2(2 * (x * y + w * z), w * w + x * x - y * y - z * z)))

    # -- np.random.random of random_force_in_plane (:355): cart, cart2, cart, cart2, ...
    def new_episode(self):
        self.draws = 0
    def random(self):
        world, k = self.draws % 2, self.draws // 2
        self.draws += 1
        r = O.philox(SEED, world, self.resets, 1, k)  # purpose 1 = push
        return float(r[0] >> 8) * 5.9604644775390625e-08
    def end_reset(self):
        self.resets += 1
        for w in self.worlds:  # the oracle's own counters, so that the final states are comparable
            w.state[0, C.S_EPISODE] = self.resets


class NumpyProxy:
    """numpy with np.random.random replaced"""
    def __init__(self, shim):
        self.random = types.SimpleNamespace(random=shim.random)
    def __getattr__(self, name):
        return getattr(np, name)


def load_reference(shim):
    gym = types.ModuleType("gym")
    gym.Env = type("Env", (object,), {})
    gym.spaces = spaces
    pyb = types.ModuleType("pybullet")
    for name in dir(shim):
        if not name.startswith("_"):
            setattr(pyb, name, getattr(shim, name))
    saved = {k: sys.modules.get(k) for k in ("gym", "gym.spaces", "pybullet")}
    sys.modules.update({"gym": gym, "gym.spaces": spaces, "pybullet": pyb})
    try:
        mod = types.ModuleType("bullet_cartpole_reference")
        exec(compile(open(REF).read(), REF, "exec"), mod.__dict__)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    mod.np = NumpyProxy(shim)
    return mod


def run(argv, n_steps, seed):
    # the reference's own argparse surface
    boot = load_reference(BulletShim(default_opts()))
    parser = argparse.ArgumentParser()
    boot.add_opts(parser)
    ref_opts = parser.parse_args(argv)
    shim = BulletShim(ref_opts)
    mod = load_reference(shim)
    cwd = os.getcwd()
    env = mod.BulletCartpole(ref_opts, discrete_actions=False)
    rng = np.random.default_rng(seed)
    out = {"obs": [], "reward": [], "done": [], "info": [], "actions": [], "kind": []}

    def record(kind, obs, reward, done, info, action):
        out["kind"].append(kind)
        out["obs"].append(np.asarray(obs, dtype=np.float32))
        out["reward"].append(float(reward))
        out["done"].append(bool(done))
        out["info"].append(1 if info.get("done_reason") == "episode length" else 0)
        out["actions"].append(np.zeros((2, 2), np.float32) if action is None else action)

    for episode in range(2):
        shim.new_episode()
        obs = env.reset()
        shim.end_reset()
        record(0, obs, 0.0, False, {}, None)
        for t in range(n_steps):
            a = rng.uniform(-1, 1, size=(2, 2)).astype(np.float32)  # action[0] drives cart, action[1] cart2 (:201,205)
            obs, reward, done, info = env.step(a)
            record(1, obs, reward, done, info, a)
        a = rng.uniform(-1, 1, size=(2, 2)).astype(np.float32)
        obs, reward, done, info = env.step(a)  # 'calling step after done????' (:179-181) when the episode is over
        record(2, obs, reward, done, info, a)
    os.chdir(cwd)
    # only world 0 is observed by the reference (state[r][0] = cart, state[r][1] = pole, :310-311); the second
    # cart-pole's final state is kept to pin that it was driven by action[1]
    res = {k: np.asarray(v) for k, v in out.items()}
    res["final_state"] = np.stack([w.state[0] for w in shim.worlds])
    forces = [(b, fx, fy) for what, b, fx, fy in shim.log if what == "force"]
    res["forces"] = np.asarray(forces, dtype=np.float64)
    order = "".join({"step": "s", "force": "f", "teleport": "t"}[what] for what, *_ in shim.log)
    res["call_order"] = np.frombuffer(order.encode(), dtype=np.uint8)
    res["argv"] = np.asarray(argv, dtype="U")
    res["state_shape"] = np.asarray(env.observation_space.shape)
    return res


CASES = {
    # name: (reference CLI flags, env.steps per episode, action seed)
    "no_theta": (["--no-random-theta", "--initial-force=55", "--max-episode-len=6"], 6, 1),
    "random_theta": (["--max-episode-len=5"], 5, 2),                        # every default: --initial-force=200, R=2, S=1
    "naf_recipe": (["--no-random-theta", "--initial-force=55", "--action-repeats=3", "--steps-per-repeat=4",
                    "--action-force=100", "--max-episode-len=4"], 4, 3),     # README.md:171-172
}

if __name__ == "__main__":
    blob = {"seed": np.asarray(SEED)}
    for name, (argv, n_steps, seed) in CASES.items():
        r = run(argv, n_steps, seed)
        for k, v in r.items():
            blob[f"{name}/{k}"] = v
        print(name, "calls", len(r["call_order"]), "obs", r["obs"].shape, "done", r["done"].astype(int).tolist())
    np.savez_compressed(os.path.join(HERE, "env_golden.npz"), **blob)
    print("written", os.path.join(HERE, "env_golden.npz"))
