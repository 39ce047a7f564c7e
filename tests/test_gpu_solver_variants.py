"""-m gpu: the pool-free k_solve_lam_once delivers the pooled k_solve_lam's results bit for bit (same per-lane arithmetic,
DESIGN 4).  The switch is read once per process, so each variant runs in a process of its own."""
import hashlib
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PROG = r'''
import hashlib, sys, torch
sys.path.insert(0, %r)
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole
n = 8192
env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined"), False, num_envs=n, device=0, dtype="fp32", seed=4)
env.reset(copy=False)
env.rollout_random(150)
g = torch.Generator(device="cuda"); g.manual_seed(9)
h = hashlib.sha256()
for t in range(20):
    obs, rew, done, info = env.step(torch.rand((n, 2), device="cuda", generator=g) * 2 - 1)
    h.update(obs.cpu().numpy().tobytes()); h.update(done.cpu().numpy().tobytes())
h.update(env.get_state().tobytes())
print("HASH", h.hexdigest(), env.solver_kind)
''' % ROOT


def _run(extra):
    env = dict(os.environ)
    env.update(extra)
    out = subprocess.run([sys.executable, "-c", PROG], capture_output=True, text=True, env=env, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("HASH")][-1].split()
    return line[1], line[2]


def test_pool_free_solve_phase_is_bit_identical(built):
    a, kind = _run({})
    assert kind == "split/impulse-space"
    b, _ = _run({"BCP_LAM_ONCE": "0"})
    assert a == b
    c, _ = _run({"BCP_GRAPH": "0", "BCP_COHORTS": "1"})  # one pipeline, stream launches: the same per-env programs
    assert a == c
