"""configs[3]-style rollout (one episode per env, in-kernel random agent) through the split pipeline and through the
one-launch k_rollout: device time and env-steps/s.  Usage: python scripts/dev/rollout_paths.py [envs]"""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole
n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
for split in ("1", "0"):
    os.environ["BCP_ROLLOUT_SPLIT"] = split
    env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined", initial_force=55.0), discrete_actions=False, num_envs=n, device=0, dtype="fp32", seed=11)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    header = {"episodes": 0}
    calls = 0
    while header["episodes"] < n and calls < 12:
        env.rollout_random(64, max_episodes=1)
        calls += 1
        hist, header, rsum = env.stats()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print("split" if split == "1" else "k_rollout", n, "envs:", header, f"{ms:.1f} ms, {header['env_steps'] / ms * 1e3:.3e} env-steps/s")
    env.close()
