import os, sys, torch
sys.path.insert(0, '/root/repo')
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole
n=65536
env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined"), False, num_envs=n, device=0, dtype="fp32", seed=0)
env.reset(copy=False)
env.rollout_random(300)
pool = torch.rand((4, n, 2), device="cuda")*2-1
for t in range(4): env.step(pool[t], copy=False)
torch.cuda.synchronize()
