"""Soak: 3 000 env.steps of the bench workload (65 536 envs, cohorts + CUDA graphs, pipelined resets) with bookkeeping checks."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole
n = 65536
env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined"), False, num_envs=n, device=0, dtype="fp32", seed=1)
env.reset(copy=False); env.clear_stats()
g = torch.Generator(device="cuda").manual_seed(0)
live = torch.zeros((), dtype=torch.float64, device="cuda"); ended = torch.zeros((), dtype=torch.int64, device="cuda")
bad = torch.zeros((), dtype=torch.int64, device="cuda")
for t in range(3000):
    a = torch.rand((n, 2), device="cuda", generator=g) * 2 - 1
    obs, rew, done, info = env.step(a, copy=False)
    live += rew.sum(); ended += ((rew > 0) & done).sum()
    if t % 50 == 0:
        q = obs[..., 3:7]
        bad += (~torch.isfinite(obs)).sum() + ((q.norm(dim=-1) - 1).abs() > 1e-3).sum()
hist, hd, _ = env.stats()
print(json.dumps({"steps": 3000, "live_env_steps": float(live.item()), "episodes": int(ended.item()), "stats_env_steps": hd["env_steps"], "stats_episodes": hd["episodes"],
                  "hist_sum": int(hist.sum()), "bad_values": int(bad.item()), "mean_len": float((hist * range(len(hist))).sum() / max(1, hist.sum()))}))
assert int(bad.item()) == 0 and hd["env_steps"] == int(live.item()) and hd["episodes"] == int(ended.item()) == int(hist.sum())
