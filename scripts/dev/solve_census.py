import os, sys, collections, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole
B = 65536
env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined"), discrete_actions=False, num_envs=B, device=0, dtype="fp32", seed=0)
env.reset(copy=False); env.rollout_random(400)
g = torch.Generator(device="cuda").manual_seed(0)
tot = collections.Counter(); longs = collections.Counter(); its = collections.defaultdict(list)
for t in range(12):
    a = torch.rand((B, 2), device="cuda", generator=g) * 2 - 1
    o, r, d, info = env.step(a, copy=False)
    si = env.last_solve_info()
    reason = info["done_reason"].cpu().numpy()
    for key, n in zip(*np.unique(si[:, :3], axis=0, return_counts=True)):
        tot[tuple(key)] += n
    m = si[:, 3] >= 50
    for key, n in zip(*np.unique(si[m][:, :3], axis=0, return_counts=True)):
        longs[tuple(key)] += n
    if t == 11:
        print("resetting share among long:", (reason[m] == 4).mean(), "overall resetting:", (reason == 4).mean())
        print("iteration histogram:", np.bincount(np.minimum(si[:, 3], 50), minlength=51).tolist())
print("regime (gc,cp,gp): count, long count, long frac")
for k, n in sorted(tot.items(), key=lambda kv: -kv[1])[:25]:
    print(k, n, longs.get(k, 0), round(longs.get(k, 0) / n, 4))
print("total long frac", sum(longs.values()) / sum(tot.values()))
gen = {k: (n, longs.get(k, 0)) for k, n in tot.items() if k[1] > 0 and k[2] > 0}
print("generic regimes (cp>0 and gp>0): envs", sum(v[0] for v in gen.values()), "of", sum(tot.values()), "long", sum(v[1] for v in gen.values()), "per call", sum(v[0] for v in gen.values()) / 12)
