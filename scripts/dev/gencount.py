import sys, torch
sys.path.insert(0, '/root/repo')
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole
n=65536
env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined"), False, num_envs=n, device=0, dtype="fp32", seed=0)
env.reset(copy=False)
env.rollout_random(400)
st = env.get_state()
import numpy as np
from cartpoleplusplus_b200 import config as C
print("state shape", st.shape)
ids = env.get_ids() if hasattr(env, "get_ids") else None
tr = None
try:
    env.solve_trace(True)
    a = torch.rand((n,2), device="cuda")*2-1
    env.step(a)
    tr = env.solve_trace()
except Exception as e:
    print("trace err", e)
if tr is not None:
    tr = np.asarray(tr)
    print("trace shape", tr.shape)
    # words: counts gc, cp, gp, iters
    for k in range(tr.shape[0]):
        t = tr[k]
        gc, cp, gp, it = t[:,0], t[:,1], t[:,2], t[:,3]
        gen = (cp>0)&(gp>0)
        print("substep", k, "generic", int(gen.sum()), "iters of generic: mean", float(it[gen].mean()) if gen.any() else 0, "at 50:", int((it[gen]>=50).sum()), "| stack", int(((gp==0)&(cp>0)).sum()), "split", int((cp==0).sum()))
