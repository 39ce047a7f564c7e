import json, sys, os
sys.path[:0] = ["/root/repo", "/root/repo/tests"]
import numpy as np
import parity_util as PU
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200 import config as C
from cartpoleplusplus_b200.env import BatchedBulletCartpole
from oracle import Oracle
n=65536
ov = dict(initial_force=200.0, action_force=100.0, auto_reset="pipelined")
env = BatchedBulletCartpole(default_opts(**ov), discrete_actions=False, num_envs=n, device=0, dtype="fp32", seed=17)
ora = Oracle(env.cfg, dtype="fp64")
state = PU.stationary_mix(env, 400)
rng = np.random.default_rng(23)
acts = rng.uniform(-1, 1, size=(n, 2)).astype(np.float32)
rep = PU.classify_step(env, ora, state, acts, rng=rng)
dev=rep.pop("dev"); spread=rep.pop("spread")
print(json.dumps(rep)[:3000])
un=rep["unexplained"]
# redo to get details
env.set_state(state); ora.state[:]=state
import torch
env.solve_trace(True)
env.step(torch.from_numpy(acts).cuda()); torch.cuda.synchronize()
tr=env.solve_trace(); sg=env.get_state()
r=ora.step(acts, with_info=True); so=ora.state
for e in un[:25]:
    print("env",e,"dev %.3e spread %.3e"%(dev[e],spread[e]),"reason",r[3][e],"reset_left",state[e,C.S_RESET_LEFT],"gpu trace",tr[:,e,:].tolist(),"ora info",r[4][e].tolist())
    d=np.abs(sg[e]-so[e]); k=np.argsort(d[:32])[::-1][:4]
    print("   worst fields",[(int(i),float(d[i])) for i in k], "pole z", so[e,C.S_POLE+2], "cart xy", so[e,0:2])
out = np.flatnonzero(dev > 1e-4)
os.makedirs("gpurun_out", exist_ok=True)
np.savez("gpurun_out/outliers_65536.npz", idx=out, state=state[out], acts=acts[out], gpu=sg[out], ora=so[out], trace=tr[:, out, :], info=r[4][out], dev=dev[out], spread=spread[out])
