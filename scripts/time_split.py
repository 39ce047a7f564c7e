import os, sys, time, torch, numpy as np
sys.path.insert(0, '/root/repo')
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole
def run(split, n=65536, warm_calls=300, K=20, **ov):
    os.environ["BCP_SPLIT"] = str(split)
    env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined", **ov), False, num_envs=n, device=0, dtype="fp32", seed=0)
    env.reset(copy=False)
    t0=time.time(); env.rollout_random(warm_calls); torch.cuda.synchronize(); tw=time.time()-t0
    env.clear_stats()
    pool = torch.rand((K, n, 2), device="cuda")*2-1
    for t in range(3): env.step(pool[t], copy=False)
    torch.cuda.synchronize(); env.clear_stats()
    ev=[torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for t in range(K): env.step(pool[t], copy=False)
    ev[1].record(); torch.cuda.synchronize()
    ms=ev[0].elapsed_time(ev[1])/K
    hist, hdr, _ = env.stats()
    live = hdr["env_steps"]
    print(f"split={split} {ov}: warm {tw:.2f}s; {ms:.3f} ms/step; live frac {live/(n*K):.3f}; live env-steps/s {live/(ms*K*1e-3):.3e}; slot-steps/s {n/(ms*1e-3):.3e}; episodes {hdr['episodes']}")
    env.close()
for split in (0,1):
    run(split)
    run(split, initial_force=55.0, cart_friction=1.0, steps_per_repeat=5)
