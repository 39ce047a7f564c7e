import numpy as np, torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from cartpoleplusplus_b200.env import BatchedBulletCartpole
from cartpoleplusplus_b200.lqr import lqr_opts, BatchedLqrCartpole
n = 4096
for fr in (1.0, 0.3):
    env = BatchedBulletCartpole(lqr_opts(cart_friction=fr), discrete_actions=False, num_envs=n, device=0, dtype="fp32", seed=3)
    lq = BatchedLqrCartpole(env)
    lq.reset()
    G = 8
    K = torch.zeros((n, 2, 8), device="cuda")
    scales = [0.0, 5.0, 20.0, 60.0, 150.0, -20.0, 400.0, 1000.0]
    for g, sc in enumerate(scales):
        sl = slice(g * n // G, (g + 1) * n // G)
        K[sl, 0, lq.PITCH], K[sl, 0, lq.PITCH_DOT] = -sc, -sc / 10
        K[sl, 1, lq.ROLL], K[sl, 1, lq.ROLL_DOT] = sc, sc / 10
    gen = torch.Generator(device="cuda").manual_seed(0)
    kick = (torch.rand((n // G, 2), device="cuda", generator=gen) * 2 - 1) * 20.0
    kick = kick.repeat(G, 1)
    out = lq.rollout(K, 1200, disturbance_fn=lambda t: kick if t < 5 else None, stop_when_all_done=False)
    s = out["steps_survived"].float().view(G, -1)
    print("friction", fr, "scales", scales)
    print(" mean survived", s.mean(1).cpu().numpy().round(1), "median", s.median(1).values.cpu().numpy())
    print(" reasons", [np.bincount(out["first_done_reason"].view(G, -1)[g].cpu().numpy(), minlength=3).tolist() for g in range(G)])
    env.close()
