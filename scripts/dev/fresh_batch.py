"""ms per env.step over the first calls after a synchronous reset of the whole batch (every env in the same episode phase:
the worst case for regime waves), in windows of 20 calls.  Usage: python scripts/dev/fresh_batch.py [envs] [calls]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 400
env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined"), False, num_envs=n, device=0, dtype="fp32", seed=0)
env.reset(copy=False)
g = torch.Generator(device="cuda"); g.manual_seed(1)
pool = torch.rand((16, n, 2), device="cuda", generator=g) * 2 - 1
ev = [torch.cuda.Event(enable_timing=True) for _ in range(calls + 1)]
torch.cuda.synchronize()
ev[0].record()
live = []
for t in range(calls):
    _, r, _, _ = env.step(pool[t % 16], copy=False)
    ev[t + 1].record()
    if t % 20 == 19:
        live.append(float(r.mean().item()))
torch.cuda.synchronize()
ms = [ev[t].elapsed_time(ev[t + 1]) for t in range(calls)]
for w in range(0, calls, 20):
    seg = ms[w:w + 20]
    print(f"calls {w:4d}-{w + 19:4d}: mean {sum(seg) / len(seg):.3f} ms  max {max(seg):.3f} ms  live fraction at end {live[w // 20]:.2f}")
if os.environ.get("FRESH_VERBOSE"):
    big = [(t, round(ms[t], 3)) for t in range(calls) if ms[t] > 0.6]
    print("calls above 0.6 ms:", big)
