"""Experiment: does running two half batches concurrently on two streams beat one full batch?  (k_prep and the solve
phase are both latency-bound; staggered, one half's solve tail could hide under the other half's k_prep.)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole

def mk(n, off):
    env = BatchedBulletCartpole(default_opts(action_force=100.0, auto_reset="pipelined"), False, num_envs=n, device=0, dtype="fp32", seed=0, env_id_offset=off)
    env.reset(copy=False); env.rollout_random(400); return env

def run(parts, K=200):
    n = sum(e.num_envs for e in parts)
    streams = [torch.cuda.Stream() for _ in parts]
    acts = [torch.rand((8, e.num_envs, 2), device="cuda") * 2 - 1 for e in parts]
    torch.cuda.synchronize()
    for rep in range(2):
        live = [torch.zeros((), device="cuda", dtype=torch.float64) for _ in parts]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for s in streams: s.wait_stream(torch.cuda.current_stream())
        for t in range(K):
            for i, (e, s) in enumerate(zip(parts, streams)):
                with torch.cuda.stream(s):
                    e.step(acts[i][t % 8], copy=False)
                    live[i] += e.reward.sum()
        for s in streams: torch.cuda.current_stream().wait_stream(s)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    tot = sum(float(l.item()) for l in live)
    return tot / (ms * 1e-3), ms / K

for split in ([65536], [32768, 32768], [16384] * 4, [131072], [65536, 65536]):
    parts = []; off = 0
    for n in split:
        parts.append(mk(n, off)); off += n
    v, ms = run(parts)
    print(split, "env-steps/s %.4g" % v, "ms per combined step %.4f" % ms, flush=True)
    for e in parts: e.close()
