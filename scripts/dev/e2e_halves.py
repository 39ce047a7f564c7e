"""Where the double-buffered host loop spends its time: begin (enqueue) vs end (wait + patch) per half step."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from cartpoleplusplus_b200 import default_opts
from cartpoleplusplus_b200.env import BatchedBulletCartpole
n = 65536
opts = default_opts(action_force=100.0, auto_reset="pipelined")
def mk(m, off):
    e = BatchedBulletCartpole(opts, False, num_envs=m, device=0, dtype="fp32", seed=0, env_id_offset=off)
    e.reset(copy=False); e.rollout_random(400); return e
for parts in ([n], [n // 2, n // 2], [n // 4] * 4):
    envs = []; off = 0
    for m in parts:
        envs.append(mk(m, off)); off += m
    acts = [torch.empty((m, 2)).pin_memory().uniform_(-1, 1) for m in parts]
    for e, a in zip(envs, acts):
        e.step_host(a); e.step_host(a)
    torch.cuda.synchronize()
    K = 100
    tb = te = tr = 0.0; live = 0
    t0 = time.perf_counter()
    if len(envs) == 1:
        for t in range(K):
            a0 = time.perf_counter(); envs[0].step_host_begin(acts[0]); a1 = time.perf_counter()
            o, r, d, i = envs[0].step_host_end(); a2 = time.perf_counter()
            live += int(r.sum().item()); a3 = time.perf_counter()
            tb += a1 - a0; te += a2 - a1; tr += a3 - a2
    else:
        k = len(envs)
        for j in range(k - 1):
            envs[j].step_host_begin(acts[j])
        for t in range(K):
            for j in range(k):
                nxt = (j + k - 1) % k
                a0 = time.perf_counter(); envs[nxt].step_host_begin(acts[nxt]); a1 = time.perf_counter()
                o, r, d, i = envs[j].step_host_end(); a2 = time.perf_counter()
                live += int(r.sum().item()); a3 = time.perf_counter()
                tb += a1 - a0; te += a2 - a1; tr += a3 - a2
        for j in range(k - 1):
            envs[j].step_host_end()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(parts, "e2e env-steps/s %.4g" % (live / dt), "per full step ms: total %.3f begin %.3f end-wait %.3f read %.3f" % (dt / K * 1e3, tb / K * 1e3, te / K * 1e3, tr / K * 1e3), flush=True)
    for e in envs: e.close()
