"""Where the end-to-end step (step_host) spends its time: H2D, kernels, D2H, host read."""
import argparse, sys, time
import torch
sys.path.insert(0, ".")
import bench
from cartpoleplusplus_b200.env import BatchedBulletCartpole

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=65536)
a = ap.parse_args()
args = argparse.Namespace(envs=a.envs, dtype="fp32", seed=0, initial_force=200.0, repeats=2, steps_per_repeat=1,
                          cart_friction=None, mix_calls=400)
try:
    opts = bench.workload_opts(args, auto_reset="pipelined")
except Exception as e:  # keep the helper usable if bench's argument names change
    print("workload_opts failed:", e)
    raise
n = a.envs
env = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=n, device=0, dtype="fp32", seed=0)
env.reset(copy=False)
env.rollout_random(400)
torch.cuda.synchronize()
ha = torch.empty((n, 2), dtype=torch.float32).pin_memory().uniform_(-1, 1)
for _ in range(3):
    env.step_host(ha)


def ev():
    return torch.cuda.Event(enable_timing=True)


K = 20
acc = {"h2d": 0.0, "kern": 0.0, "d2h": 0.0}
wall = 0.0
host_read = 0.0
for _ in range(K):
    e = [ev() for _ in range(4)]
    t0 = time.perf_counter()
    e[0].record()
    env._d_act.copy_(ha, non_blocking=True)
    e[1].record()
    env.step(env._d_act, copy=False)
    e[2].record()
    env._h_out.copy_(env._out, non_blocking=True)
    e[3].record()
    torch.cuda.current_stream().synchronize()
    t1 = time.perf_counter()
    s = int(env._h_reward.sum().item())
    t2 = time.perf_counter()
    wall += t1 - t0
    host_read += t2 - t1
    acc["h2d"] += e[0].elapsed_time(e[1])
    acc["kern"] += e[1].elapsed_time(e[2])
    acc["d2h"] += e[2].elapsed_time(e[3])
print({k: round(v / K, 4) for k, v in acc.items()}, "ms; wall", round(wall / K * 1e3, 4), "ms; host read",
      round(host_read / K * 1e3, 4), "ms; d2h bytes", env._out.numel(), "GB/s", env._out.numel() / (acc["d2h"] / K * 1e-3) / 1e9)
# launch-side only: how long does the host take to enqueue a step?
t0 = time.perf_counter()
for _ in range(K):
    env.step(env._d_act, copy=False)
t1 = time.perf_counter()
torch.cuda.synchronize()
print("enqueue per step", round((t1 - t0) / K * 1e3, 4), "ms")

import time as _t
torch.cuda.synchronize()
t0 = _t.perf_counter()
for _ in range(50):
    o, r, d, _ = env.step_host(ha)
    s_ = int(r.sum().item())
t1 = _t.perf_counter()
print("step_host (bcp_step_host) wall per step incl. reward read:", round((t1 - t0) / 50 * 1e3, 4), "ms")

import ctypes
vp = ctypes.c_void_p
lib = env._lib
st = env._stream()
ta = tb = 0.0
e0, e1 = ev(), ev()
gpu = 0.0
for _ in range(50):
    t0 = _t.perf_counter()
    e0.record()
    lib.bcp_step_host(env._h, vp(ha.data_ptr()), vp(env._h_obs_rm.data_ptr()), vp(env._h_reward.data_ptr()), vp(env._h_done.data_ptr()), vp(env._h_reason.data_ptr()), st)
    e1.record()
    t1 = _t.perf_counter()
    torch.cuda.current_stream().synchronize()
    t2 = _t.perf_counter()
    lib.bcp_step_host_finish(env._h, vp(env._h_obs_rm.data_ptr()), vp(env._h_reason.data_ptr()), st)
    t3 = _t.perf_counter()
    ta += t1 - t0; tb += t2 - t1; gpu += e0.elapsed_time(e1)
    if _ == 0: first = (t3 - t2)
print("bcp_step_host enqueue", round(ta / 50 * 1e3, 4), "wait", round(tb / 50 * 1e3, 4), "gpu e0->e1", round(gpu / 50, 4), "finish (patch) first", round(first * 1e3, 4), "ms")
