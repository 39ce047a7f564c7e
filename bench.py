#!/usr/bin/env python
"""bench.py — env-steps/sec of the batched BulletCartpole step path on B200.

A "step" is one `env.step` over the whole batch (BASELINE.json configs[2]: 65536 envs per
GPU, continuous 2-D force, --action-force=100, uniform random actions, low-dim (R,2,7)
obs) followed by the reset of the envs that finished (130 substeps each, compacted into
dense warps).  Contract and JSON keys: see the task statement / DESIGN.md section 6.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nproc-per-node N bench.py --gpus N ...        (one rank per GPU, NCCL)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from cartpoleplusplus_b200 import default_opts, make_config  # noqa: E402
from cartpoleplusplus_b200 import config as C  # noqa: E402

WORKLOAD = "configs[2]: 65536 envs/GPU, continuous 2-D force, action_force=100, U(-1,1)^2 actions, (R,2,7) obs"


def workload_opts(args):
    return default_opts(action_force=100.0, initial_force=args.initial_force, action_repeats=args.repeats,
                        steps_per_repeat=args.steps_per_repeat)


# ---------------------------------------------------------------------------------------
# flop / byte model of one substep (FMA = 2 flops, useful work only)
# ---------------------------------------------------------------------------------------
F_FIXED = 2400.0        # poses->matrices, 2x forward dynamics, 3x SAT (+clip), manifolds, 2x integrate
F_SETUP_2, F_SETUP_1 = 250.0, 150.0   # contact-row setup: two dynamic bodies / one
F_SWEEP_2, F_SWEEP_1 = 60.0, 36.0     # one PGS row update: two dynamic bodies / one


def flops_per_substep(rows_cart_pole, rows_ground, iters):
    return (F_FIXED + rows_cart_pole * F_SETUP_2 + rows_ground * F_SETUP_1
            + iters * (rows_cart_pole * F_SWEEP_2 + rows_ground * F_SWEEP_1))


def bytes_per_env_step(cfg, mean_points):
    """Algorithmic HBM bytes of one env.step launch per env: state in + out, action, obs, flags."""
    sz = 8 if cfg.dtype == C.DTYPE_F64 else 4
    state = (32 + mean_points * C.MANI_POINT) * sz + 6 * 4
    act = 4 if cfg.discrete_actions else 8
    return 2 * state + act + cfg.action_repeats * 56 + 4 + 1 + 1


# ---------------------------------------------------------------------------------------
# CPU baseline: the oracle (a port - pybullet is unavailable on this box), all host threads
# ---------------------------------------------------------------------------------------
def cpu_baseline(args, steps=None, n_envs=None):
    from oracle import Oracle
    cores = os.cpu_count() or 1
    n = n_envs or 128 * cores
    steps = steps or 12
    opts = workload_opts(args)
    cfg = make_config(opts, False, num_envs=n, dtype="fp64", seed=args.seed)
    ora = Oracle(cfg, dtype="fp64")
    rng = np.random.default_rng(args.seed)
    ora.reset()
    acts = rng.uniform(-1, 1, size=(steps, n, 2)).astype(np.float32)
    t0 = time.perf_counter()
    resets = 0
    for t in range(steps):
        _, _, done, _ = ora.step(acts[t])
        if done.any():
            ora.reset(done.astype(np.uint8))
            resets += int(done.sum())
    dt = time.perf_counter() - t0
    return {"value": n * steps / dt, "unit": "env-steps/s", "cores": cores, "kind": "port",
            "sample": f"{n} envs x {steps} env.steps (+{resets} resets of 130 substeps) of the same workload, "
                      f"fp64 C oracle, OpenMP over {cores} threads; pybullet: unavailable on this box",
            "seconds": dt, "resets": resets}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        rows = [r for r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons,
                "power_w_max": max(float(r[2]) for r in rows), "samples": len(rows)}


# ---------------------------------------------------------------------------------------
# the measured arm
# ---------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from cartpoleplusplus_b200 import _lib
    from cartpoleplusplus_b200.env import BatchedBulletCartpole

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    n = args.envs
    opts = workload_opts(args)
    env = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=n, device=local, dtype=args.dtype,
                                seed=args.seed, env_id_offset=rank * n)
    K, W = args.steps, args.warmup
    gen = torch.Generator(device=dev)
    gen.manual_seed(args.seed + rank)
    pool = torch.rand((W + K, n, 2), device=dev, generator=gen) * 2 - 1  # inputs resident in HBM
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > L2 (126 MB)
    env.reset(copy=False)

    def one_step(t, ev=None):
        flush.zero_()  # untimed L2 flush between steps
        if ev:
            ev[0].record()
        env.step(pool[t], copy=False)
        if ev:
            ev[1].record()
        env.reset(mask=env.done, copy=False)
        if ev:
            ev[2].record()

    for t in range(W):
        one_step(t)
    torch.cuda.synchronize()
    env.clear_stats()
    launches0 = env.launch_count
    sampler = ClockSampler(local)
    sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
    wall0 = time.perf_counter()
    for t in range(K):
        one_step(W + t, evs[t])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    wall = time.perf_counter() - wall0
    sampler.stop_flag = True
    step_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    reset_ms = sum(e[1].elapsed_time(e[2]) for e in evs)
    total_ms = step_ms + reset_ms
    launches = env.launch_count - launches0
    hist, header, _ = env.stats()

    # ---- e2e: same metric through step_host (pinned host buffers, H2D + D2H inside) ----
    host_actions = [torch.empty((n, 2), dtype=torch.float32).pin_memory() for _ in range(2)]
    for h in host_actions:
        h.uniform_(-1, 1)
    Ke = max(3, min(K, 20))
    for t in range(2):
        env.step_host(host_actions[t % 2], auto_reset=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for t in range(Ke):
        env.step_host(host_actions[t % 2], auto_reset=True)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0

    # ---- max over ranks, whole-job aggregate; the one collective: histogram allreduce ----
    tt = torch.tensor([total_ms, step_ms, reset_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    hist_t = torch.from_numpy(hist).to(dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(hist_t, op=dist.ReduceOp.SUM)
    total_ms, step_ms, reset_ms, e2e_ms = (float(v) for v in tt.tolist())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (k_step): flops from a sampled state census ----
    from oracle import Oracle
    ns = min(n, 4096)
    st = env.get_state()[:ns].copy()
    st[:, C.S_DONE] = 0
    scfg = C.copy_config(env.cfg)
    scfg.num_envs = ns
    ora = Oracle(scfg, dtype="fp64")
    ora.state[:] = st
    info = ora.substep(with_info=True)
    iters = float(info[:, 0].mean())
    rows_gc, rows_cp, rows_gp = (float(info[:, 3 + k].mean()) for k in range(3))
    f_sub = flops_per_substep(rows_cp, rows_gc + rows_gp, iters)
    RS = env.cfg.action_repeats * env.cfg.steps_per_repeat
    flops_launch = f_sub * RS * n
    step_launch_ms = step_ms / K
    lib = _lib.load()
    import ctypes
    peak32, peak64 = ctypes.c_double(), ctypes.c_double()
    lib.bcp_bench_fma_peak(local, 0, 4096, ctypes.byref(peak32))
    lib.bcp_bench_fma_peak(local, 1, 1024, ctypes.byref(peak64))
    pipe_peak = peak64.value if args.dtype == "fp64" else peak32.value
    achieved_tf = flops_launch / (step_launch_ms * 1e-3) / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    bytes_launch = bytes_per_env_step(env.cfg, rows_gc + rows_cp + rows_gp) * n
    achieved_gbs = bytes_launch / (step_launch_ms * 1e-3) / 1e9
    value = world * n * K / (total_ms * 1e-3)
    out = {
        "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.dtype == "fp32" else "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "envs_per_gpu": n, "action_repeats": env.cfg.action_repeats,
                   "steps_per_repeat": env.cfg.steps_per_repeat, "initial_force": args.initial_force,
                   "termination": "angle/position/200-step rule on", "auto_reset": "compacted reset kernel per step",
                   "l2": "flushed between steps (256 MiB memset, untimed)", "parallelism": f"env-sharded x{world}",
                   "seed": args.seed},
        "substeps_per_s": (world * n * K * RS + int(hist_t.sum().item()) * (env.cfg.settle_steps + env.cfg.initial_force_steps)) / (total_ms * 1e-3),
        "step_kernel_ms": step_launch_ms, "reset_ms_per_step": reset_ms / K,
        "episodes_finished": int(hist_t.sum().item()),
        "e2e": {"value": world * n * Ke / (e2e_ms * 1e-3), "unit": "env-steps/s",
                "h2d_bytes_per_step": env.h2d_bytes_per_step, "d2h_bytes_per_step": env.d2h_bytes_per_step},
        "gpu_launches": int(launches),
        "clocks": sampler.summary(),
        "roofline": {"bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved_gbs / hbm_peak, "traffic": None,
                     "note": "schema-mandated HBM view of k_step; the kernel is FP32-pipe/latency bound, see roofline_pipe"},
        "roofline_pipe": {"bound": "fp32-cuda-core" if args.dtype == "fp32" else "fp64-cuda-core",
                          "achieved": achieved_tf, "peak": pipe_peak, "unit": "TFLOP/s", "frac": achieved_tf / pipe_peak,
                          "peak_source": "bcp_bench_fma_peak (FMA chains, measured in this run)",
                          "flops_per_substep": f_sub, "mean_solver_iterations": iters,
                          "mean_rows": {"ground_cart": rows_gc, "cart_pole": rows_cp, "ground_pole": rows_gp}},
    }
    if not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline(args)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    """--impl reference: the reference's CPU path.  pybullet is unavailable on this box, so this
    times the oracle port (the one other place bench.py may execute oracle/), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    K, W = args.steps, args.warmup
    n = 64 * cores
    per = max(2, min(8, 200 // max(1, K + W)))  # env.steps per bench "step" (bounded sample)
    vals = []
    for t in range(W + K):
        b = cpu_baseline(args, steps=per, n_envs=n)
        if t >= W:
            vals.append(b)
    total_steps = sum(n * per for _ in vals)
    total_s = sum(b["seconds"] for b in vals)
    v = total_steps / total_s
    print(json.dumps({
        "impl": "reference", "metric": "env-steps/sec", "value": v, "unit": "env-steps/s",
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": K, "warmup": W,
        "ms_per_step": 1e3 * total_s / max(1, K), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample_envs": n, "env_steps_per_bench_step": per},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                         "sample": f"{n} envs x {per} env.steps per bench step incl. resets; C oracle (fp64, OpenMP); "
                                   "pybullet: unavailable on this box"},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs", type=int, default=65536, help="envs per GPU")
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--initial-force", type=float, default=200.0)
    ap.add_argument("--repeats", type=int, default=2)
    ap.add_argument("--steps-per-repeat", type=int, default=1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
