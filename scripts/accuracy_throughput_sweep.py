#!/usr/bin/env python
"""FP64 parity mode vs FP32 fast mode on the same batch (BASELINE.json configs[4]: discrete
random actions): how far the fp32 poses drift from the fp64 ones as a function of env.steps
from identical start states with identical action sequences, and what each mode costs.

One JSON line on stdout.  Needs a GPU.  `--envs` per mode AND PER GPU (default 65536 discrete envs).
Under torchrun (configs[4] at size: `torchrun --nproc-per-node 8 scripts/accuracy_throughput_sweep.py --envs 1048576` =
8 M envs) every rank sweeps its own shard (global env ids by rank); rank 0 merges: throughputs add up over the slowest
rank's time, drift quantiles are taken over the ranks' (max of the maxima and q99s, median of the medians)."""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cartpoleplusplus_b200 import default_opts  # noqa: E402
from cartpoleplusplus_b200.env import BatchedBulletCartpole  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=65536)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--initial-force", type=float, default=55.0)
    ap.add_argument("--cart-friction", type=float, default=0.0)
    ap.add_argument("--seed", type=int, default=0)
    args = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    opts = default_opts(initial_force=args.initial_force, cart_friction=args.cart_friction)
    envs = {m: BatchedBulletCartpole(opts, discrete_actions=True, num_envs=args.envs, device=local, dtype=m, seed=args.seed,
                                     env_id_offset=rank * args.envs)
            for m in ("fp64", "fp32")}
    # identical start: the fp64 reset state, rounded into the fp32 env
    envs["fp64"].reset(copy=False)
    envs["fp32"].reset(copy=False)
    envs["fp32"].set_state(envs["fp64"].get_state())
    g = torch.Generator(device=dev)
    g.manual_seed(args.seed + rank)
    acts = torch.randint(0, 5, (args.steps, args.envs), device=dev, dtype=torch.int32, generator=g)
    checkpoints = sorted({1, 2, 5, 10, 20, 40, args.steps})
    drift, ms = [], {"fp64": 0.0, "fp32": 0.0}
    alive = torch.ones(args.envs, dtype=torch.bool, device=dev)
    for t in range(args.steps):
        out = {}
        for m, env in envs.items():
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            obs, _, done, _ = env.step(acts[t], copy=True)
            e1.record()
            e1.synchronize()
            ms[m] += e0.elapsed_time(e1)
            out[m] = (obs, done)
        alive &= ~(out["fp64"][1] | out["fp32"][1])   # compare envs still running in both modes
        if (t + 1) in checkpoints and bool(alive.any()):
            d = (out["fp64"][0][alive] - out["fp32"][0][alive]).abs().amax(dim=(1, 2, 3)).double().cpu().numpy()
            drift.append({"env_steps": t + 1, "envs_compared": int(alive.sum()), "median": float(np.median(d)),
                          "q99": float(np.quantile(d, 0.99)), "max": float(d.max())})
    res = {"what": "fp32 fast mode vs fp64 parity mode, identical start states and discrete random actions",
           "envs": args.envs, "steps": args.steps, "initial_force": args.initial_force, "cart_friction": args.cart_friction,
           "pose_abs_error_vs_fp64": drift,
           "env_steps_per_s": {m: args.envs * args.steps / (ms[m] * 1e-3) for m in ms},
           "ms_per_env_step_call": {m: ms[m] / args.steps for m in ms},
           "note": "slot throughput over the first steps after a reset (all envs live, no resets in the loop)"}
    if world > 1:
        parts = [None] * world
        dist.all_gather_object(parts, {"drift": drift, "ms": ms})
        if rank == 0:
            slow = {m: max(p["ms"][m] for p in parts) for m in ms}
            merged = []
            for k, d0 in enumerate(drift):
                rows = [p["drift"][k] for p in parts if len(p["drift"]) > k]
                merged.append({"env_steps": d0["env_steps"], "envs_compared": sum(r["envs_compared"] for r in rows),
                               "median": float(np.median([r["median"] for r in rows])), "q99": max(r["q99"] for r in rows),
                               "max": max(r["max"] for r in rows)})
            res.update({"n_gpus": world, "envs_total": world * args.envs, "pose_abs_error_vs_fp64": merged,
                        "env_steps_per_s": {m: world * args.envs * args.steps / (slow[m] * 1e-3) for m in ms},
                        "ms_per_env_step_call": {m: slow[m] / args.steps for m in ms}})
        dist.barrier()
        dist.destroy_process_group()
        if rank != 0:
            return
    print(json.dumps(res))


if __name__ == "__main__":
    main()
