#!/usr/bin/env python
"""bench.py — env-steps/sec of the batched BulletCartpole step path on B200.

Workload (BASELINE.json configs[2]): 65536 envs per GPU, continuous 2-D force,
--action-force=100, uniform random actions, low-dim (R,2,7) observations, the reference's
own models and defaults otherwise (R=2, S=1, --initial-force=200, termination rule on).
Envs auto-reset in the pipelined mode: a finished env runs its
130-substep reset R*S substeps per call next to the live envs, so every reset substep is
inside the timed region.  A bench "step" is one `env.step` over the whole batch; `value`
counts only the env.steps actually taken (reward 1) - slots that are resetting do not count.
Before timing, the batch is rolled to its stationary mix of episode phases (untimed).

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

WORKLOAD = ("configs[2]: 65536 envs/GPU, continuous 2-D force, action_force=100, U(-1,1)^2 actions, "
            "(R,2,7) obs, reference models/defaults")


def workload_opts(args, **extra):
    return default_opts(action_force=100.0, initial_force=args.initial_force, action_repeats=args.repeats,
                        steps_per_repeat=args.steps_per_repeat, cart_friction=args.cart_friction,
                        no_carry_over_reset=args.no_carry_over_reset, reset_settle_cache=args.reset_settle_cache, **extra)


# ---------------------------------------------------------------------------------------
# Algorithmic work per unit.
#  flops: profiles/flops_model.json - COUNTED on the parity substep (the oracle's own operation sequence) by
#         scripts/count_flops.py: flops(substep) = fixed[gc,cp,gp] + iterations * pgs_per_iteration[gc,cp,gp], FMA = 2,
#         useful work only.  Evaluated here on the kernel's own solve records (contact counts, iterations) - no oracle.
#  bytes: SURVEY.md 8d - per env-step of the step-per-launch API: action in, (R,2,7) obs + reward + done + reason out, and
#         the env's state once in and once out (26 body scalars + 6 pending force / torque + 3 x 4 x 11 manifold scalars,
#         8 counters): what a single-launch step would have to move.
# ---------------------------------------------------------------------------------------
def load_flops_model(cfg):
    path = os.path.join(ROOT, "profiles", "flops_model.json")
    blob = json.load(open(path))
    friction = (cfg.cart.friction * cfg.pole.friction) != 0.0
    name = "upstream_style_friction" if friction else "configs2_fork_defaults"
    return name, blob["workloads"][name]


def substep_flops(model, counts):
    """counts (m, 4) int: gc, cp, gp contact points and PGS iterations of m env-substeps -> (fixed flops, pgs flops) per row"""
    fixed = np.empty(len(counts))
    pgs = np.empty(len(counts))
    keys, inv = np.unique(counts[:, :3], axis=0, return_inverse=True)
    inv = inv.reshape(-1)
    for k, key in enumerate(keys):
        r = model["regimes"].get("%d,%d,%d" % tuple(int(v) for v in key), model["fallback"])
        m = inv == k
        fixed[m] = r["fixed_total"]
        pgs[m] = r["pgs_per_iteration"] * counts[m, 3]
    return fixed, pgs


def algorithmic_bytes_per_env_step(cfg):
    state = (26 + 6 + 3 * 4 * 11) * 4 + 8 * 4
    action = 4 if cfg.discrete_actions else 8
    return 2 * state + action + cfg.action_repeats * 56 + 4 + 1 + 1


# ---------------------------------------------------------------------------------------
# CPU baseline: the oracle (a port - pybullet is unavailable on this box), all host threads
# ---------------------------------------------------------------------------------------
def cpu_rollout(args, n_envs, n_steps, warm_steps):
    """Oracle random-agent rollout with synchronous in-place resets: n_envs x n_steps env.steps
    (+ the 130-substep resets they trigger), OpenMP over all host threads.  Same workload,
    same action distribution, from the stationary phase mix."""
    from oracle import Oracle
    from oracle.oracle import set_threads
    set_threads(os.cpu_count() or 1)  # torchrun exports OMP_NUM_THREADS=1; the baseline uses every host thread
    cfg = make_config(workload_opts(args), False, num_envs=n_envs, dtype="fp64", seed=args.seed)
    ora = Oracle(cfg, dtype="fp64")
    ora.state[:, C.S_DONE] = 1  # rollout_random resets every env first
    ora.rollout_random(warm_steps)
    h0 = ora.header.copy()
    t0 = time.perf_counter()
    ora.rollout_random(n_steps)
    dt = time.perf_counter() - t0
    steps = int(ora.header[1] - h0[1])
    resets = int(ora.header[3] - h0[3])
    return steps, resets, dt


def cpu_baseline(args, budget_s=None):
    budget_s = budget_s or args.cpu_budget
    cores = os.cpu_count() or 1
    n = 256 * cores
    steps, resets, dt = cpu_rollout(args, n, 8, 40)  # probe
    rate = steps / dt
    n_steps = int(max(8, min(4000, budget_s * rate / n)))
    steps, resets, dt = cpu_rollout(args, n, n_steps, 40)
    return {"value": steps / dt, "unit": "env-steps/s", "cores": cores, "kind": "port",
            "sample": f"{n} envs x {n_steps} env.steps (+{resets} resets of 130 substeps) of the same workload from its "
                      f"stationary phase mix, fp64 C oracle, OpenMP over {cores} threads; pybullet: unavailable on this box",
            "seconds": dt, "resets": resets}


SECONDARY = {
    # the workload where the pole is actually balanced on the cart (two-body friction rows live, ~31 PGS iterations per substep)
    "upstream_style_workload": (["--cart-friction", "1.0", "--steps-per-repeat", "5", "--initial-force", "55", "--mix-calls", "60"],
                                "cart friction 1.0, R*S = 10, --initial-force=55 (SURVEY 8d asks for both workloads, each labelled)"),
    # configs[2] with the opt-in fast reset: a DEVIATION from pybullet's reset (the pending force and the contact caches are
    # cleared, --no-carry-over-reset), under which the 100 settle substeps are the same computation in every env and are
    # replayed from a per-handle cache (--reset-settle-cache): a reset costs its 30 push substeps.  Never the headline.
    "fast_reset_workload": (["--no-carry-over-reset", "--reset-settle-cache"],
                            "configs[2] with --no-carry-over-reset --reset-settle-cache: deviates from pybullet's reset semantics (opt-in)"),
}


def secondary_workload(args, name):
    """A labelled workload measured by this same script in a child process and reported BESIDE the headline, never as it."""
    extra, what = SECONDARY[name]
    cmd = [sys.executable, os.path.abspath(__file__), *extra, "--steps", "10", "--warmup", "3", "--envs", str(args.envs),
           "--no-secondary", "--cpu-budget", "6"]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        d = json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:
        return {"unavailable": str(e)[:200]}
    keep = ("value", "unit", "ms_per_step", "steps", "live_fraction", "substeps_per_s", "kernel_ms_per_step", "e2e", "cpu_baseline")
    out = {"what": what}
    out.update({k: d.get(k) for k in keep})
    out["config"] = {k: d["config"].get(k) for k in ("envs_per_gpu", "action_repeats", "steps_per_repeat", "initial_force", "cart_friction", "pipeline", "reset")}
    out["roofline"] = {k: d["roofline"].get(k) for k in ("bound", "whole_step_frac", "whole_step_achieved", "peak", "unit")}
    out["census"] = d.get("census")
    return out


class ClockSampler(threading.Thread):
    """SM clock, power and throttle reasons sampled DURING the timed regions.

    NVML in-process (about a millisecond per sample, so that even a 15 ms timed region gets several);
    the nvidia-smi query of the profiling recipe is the fallback when NVML cannot be loaded."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index, uuid=None, interval=0.002):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.interval = index, [], False, interval
        self.nvml = self.handle = None
        self.source = "nvidia-smi"
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode()) if uuid else None
            except Exception:
                self.handle = None
            if self.handle is None:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nvml, self.source = pynvml, "nvml"
        except Exception:
            self.nvml = None

    def sample(self):
        """-> [sm_mhz, sm_max_mhz, power_w, hw_slowdown, hw_thermal, sw_thermal, sw_power_cap] (flags 0/1)"""
        if self.nvml is not None:
            n, h = self.nvml, self.handle
            sm = n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)
            mx = n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM)
            try:
                pw = n.nvmlDeviceGetPowerUsage(h) / 1000.0
            except Exception:
                pw = 0.0
            r = n.nvmlDeviceGetCurrentClocksEventReasons(h)
            bits = [n.nvmlClocksEventReasonHwSlowdown, n.nvmlClocksEventReasonHwThermalSlowdown,
                    n.nvmlClocksEventReasonSwThermalSlowdown, n.nvmlClocksEventReasonSwPowerCap]
            return [float(sm), float(mx), pw] + [1 if r & b else 0 for b in bits]
        out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                              "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
        c = [x.strip() for x in out.strip().split(",")]
        return [float(c[0]), float(c[1]), float(c[2])] + [1 if x.lower().startswith("active") else 0 for x in c[3:7]]

    def run(self):
        while not self.stop_flag:
            try:
                self.rows.append(self.sample())
            except Exception:
                pass
            time.sleep(self.interval if self.nvml is not None else 0.05)

    def summary(self):
        self.stop_flag = True
        if self.is_alive():
            self.join(timeout=6)
        rows = [r for r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"], "source": self.source}
        sm = sorted(r[0] for r in rows)
        reasons = [n for i, n in enumerate(self.NAMES) if any(r[3 + i] for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": rows[0][1], "reasons": reasons,
                "power_w_max": max(r[2] for r in rows), "samples": len(rows), "source": self.source}


def pin_near_gpu(local, world):
    """Host placement of this rank: its threads on the cores of the GPU's NUMA node (an even share of them when several
    ranks live on the node), so that the pinned buffers it allocates next are first-touched there and the D2H / H2D copies
    and the host-side reads of the outputs do not cross sockets.  torchrun exports OMP_NUM_THREADS=1; the torch thread
    count is set to the share.  Returns a description for the JSON line."""
    import torch
    try:
        bus = torch.cuda.get_device_properties(local).pci_bus_id
        dom = torch.cuda.get_device_properties(local).pci_domain_id
        dev = torch.cuda.get_device_properties(local).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, dev)
        node = int(open(path).read().strip())
        if node < 0:
            node = 0
        cpus = []
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus += list(range(int(a), int(b or a) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return {"numa_node": node, "pinned": False}
        # ranks sharing the node take disjoint slices
        same = [r for r in range(world) if _gpu_numa(r) == node] if world > 1 else [local]
        k, m = same.index(local) if local in same else 0, max(1, len(same))
        share = allowed[k * len(allowed) // m:(k + 1) * len(allowed) // m] or allowed
        os.sched_setaffinity(0, share)
        # the host side of a step is a 65 536-float sum and a byte scan: one thread does it; more threads only spin on
        # the few cores a rank has when several ranks share the host (N = 8 on 32 cores: 4 each)
        # (measured: N = 2, e2e / value 0.57 with 4 torch threads per rank, 0.72 with 1; N = 1 with 16 cores to itself: 0.78 vs 0.74)
        torch.set_num_threads(int(os.environ.get("BCP_BENCH_TORCH_THREADS", "1" if world > 1 else str(max(1, min(4, len(share)))))))
        return {"numa_node": node, "pinned": True, "cpus": len(share), "torch_threads": torch.get_num_threads()}
    except Exception as e:  # no sysfs / no permission: run unpinned and say so
        return {"pinned": False, "why": str(e)[:80]}


def _gpu_numa(idx):
    import torch
    try:
        p = torch.cuda.get_device_properties(idx)
        return max(0, int(open("/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)).read().strip()))
    except Exception:
        return 0


# ---------------------------------------------------------------------------------------
# BASELINE.json configs[3]: 1 M envs sharded over the ranks, --initial-force=55, episode-length deciles
# ---------------------------------------------------------------------------------------
def run_configs3(args, rank, world, local):
    """This rank's shard of configs[3]: envs [rank*B/G, (rank+1)*B/G) of B = 1 048 576 continuous envs with
    --initial-force=55 play ONE episode each under the in-kernel random agent (bcp_rollout_random: actions drawn from
    Philox keyed by GLOBAL env id, so the episodes do not depend on the sharding).  Returns the shard's histogram and
    counters; the caller merges them in the bench's single all_reduce."""
    import torch
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    from cartpoleplusplus_b200.parallel import shard
    total = args.configs3_envs
    off, n = shard(total, rank, world)
    opts = default_opts(action_force=100.0, initial_force=55.0, auto_reset="pipelined")
    env = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=n, device=local, dtype=args.dtype, seed=args.seed, env_id_offset=off)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    header = {"episodes": 0}
    calls = 0
    while header["episodes"] < n and calls < 64:   # reset (65 calls of R*S substeps) + at most 200 steps per episode
        env.rollout_random(64, max_episodes=1)
        calls += 1
        hist, header, rsum = env.stats()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if header["episodes"] != n:
        raise RuntimeError(f"configs[3]: {header['episodes']} of {n} episodes finished")
    env.close()
    return {"hist": hist, "episodes": header["episodes"], "env_steps": header["env_steps"], "reward_sum": rsum, "ms": ms, "envs_total": total}


# ---------------------------------------------------------------------------------------
# the measured arm
# ---------------------------------------------------------------------------------------
def run_ours(args):
    import ctypes
    import torch
    import torch.distributed as dist
    from cartpoleplusplus_b200 import _lib
    from cartpoleplusplus_b200.env import BatchedBulletCartpole

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    placement = pin_near_gpu(local, world)  # before any pinned buffer is allocated: first touch puts it on the GPU's node
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    n = args.envs
    opts = workload_opts(args, auto_reset="pipelined")
    env = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=n, device=local, dtype=args.dtype,
                                seed=args.seed, env_id_offset=rank * n)
    K, W = args.steps, args.warmup
    RS = env.cfg.action_repeats * env.cfg.steps_per_repeat
    gen = torch.Generator(device=dev)
    gen.manual_seed(args.seed + rank)
    pool = torch.rand((min(W + K, 64), n, 2), device=dev, generator=gen) * 2 - 1  # inputs resident in HBM (cycled)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > L2 (126 MB)
    env.reset(copy=False)
    # untimed: roll the batch to its stationary mix of episode phases (random agent, in-kernel actions)
    env.rollout_random(args.mix_calls)
    torch.cuda.synchronize()

    def one_step(t, ev=None):
        flush.zero_()  # untimed L2 flush between steps
        if ev:
            ev[0].record()
        env.step(pool[t % len(pool)], copy=False)
        if ev:
            ev[1].record()

    for t in range(W):
        one_step(t)
    torch.cuda.synchronize()
    launches0 = env.launch_count
    try:
        uuid = str(torch.cuda.get_device_properties(local).uuid)
        uuid = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
    except Exception:
        uuid = None
    # runs through both timed regions (device-timed steps, then e2e); 2 ms between samples, 10 ms when several ranks share
    # the host's few cores with their own agent loops
    sampler = ClockSampler(local, uuid, interval=0.002 if world == 1 else 0.01)
    sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    # The batch's episode phases move in waves (every env was reset together; the mix calls decorrelate them only
    # partly), so a short contiguous window reads high or low depending on where it falls (round 1: 1.38e8 over 20
    # steps, 1.20e8 over 200).  The K timed steps are therefore spread over a fixed span of >= 200 env.steps: between
    # two timed steps the batch advances `stride - 1` untimed steps (same call, same inputs, not timed, not counted).
    stride = max(1, -(-args.span // K))
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(K)]
    live_acc = torch.zeros((), dtype=torch.float64, device=dev)   # env.steps taken by the TIMED calls (reward 1), summed on the device
    t_all = W
    env.clear_stats()
    for t in range(K):
        for _ in range(stride - 1):
            one_step(t_all)
            t_all += 1
        one_step(t_all, evs[t])
        live_acc += env.reward.sum()
        t_all += 1
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    calls = K * stride
    hist, header, _ = env.stats()   # episode statistics of the whole span
    live_steps = int(round(float(live_acc.item())))
    total_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    launches = (env.launch_count - launches0) // stride

    # ---- per-kernel launch times, untimed extra calls: every kernel bracketed by CUDA events on its launch stream.  While
    # profiling, the handle's cohorts run one after the other (a kernel's time is then its own), so these calls are slower
    # than the timed ones, where the cohorts overlap: the kernels' SHARES are what the roofline block uses ----
    prof_calls = min(20, K)
    env.profile_begin()
    for t in range(prof_calls):
        one_step(t_all + t)
    torch.cuda.synchronize()
    prof = env.profile_end()
    prof_ms = prof["ms_prep"] + prof["ms_solve"]

    # ---- e2e: same metric through the host-buffer API (pinned host buffers; the H2D copy of every step's actions and the D2H
    # copy of its observations, rewards and done flags are inside the timed region, and the host reads every reward).
    # The batch is driven the way a throughput-oriented host loop drives it - as two half batches, double buffered
    # (step_host_begin / step_host_end): while one half's outputs cross PCIe and are read, the other half steps ----
    # one GPU: four parts (1.33e8 against 1.18e8 with two); several ranks on one host: the copies share the host's PCIe and
    # the loop is bytes-bound - fewer, larger copies win (N = 2: 2.40e8 with two parts, 1.62e8 with four)
    parts = args.e2e_parts if args.e2e_parts > 0 else (4 if world == 1 else 2)
    nh = [n // parts] * (parts - 1) + [n - (n // parts) * (parts - 1)]
    halves, off = [], 0
    for i in range(parts):
        he = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=nh[i], device=local, dtype=args.dtype, seed=args.seed,
                                   env_id_offset=rank * n + off)
        off += nh[i]
        he.reset(copy=False)
        he.rollout_random(args.mix_calls)
        halves.append(he)
    host_actions = [[torch.empty((nh[i], 2), dtype=torch.float32).pin_memory().uniform_(-1, 1) for _ in range(2)] for i in range(parts)]
    Ke = max(3, K)
    for t in range(2 * max(1, (max(W, 3) + 1) // 2)):  # W untimed steps (an even number: both buffer / graph parities) through the same API
        for i in range(parts):
            halves[i].step_host(host_actions[i][t % 2])
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e2e_live = 0
    t0 = time.perf_counter()
    # every part is always either in flight or being read: begin(0..parts-2) up front, then per part: begin the one that was
    # read longest ago, wait for the oldest in flight, read its rewards
    for i in range(parts - 1):
        halves[i].step_host_begin(host_actions[i][0])
    for t in range(Ke):
        for i in range(parts):
            j = (i + parts - 1) % parts          # the part whose previous results were consumed longest ago
            tj = t if j >= i else t + 1          # its step index
            if tj < Ke:
                halves[j].step_host_begin(host_actions[j][tj % 2])
            _, rew, _, _ = halves[i].step_host_end()
            e2e_live += int(rew.sum().item())
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    sampler.stop_flag = True
    h2d_bytes = sum(h.h2d_bytes_per_step for h in halves)
    d2h_bytes = sum(h.d2h_bytes_per_step for h in halves)

    # ---- configs[3] (N >= 2 only): 1 048 576 envs sharded over the ranks, --initial-force=55, a fixed episode budget (one
    # episode per env) through bcp_rollout_random; its statistics ride in the same single collective below ----
    c3 = None
    if world > 1 and not args.no_configs3:
        c3 = run_configs3(args, rank, world, local)

    # ---- THE one collective of the path: a single all_reduce(SUM) of float64 - per-rank time slots (max taken locally),
    # counts, the episode-length histograms and the reward sum (all exact below 2^53) ----
    nb = env.max_episode_len + 1
    slots = np.zeros(3 * world)
    slots[3 * rank:3 * rank + 3] = [total_ms, e2e_s * 1e3, c3["ms"] if c3 else 0.0]
    counts = [live_steps, e2e_live, header["episodes"]] + ([c3["episodes"], c3["env_steps"], c3["reward_sum"]] if c3 else [0, 0, 0.0])
    packed = torch.from_numpy(np.concatenate([slots, counts, hist.astype(np.float64), c3["hist"].astype(np.float64) if c3 else np.zeros(nb)])).to(dev)
    if world > 1:
        dist.all_reduce(packed, op=dist.ReduceOp.SUM)
    packed = packed.cpu().numpy()
    slots = packed[:3 * world].reshape(world, 3)
    total_ms, e2e_ms, c3_ms = (float(slots[:, k].max()) for k in range(3))
    live_all, e2e_live_all, episodes_all, c3_eps, c3_steps = (int(round(v)) for v in packed[3 * world:3 * world + 5])
    c3_rsum = float(packed[3 * world + 5])
    hist_all = np.rint(packed[3 * world + 6:3 * world + 6 + nb]).astype(np.int64)
    c3_hist = np.rint(packed[3 * world + 6 + nb:]).astype(np.int64)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- census: what every env's solve did in each of the R*S substeps of one more (untimed) call, read back from the
    # pipeline's own solve records on the GPU (bcp_solve_trace) ----
    split = prof["n_solve"] > 0
    model_name, model = load_flops_model(env.cfg)
    lib = _lib.load()
    peak32, peak64 = ctypes.c_double(), ctypes.c_double()
    lib.bcp_bench_fma_peak(local, 0, 4096, ctypes.byref(peak32))
    lib.bcp_bench_fma_peak(local, 1, 1024, ctypes.byref(peak64))
    pipe_peak = peak64.value if args.dtype == "fp64" else peak32.value
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = {}
    try:  # dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed ncu --set full captures
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        pass
    kernels, census = {}, None
    if split:
        env.solve_trace(True)
        one_step(t_all)
        torch.cuda.synchronize()
        tr = env.solve_trace().astype(np.int64)                      # (RS, n, 4): gc, cp, gp points, iterations
        busy = ((env.reward > 0) | (env.reason == C.REASON_RESETTING) | (env.reason == C.REASON_FRESH)).cpu().numpy()
        env.solve_trace(False)
        cnt = tr[:, busy, :].reshape(-1, 4)
        fixed, pgs = substep_flops(model, cnt)
        busy_frac = float(busy.mean())
        flops_call = float(fixed.sum() + pgs.sum())                   # one env.step of the whole batch (this rank)
        census = {"source": "bcp_solve_trace: every env's solve record in each substep of one more call after the timed span",
                  "busy_fraction": busy_frac, "mean_solver_iterations": float(cnt[:, 3].mean()),
                  "mean_contact_points": {"ground_cart": float(cnt[:, 0].mean()), "cart_pole": float(cnt[:, 1].mean()), "ground_pole": float(cnt[:, 2].mean())},
                  "flops_per_busy_substep": {"fixed": float(fixed.mean()), "pgs": float(pgs.mean())},
                  "flops_model": f"profiles/flops_model.json[{model_name}] (counted on the parity substep, FMA = 2)"}
        kernels["k_prep"] = {"ms_per_step": prof["ms_prep"] / prof_calls, "launch_ms": prof["ms_prep"] / max(1, prof["n_prep"]),
                             "flops_per_step": float(fixed.sum()), "what": "narrowphase, manifolds, dynamics, row setup, integrate"}
        kernels["k_solve"] = {"ms_per_step": prof["ms_solve"] / prof_calls, "launch_ms": prof["ms_solve"] / max(1, prof["n_solve"]),
                              "flops_per_step": float(pgs.sum()), "what": "projected Gauss-Seidel (k_solve_lam + k_solve_tail + side-stream k_solve)"}
    else:  # the one-launch kernels keep no per-env solve record: the model's workload mean stands in
        flops_call = model["mean_flops_per_substep"] * RS * n
        kernels["k_step"] = {"ms_per_step": prof["ms_prep"] / prof_calls, "launch_ms": prof["ms_prep"] / max(1, prof["n_prep"]),
                             "flops_per_step": flops_call, "what": "whole env.step in one launch"}
    step_ms = total_ms / K
    for name, kd in kernels.items():
        kd["achieved_tflops"] = kd["flops_per_step"] / (kd["ms_per_step"] * 1e-3) / 1e12
        kd["frac_of_fp32_peak"] = kd["achieved_tflops"] / pipe_peak
        kd["share_of_step"] = kd["ms_per_step"] * prof_calls / prof_ms   # of the kernels' own (serialised) time
        t = traffic.get(name if args.dtype == "fp32" else name + "_f64")
        kd["dram_bytes_per_launch_ncu"] = t.get("bytes_per_launch") if isinstance(t, dict) and int(t.get("num_envs", n)) == n else None
    dom_name = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
    dom = kernels[dom_name]
    alg_bytes = algorithmic_bytes_per_env_step(env.cfg) * n * busy_frac if split else algorithmic_bytes_per_env_step(env.cfg) * n
    t_step = traffic.get("env_step") if args.dtype == "fp32" else None
    dram_step = t_step.get("bytes_per_step") if isinstance(t_step, dict) and int(t_step.get("num_envs", n)) == n else None
    whole_tf = flops_call / (step_ms * 1e-3) / 1e12
    value = live_all / (total_ms * 1e-3)
    out = {
        "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.dtype == "fp32" else "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "envs_per_gpu": n, "action_repeats": env.cfg.action_repeats,
                   "steps_per_repeat": env.cfg.steps_per_repeat, "initial_force": args.initial_force,
                   "cart_friction": env.cfg.cart.friction,
                   "reset": "pybullet's (pending force and contact caches carried over; 100 settle + 30 push substeps)" if env.cfg.carry_over_reset
                            else ("--no-carry-over-reset" + (" --reset-settle-cache (settle phase replayed: 30 push substeps per reset)" if env.cfg.reset_settle_cache else "")),
                   "termination": "angle/position/200-step rule on",
                   "auto_reset": "pipelined: a finished env runs its 130-substep reset R*S substeps per call; only env.steps "
                                 "actually taken are counted",
                   "phase_mix": f"stationary ({args.mix_calls} untimed random-agent calls before timing)",
                   "window": f"the {K} timed steps are every {stride}th env.step of a span of {calls} (episode phases move in waves; "
                             "the steps in between run untimed and uncounted)" if stride > 1 else f"{K} consecutive env.steps",
                   "l2": "flushed between steps (256 MiB memset, untimed)", "parallelism": f"env-sharded x{world}", "pipeline": env.solver_kind,
                   "seed": args.seed, "host_placement": placement},
        "live_fraction": live_all / (world * n * K),
        "slot_steps_per_s": world * n * K / (total_ms * 1e-3),
        "substeps_per_s": world * n * K * RS / (total_ms * 1e-3),
        "episodes_finished_in_span": episodes_all,
        "mean_episode_len": float((hist_all * np.arange(len(hist_all))).sum() / max(1, int(hist_all.sum()))),
        "kernel_ms_per_step": {k: v["ms_per_step"] for k, v in kernels.items()},
        "kernel_ms_note": f"serialised (cohorts one after the other, {prof_calls} untimed profiled calls); the timed steps overlap the cohorts",
        "e2e": {"value": e2e_live_all / (e2e_ms * 1e-3), "unit": "env-steps/s",
                "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                "how": f"step_host_begin / step_host_end on {parts} parts of the batch, multi-buffered; every step's H2D and D2H inside the timed region"},
        "gpu_launches": int(launches),
        "clocks": sampler.summary(),
        # BASELINE.json's metric is "% FP32 peak": the path is bound by the FP32 CUDA-core pipe (no contraction, arithmetic
        # intensity far above the ridge), so the roofline is flops - counted, algorithmic - over the measured FMA peak.
        # `achieved` / `frac`: counted flops of the whole env.step over the TIMED steps' device time - the cohorts' kernels overlap in
        # the timed region, so a single kernel's time there is not separable; `dominant_kernel` holds that kernel's own figures from
        # the serialised profile pass (launches of one cohort each, run alone)
        "roofline": {"bound": "fp32", "kernel": dom_name, "achieved": whole_tf, "peak": pipe_peak, "unit": "TFLOP/s",
                     "frac": whole_tf / pipe_peak, "whole_step_frac": whole_tf / pipe_peak, "whole_step_achieved": whole_tf,
                     "dominant_kernel": {"name": dom_name, "achieved": dom["achieved_tflops"], "frac": dom["frac_of_fp32_peak"],
                                         "launch_ms": dom["launch_ms"], "share_of_serialised_kernel_time": dom["share_of_step"]},
                     "flops_per_env_step_call": flops_call / n,
                     "peak_source": "bcp_bench_fma_peak (dependent-free FMA chains on every SM, measured in this run; "
                                    "MEASURED_PEAKS.json has no CUDA-core entry)",
                     # DRAM bytes of ALL launches of one env.step (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum,
                     # profiles/traffic.json) beside SURVEY 8d's algorithmic bytes (state once in, once out, action, outputs)
                     "traffic": dram_step, "algorithmic_bytes": alg_bytes,
                     "traffic_ratio": (dram_step / alg_bytes) if dram_step else None,
                     "hbm": {"achieved_gbs": alg_bytes / (step_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                             "frac": alg_bytes / (step_ms * 1e-3) / 1e9 / hbm_peak,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s"}},
        "roofline_kernels": kernels,
        "census": census,
    }
    if c3:
        from cartpoleplusplus_b200 import protocol
        import hashlib
        out["configs3"] = {
            "workload": "configs[3]: 1 048 576 envs continuous sharded over the ranks, --initial-force=55, one episode per env "
                        "(bcp_rollout_random, in-kernel Philox actions keyed by global env id), one allreduce",
            "envs_total": c3["envs_total"], "envs_per_gpu": c3["envs_total"] // world, "episodes": c3_eps, "env_steps": c3_steps,
            "value": c3_steps / (c3_ms * 1e-3), "unit": "env-steps/s", "ms": c3_ms, "reward_sum": c3_rsum,
            "deciles": [float(x) for x in protocol.deciles_from_histogram(c3_hist)],
            "histogram_sha256": hashlib.sha256(c3_hist.astype("<i8").tobytes()).hexdigest(),
            "note": "trajectories are keyed by global env id: deciles and hash must be identical for every N"}
    if not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline(args)
    if world == 1 and not args.no_secondary and args.cart_friction is None:
        for name in SECONDARY:
            out[name] = secondary_workload(args, name)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    """--impl reference: the reference's CPU path.  pybullet is unavailable on this box, so this
    times the oracle port (the one other place bench.py may execute oracle/), all host threads.
    Each bench step is a bounded sample: every env of a host-sized batch takes `per` env.steps."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    K, W = args.steps, args.warmup
    n = 256 * cores
    probe_steps, _, probe_dt = cpu_rollout(args, n, 8, 40)
    rate = probe_steps / probe_dt
    per = int(max(2, min(200, 90.0 * rate / n / max(1, K + W))))  # whole run ~90 s of CPU work
    tot_steps, tot_s, tot_resets = 0, 0.0, 0
    for t in range(W + K):
        steps, resets, dt = cpu_rollout(args, n, per, 40)
        if t >= W:
            tot_steps += steps
            tot_s += dt
            tot_resets += resets
    v = tot_steps / tot_s
    print(json.dumps({
        "impl": "reference", "metric": "env-steps/sec", "value": v, "unit": "env-steps/s",
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": K, "warmup": W,
        "ms_per_step": 1e3 * tot_s / max(1, K), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample_envs": n, "env_steps_per_bench_step": per},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                         "sample": f"{n} envs x {per} env.steps per bench step (+{tot_resets} resets of 130 substeps in all), "
                                   "stationary phase mix; C oracle (fp64, OpenMP); pybullet: unavailable on this box"},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def ensure_built():
    """The in-tree libraries are build products (git-ignored): compile them when a checkout arrives without
    them.  One rank compiles, the others wait for the files."""
    from cartpoleplusplus_b200 import _lib
    from oracle.oracle import oracle_lib_path
    need = [_lib.LIB_PATH, oracle_lib_path()]
    if all(os.path.exists(p) for p in need):
        return
    if int(os.environ.get("LOCAL_RANK", "0")) == 0:
        import __graft_entry__ as g
        g.build()
        return
    t0 = time.time()
    while not all(os.path.exists(p) for p in need):
        if time.time() - t0 > 1800:
            raise RuntimeError("timed out waiting for rank 0 to build libbcp.so")
        time.sleep(2.0)
    time.sleep(2.0)  # the linker's output is complete once the file is closed


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs", type=int, default=65536, help="envs per GPU")
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--initial-force", type=float, default=200.0)
    ap.add_argument("--repeats", type=int, default=2)
    ap.add_argument("--steps-per-repeat", type=int, default=1)
    ap.add_argument("--cart-friction", type=float, default=None, help="override cart.urdf's lateral_friction (0.0)")
    ap.add_argument("--mix-calls", type=int, default=400, help="untimed random-agent calls before timing")
    ap.add_argument("--span", type=int, default=200, help="env.steps the timed steps are spread over (>= --steps: every span/steps-th step is timed)")
    ap.add_argument("--no-configs3", action="store_true", help="N >= 2: skip the configs[3] run (1 M envs sharded, deciles)")
    ap.add_argument("--configs3-envs", type=int, default=1048576)
    ap.add_argument("--e2e-parts", type=int, default=0,
                    help="the e2e host loop drives the batch as this many parts, multi-buffered (0: 4 on one GPU, 2 when several ranks share the host's PCIe)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the labelled workloads reported beside the headline (upstream-style friction; fast reset)")
    ap.add_argument("--no-carry-over-reset", action="store_true", help="reset() clears the pending force and the contact caches (deviation from pybullet)")
    ap.add_argument("--reset-settle-cache", action="store_true", help="with --no-carry-over-reset: replay the settle phase from a per-handle cache")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work of the cpu_baseline sample")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    ensure_built()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
