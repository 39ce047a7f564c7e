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
                        steps_per_repeat=args.steps_per_repeat, cart_friction=args.cart_friction, **extra)


# ---------------------------------------------------------------------------------------
# flop model of the solve (FMA = 2 flops, useful work only: masked lanes,
# padding and recomputation do not count).  One row update / one cone update of two rows.
# ---------------------------------------------------------------------------------------
F_ROW1, F_ROW2 = 36.0, 61.0      # normal row touching one / two dynamic bodies
F_CONE1, F_CONE2 = 83.0, 133.0   # the two friction rows of one contact, cone-clipped


def solve_flops(info, cfg):
    """Useful flops of the PGS solve of one substep for the envs censused in `info`
    (oracle substep info rows: iterations, rows, friction rows, contacts gc/cp/gp)."""
    it = info[:, 0].astype(np.float64)
    gc, cp, gp = (info[:, 3 + k].astype(np.float64) for k in range(3))
    mu = {"gc": cfg.ground.friction * cfg.cart.friction, "cp": cfg.cart.friction * cfg.pole.friction,
          "gp": cfg.ground.friction * cfg.pole.friction}
    per_iter = gc * F_ROW1 + gp * F_ROW1 + cp * F_ROW2
    if cfg.friction_dirs == 2:
        per_iter += (gc * F_CONE1 if mu["gc"] else 0) + (gp * F_CONE1 if mu["gp"] else 0) + (cp * F_CONE2 if mu["cp"] else 0)
    return it * per_iter


def _row_words(info, cfg):
    gc, cp, gp = (info[:, 3 + k].astype(np.float64) for k in range(3))
    rpc = lambda a, b: 3 if (a * b) and cfg.friction_dirs == 2 else 1  # noqa: E731
    return (gc * rpc(cfg.ground.friction, cfg.cart.friction) * 8 + gp * rpc(cfg.ground.friction, cfg.pole.friction) * 8
            + cp * rpc(cfg.cart.friction, cfg.pole.friction) * 12)


def solve_bytes(info, cfg):
    """Algorithmic bytes k_solve moves per env (fp32): the record header (32 words) and its live
    rows (8 / 12 words per one- / two-body row) in; iteration count, velocity deltas and normal
    impulses (25 words) out."""
    return 4.0 * (32 + _row_words(info, cfg) + 25)


def prep_bytes(info, cfg, RS):
    """Algorithmic bytes k_prep moves per env and launch, averaged over the RS + 1 launches of a
    call: the env record (164 floats + 8 ints) in and out every launch; RS of them write a solve
    record (header + live rows) and RS read one back (header) and the action; the last writes the
    (R,2,7) observation, reward, done and reason."""
    state = 2.0 * (164 * 4 + 8 * 4)
    setup = 4.0 * (32 + _row_words(info, cfg))
    finish = 4.0 * 32 + 8.0
    close = cfg.action_repeats * 14 * 4 + 6.0
    return state + (RS * (setup + finish) + close) / (RS + 1)


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


def cpu_baseline(args, budget_s=12.0):
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


class ClockSampler(threading.Thread):
    """SM clock, power and throttle reasons sampled DURING the timed regions.

    NVML in-process (about a millisecond per sample, so that even a 15 ms timed region gets several);
    the nvidia-smi query of the profiling recipe is the fallback when NVML cannot be loaded."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index, uuid=None):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
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
            time.sleep(0.002 if self.nvml is not None else 0.05)

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
    pool = torch.rand((W + K, n, 2), device=dev, generator=gen) * 2 - 1  # inputs resident in HBM
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > L2 (126 MB)
    env.reset(copy=False)
    # untimed: roll the batch to its stationary mix of episode phases (random agent, in-kernel actions)
    env.rollout_random(args.mix_calls)
    torch.cuda.synchronize()

    def one_step(t, ev=None):
        flush.zero_()  # untimed L2 flush between steps
        if ev:
            ev[0].record()
        env.step(pool[t], copy=False)
        if ev:
            ev[1].record()

    for t in range(W):
        one_step(t)
    torch.cuda.synchronize()
    env.clear_stats()
    launches0 = env.launch_count
    try:
        uuid = str(torch.cuda.get_device_properties(local).uuid)
        uuid = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
    except Exception:
        uuid = None
    sampler = ClockSampler(local, uuid)  # runs through both timed regions (device-timed steps, then e2e)
    sampler.start()
    env.profile_begin()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(K)]
    for t in range(K):
        one_step(W + t, evs[t])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    prof = env.profile_end()
    total_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    launches = env.launch_count - launches0
    hist, header, _ = env.stats()
    live_steps = header["env_steps"]

    # ---- e2e: same metric through step_host (pinned host buffers, H2D + D2H inside) ----
    host_actions = [torch.empty((n, 2), dtype=torch.float32).pin_memory() for _ in range(2)]
    for h in host_actions:
        h.uniform_(-1, 1)
    Ke = max(3, K)
    for t in range(2):
        env.step_host(host_actions[t % 2])
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e2e_live = 0
    t0 = time.perf_counter()
    for t in range(Ke):
        _, rew, _, _ = env.step_host(host_actions[t % 2])
        e2e_live += int(rew.sum().item())
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    sampler.stop_flag = True

    # ---- max over ranks of the times, sums of the counts; the one collective: statistics allreduce ----
    tt = torch.tensor([total_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    cc = torch.tensor([live_steps, e2e_live, header["episodes"]], dtype=torch.int64, device=dev)
    hist_t = torch.from_numpy(hist).to(dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(cc, op=dist.ReduceOp.SUM)
        dist.all_reduce(hist_t, op=dist.ReduceOp.SUM)
    total_ms, e2e_ms = (float(v) for v in tt.tolist())
    live_all, e2e_live_all, episodes_all = (int(v) for v in cc.tolist())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- census of the batch: what the solves of the last substep did (contact points per pair,
    # PGS iterations), read back from the pipeline's own solve records on the GPU ----
    split = prof["n_solve"] > 0
    if split:
        li = env.last_solve_info().astype(np.int64)                       # (n, 4): gc, cp, gp points, iterations
        busy = ((env.reward > 0) | (env.reason == C.REASON_RESETTING)).cpu().numpy()   # envs that ran that substep
        info = np.zeros((n, 6), dtype=np.int64)
        info[:, 0] = li[:, 3]
        info[:, 3:6] = li[:, 0:3]
        info[~busy] = 0
        iters = float(info[busy, 0].mean()) if busy.any() else 0.0
        rows_gc, rows_cp, rows_gp = (float(info[busy, 3 + k].mean()) for k in range(3))
        flops_env = float(solve_flops(info, env.cfg).mean())
        bytes_env = float(solve_bytes(info, env.cfg).mean())
    else:  # the one-launch kernels keep no per-env solve record: no census, no flop view
        info, busy, iters, rows_gc, rows_cp, rows_gp, flops_env, bytes_env = None, None, None, None, None, None, None, None
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
    # per kernel: algorithmic bytes of one launch / its average launch time in the timed region
    kernels = {}
    if split:
        pb = float(prep_bytes(info, env.cfg, RS).mean())
        kernels["k_prep"] = {"ms_per_step": prof["ms_prep"] / K, "launch_ms": prof["ms_prep"] / max(1, prof["n_prep"]),
                             "bytes_per_env_launch": pb, "bytes_per_launch": pb * n}
        kernels["k_solve"] = {"ms_per_step": prof["ms_solve"] / K, "launch_ms": prof["ms_solve"] / max(1, prof["n_solve"]),
                              "bytes_per_env_launch": bytes_env, "bytes_per_launch": bytes_env * n}
    else:
        kernels["k_step"] = {"ms_per_step": prof["ms_prep"] / K, "launch_ms": prof["ms_prep"] / max(1, prof["n_prep"]),
                             "bytes_per_env_launch": 2.0 * 176 * 8 + 8, "bytes_per_launch": (2.0 * 176 * 8 + 8) * n}
    for name, kd in kernels.items():
        kd["achieved_gbs"] = kd["bytes_per_launch"] / (kd["launch_ms"] * 1e-3) / 1e9
        kd["frac_of_hbm_peak"] = kd["achieved_gbs"] / hbm_peak
        kd["share_of_step"] = kd["ms_per_step"] * K / total_ms
        t = traffic.get(name if args.dtype == "fp32" else name + "_f64")
        kd["traffic"] = t.get("bytes_per_launch") if isinstance(t, dict) and int(t.get("num_envs", n)) == n else None
    dom_name = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
    dom = kernels[dom_name]
    solve_name = "k_solve" if split else "k_step"
    solve_ms = kernels[solve_name]["launch_ms"]
    achieved_tf = flops_env * n / (solve_ms * 1e-3) / 1e12 if split else None
    value = live_all / (total_ms * 1e-3)
    out = {
        "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.dtype == "fp32" else "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "envs_per_gpu": n, "action_repeats": env.cfg.action_repeats,
                   "steps_per_repeat": env.cfg.steps_per_repeat, "initial_force": args.initial_force,
                   "cart_friction": env.cfg.cart.friction,
                   "termination": "angle/position/200-step rule on",
                   "auto_reset": "pipelined: a finished env runs its 130-substep reset R*S substeps per call; only env.steps "
                                 "actually taken are counted",
                   "phase_mix": f"stationary ({args.mix_calls} untimed random-agent calls before timing)",
                   "l2": "flushed between steps (256 MiB memset, untimed)", "parallelism": f"env-sharded x{world}", "pipeline": env.solver_kind,
                   "seed": args.seed},
        "live_fraction": live_all / (world * n * K),
        "slot_steps_per_s": world * n * K / (total_ms * 1e-3),
        "substeps_per_s": world * n * K * RS / (total_ms * 1e-3),
        "episodes_finished": episodes_all,
        "mean_episode_len": float((hist_t.cpu().numpy() * np.arange(len(hist))).sum() / max(1, int(hist_t.sum().item()))),
        "kernel_ms_per_step": {k: v["ms_per_step"] for k, v in kernels.items()},
        "e2e": {"value": e2e_live_all / (e2e_ms * 1e-3), "unit": "env-steps/s",
                "h2d_bytes_per_step": env.h2d_bytes_per_step, "d2h_bytes_per_step": env.d2h_bytes_per_step},
        "gpu_launches": int(launches),
        "clocks": sampler.summary(),
        # the dominant kernel (largest share of the step) against the measured HBM copy bandwidth
        "roofline": {"bound": "hbm", "achieved": dom["achieved_gbs"], "peak": hbm_peak, "unit": "GB/s",
                     "frac": dom["frac_of_hbm_peak"], "traffic": dom["traffic"], "kernel": dom_name,
                     "launch_ms": dom["launch_ms"], "share_of_step": dom["share_of_step"],
                     "algorithmic_bytes_per_env_launch": dom["bytes_per_env_launch"],
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s",
                     "note": "both kernels are issue-latency bound, not bandwidth bound (profiles/); "
                             "see roofline_kernels and roofline_pipe"},
        "roofline_kernels": kernels,
        "roofline_pipe": None if not split else {
            "bound": "fp32-cuda-core", "kernel": solve_name,
            "launch_ms": solve_ms, "share_of_step": kernels[solve_name]["share_of_step"],
            "achieved": achieved_tf, "peak": pipe_peak, "unit": "TFLOP/s", "frac": achieved_tf / pipe_peak,
            "peak_source": "bcp_bench_fma_peak (FMA chains, measured in this run)",
            "flops_per_env_solve": flops_env, "bytes_per_env_solve": bytes_env,
            "mean_solver_iterations": iters, "busy_fraction": float(busy.mean()),
            "mean_rows": {"ground_cart": rows_gc, "cart_pole": rows_cp, "ground_pole": rows_gp},
            "census": "bcp_last_solve_info: every env's solve record after the last timed call"},
    }
    if not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline(args)
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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    ensure_built()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
