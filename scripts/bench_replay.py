"""Device replay memory under the bench workload: BASELINE.json configs[2] (65 536 envs, continuous,
pipelined auto-reset) feeding DeviceReplayMemory every env.step, then batch gathers.

Prints one JSON line: ms per env.step with and without the replay feed, the feed's own kernel
time (CUDA events around observe), its algorithmic bytes and GB/s against the HBM peak of
MEASURED_PEAKS.json, and the same for batch().  Inputs are larger than L2 between calls (the
staging area and the ring are touched once per call; the 256 MiB flush of bench.py is applied
between timed calls as well).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from cartpoleplusplus_b200 import default_opts  # noqa: E402
from cartpoleplusplus_b200.env import BatchedBulletCartpole  # noqa: E402
from cartpoleplusplus_b200.replay import DeviceReplayMemory  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=65536)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warm", type=int, default=420)
    ap.add_argument("--batch", type=int, default=65536)
    ap.add_argument("--buffer", type=int, default=0, help="ring rows (default: envs * max_episode_len rounded up to 2^k)")
    a = ap.parse_args()
    B, L = a.envs, 200
    N = a.buffer or 1 << (B * L - 1).bit_length()
    opts = default_opts(action_force=100.0, auto_reset="pipelined")
    env = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=B, device=0, dtype="fp32", seed=0)
    W, A = env.repeats * 14, 2
    rm = DeviceReplayMemory(N, (env.repeats, 2, 7), A, num_envs=B, max_episode_len=L, device=0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    g = torch.Generator(device="cuda").manual_seed(0)
    acts = [torch.rand((B, 2), device="cuda", generator=g) * 2 - 1 for _ in range(8)]
    obs = env.reset(copy=False)
    rm.begin(obs)
    for t in range(a.warm):
        o, r, d, info = env.step(acts[t % 8], copy=False)
        rm.observe(acts[t % 8], o, r, d, info["done_reason"])
    i0 = rm.info()
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    # events are recorded for every call and read after one final sync: the host runs ahead of the
    # device, so the deltas are device time without launch gaps
    marks = []
    for t in range(a.steps):
        flush.zero_()
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        o, r, d, info = env.step(acts[t % 8], copy=False)
        e1.record()
        rm.observe(acts[t % 8], o, r, d, info["done_reason"])
        e2.record()
        marks.append((e0, e1, e2))
    torch.cuda.synchronize()
    t_step = sum(m[0].elapsed_time(m[1]) for m in marks)
    t_obs = sum(m[1].elapsed_time(m[2]) for m in marks)
    obs_ms = sorted(m[1].elapsed_time(m[2]) for m in marks)
    # marginal cost in a free-running loop: K x (flush, step) against K x (flush, step, observe), one sync each
    def loop(feed):
        e0, e1 = ev(), ev()
        e0.record()
        for t in range(a.steps):
            flush.zero_()
            o, r, d, info = env.step(acts[t % 8], copy=False)
            if feed:
                rm.observe(acts[t % 8], o, r, d, info["done_reason"])
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / a.steps
    ms_loop_feed = loop(True)
    ms_loop_plain = loop(False)
    i1 = rm.info()
    trans = (i1[">add"] - i0[">add"]) / 2  # two fed loops of a.steps calls
    eps = (i1[">add_episode"] - i0[">add_episode"]) / 2
    # algorithmic bytes: observe per env (obs f32 in, f16 state + action + reward out, flags), commit per transition
    # (staged state/action/reward in, pool state + ring row out, free-slot entries), per episode one more state
    obs_bytes = B * (4 * W + 4 * A + 4 + 2 + 8 + 2 * W + 4 * A + 4 + 4)
    commit_bytes = trans * (2 * W + 4 * A + 4 + 8 + 2 * W + 4 * (A + 4) + 8) / a.steps + eps * (4 * W + 24) / a.steps
    # one commit of a whole wave: every env finishes a max_episode_len-step episode in the same call
    dflag = torch.zeros(B, dtype=torch.uint8, device="cuda")  # one buffer: the call's CUDA graph is keyed by the pointers
    rew1 = torch.ones(B, dtype=torch.float32, device="cuda")
    rm.begin(o)
    for t in range(L - 1):
        rm.observe(acts[0], o, rew1, dflag)
    dflag.fill_(1)
    flush.zero_()
    w0, w1 = ev(), ev()
    w0.record()
    rm.observe(acts[0], o, rew1, dflag)
    w1.record()
    torch.cuda.synchronize()
    ms_wave = w0.elapsed_time(w1)
    iw = rm.info()
    wave_bytes = B * (4 * W + 4 * A + 4 + 2 + 8 + 2 * W + 4 * A + 4 + 4) + \
        iw["last_commit_transitions"] * (2 * W + 4 * A + 4 + 8 + 2 * W + 4 * (A + 4) + 8 + 2 * (4 + 4 + 4)) + iw["last_commit_episodes"] * (4 * W + 24)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6539.5))
    ms_obs = t_obs / a.steps
    # batch gather
    n = a.batch
    for _ in range(3):
        rm.batch(n)
    reps = 20
    bm = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = ev(), ev()
        e0.record()
        rm.batch(n)
        e1.record()
        bm.append((e0, e1))
    torch.cuda.synchronize()
    tb = sum(m[0].elapsed_time(m[1]) for m in bm)
    ms_batch = tb / reps
    batch_bytes = n * (8 + 8 + 2 * (2 * W) * 2 + 4 * A * 2 + 16 + 8)
    out = {
        "workload": f"{B} envs continuous pipelined auto-reset -> DeviceReplayMemory({N} rows, state (R,2,7) f16, action_dim 2)",
        "ms_env_step": t_step / a.steps, "ms_replay_observe": ms_obs, "replay_share_of_step": ms_obs / (t_step / a.steps + ms_obs),
        "ms_loop_step_plus_observe": ms_loop_feed, "ms_loop_step_only": ms_loop_plain,
        "ms_observe_marginal": ms_loop_feed - ms_loop_plain,
        "transitions_per_call": trans / a.steps, "episodes_per_call": eps / a.steps,
        "replay_kernels_per_call": 7,
        "observe_commit_algorithmic_bytes_per_call": obs_bytes + commit_bytes,
        "observe_commit_gbs": (obs_bytes + commit_bytes) / (ms_obs * 1e-3) / 1e9,
        "wave_commit": {"transitions": iw["last_commit_transitions"], "episodes": iw["last_commit_episodes"], "ms": ms_wave,
                        "algorithmic_bytes": wave_bytes, "gbs": wave_bytes / (ms_wave * 1e-3) / 1e9,
                        "frac_of_hbm_peak": wave_bytes / (ms_wave * 1e-3) / 1e9 / peak,
                        "note": "every env finishes a 200-step episode in one call; the ring is full, so every row also evicts"},
        "ms_batch": ms_batch, "batch_rows": n, "batch_algorithmic_bytes": batch_bytes,
        "batch_gbs": batch_bytes / (ms_batch * 1e-3) / 1e9, "batch_frac_of_hbm_peak": batch_bytes / (ms_batch * 1e-3) / 1e9 / peak,
        "hbm_peak_gbs": peak, "size": i1["size"], "full": i1["full"], "free_slots": i1["free_slots"],
        "launches": rm.launch_count,
        "ms_replay_observe_min_median_max": [obs_ms[0], obs_ms[len(obs_ms) // 2], obs_ms[-1]],
    }
    print(json.dumps(out))


if __name__ == "__main__":
    main()
