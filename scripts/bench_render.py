"""Raw-pixel observation kernel at the bench batch: ms per launch and GB/s of image written against the HBM peak
of MEASURED_PEAKS.json.  One JSON line."""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from cartpoleplusplus_b200.render import PixelRenderer  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=65536)
    ap.add_argument("--cameras", type=int, default=1)
    ap.add_argument("--repeats", type=int, default=2)
    ap.add_argument("--size", type=int, default=50)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    n = a.envs
    r = PixelRenderer(n, a.repeats, a.cameras, a.size, a.size)
    g = torch.Generator(device="cuda").manual_seed(0)
    obs = torch.zeros((n, a.repeats, 2, 7), device="cuda")
    obs[..., 6] = 1
    obs[:, :, 0, 2], obs[:, :, 1, 2] = 0.075, 0.35
    obs[:, :, :, :2] = (torch.rand((n, 1, 1, 2), device="cuda", generator=g) - 0.5) * 0.5
    q = torch.randn((n, a.repeats, 4), device="cuda", generator=g) * torch.tensor([0.2, 0.2, 0.05, 0.0], device="cuda")
    q[..., 3] = 1
    obs[:, :, 1, 3:] = q / q.norm(dim=-1, keepdim=True)
    for _ in range(3):
        r.render(obs)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.reps)]
    for e0, e1 in ev:  # the image (2 GB at the defaults) is far larger than L2: no flush needed between launches
        e0.record()
        r.render(obs)
        e1.record()
    torch.cuda.synchronize()
    ms = sorted(e0.elapsed_time(e1) for e0, e1 in ev)
    med = ms[len(ms) // 2]
    out_bytes = r.pixels.numel() * 2
    peak = 6539.5
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak))
    except Exception:
        pass
    gbs = (out_bytes + obs.numel() * 4) / (med * 1e-3) / 1e9
    print(json.dumps({"workload": f"{n} envs, {a.size}x{a.size}, C={a.cameras}, R={a.repeats}: (B,H,W,3,C,R) float16",
                      "ms_median": med, "ms_min": ms[0], "ms_max": ms[-1], "algorithmic_bytes": out_bytes + obs.numel() * 4,
                      "gbs": gbs, "hbm_peak_gbs": peak, "frac_of_hbm_peak": gbs / peak,
                      "images_per_s": n * a.cameras * a.repeats / (med * 1e-3)}))


if __name__ == "__main__":
    main()
