"""Calibration of the oracle's [BULLET-RECALL] constants against ALL FOUR random-agent rows the reference publishes
(README.md:77-92; upstream's numbers, n = 100 episodes each) - jointly, not one answer at a time.

TEST TOOL (runs oracle/ on the host cores).  For every combination of the recalled engine constants and of the settings the
README rows were (probably) produced with - none of which is derivable from this tree - the fp64 oracle plays the four
protocols and the script reports, per combination, a distance to the README deciles:
    distance = mean over the 4 rows of  mean_k |log(oracle decile_k / README decile_k)|     (k = 0..10)
(log ratio: a factor-2 error counts the same at the 10 % and the 90 % decile).  Output: profiles/oracle_calibration.json,
best combinations first, with the deciles of the leaders and of the current defaults."""
import dataclasses
import itertools
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cartpoleplusplus_b200 import config as C  # noqa: E402
from cartpoleplusplus_b200 import default_opts, make_config, protocol, urdf  # noqa: E402
from oracle import Oracle  # noqa: E402

ROWS = [(0.0, (0,)), (0.0, (0, 1, 2, 3, 4)), (55.0, (0,)), (55.0, (0, 1, 2, 3, 4))]

GRID = {
    "inertia_scaling": [1.0, 3.0],        # <inertia_scaling value="3.0"> honoured or not
    "damping": [0.0, 0.04],               # btMultiBody linear / angular damping
    "erp2": [0.08, 0.2],
    "warmstart": [0.0, 0.1, 0.85],
    "friction": ["2dir-cone", "1dir"],
    "RS": [(2, 5), (3, 4)],               # upstream's defaults (recalled) / README NAF recipe
    "frame": ["LINK", "WORLD"],
    "cart_friction": [1.0, 0.5],          # upstream cart.urdf (recalled) / Bullet's default
    "pole_width": [0.01, 0.02, 0.05],     # fork (pole.urdf:35); the value pole.urdf:15-19's comment says it replaced; upstream's look (cartpole.png)
}


def deciles_for(combo, episodes, seed=0):
    ground, cart, pole = urdf.load_scene(None)
    cart = dataclasses.replace(cart, inertia_scaling=combo["inertia_scaling"], lateral_friction=combo["cart_friction"])
    w = combo["pole_width"]
    pole = dataclasses.replace(pole, inertia_scaling=combo["inertia_scaling"], box_size=(w, w, pole.box_size[2]))
    phys = C.PhysicsParams(linear_damping=C._f32(combo["damping"]), angular_damping=C._f32(combo["damping"]), erp2=combo["erp2"],
                           warmstart_factor=combo["warmstart"], friction_dirs=2 if combo["friction"] == "2dir-cone" else 1,
                           cone_friction=1 if combo["friction"] == "2dir-cone" else 0)
    out = []
    for (force, actions), n_ep in zip(ROWS, episodes):
        per_env = 4 if n_ep >= 64 else 1
        n = max(1, n_ep // per_env)
        opts = default_opts(initial_force=force, action_repeats=combo["RS"][0], steps_per_repeat=combo["RS"][1], theta_per_reset=True,
                            world_frame_force=combo["frame"] == "WORLD")
        cfg = make_config(opts, True, num_envs=n, dtype="fp64", seed=seed, physics=phys, scene=(ground, cart, pole))
        ora = Oracle(cfg, dtype="fp64")
        ora.state[:, C.S_DONE] = 1
        mask = sum(1 << a for a in actions)
        ora.rollout_random(per_env * 201 + 1, action_mask=mask, max_episodes=per_env)
        out.append(protocol.deciles_from_histogram(ora.hist))
    return out


def distance(dec):
    d = []
    for (force, actions), got in zip(ROWS, dec):
        want = np.asarray(protocol.README_DECILES[(force, tuple(actions))], dtype=float)
        d.append(float(np.mean(np.abs(np.log(np.maximum(got, 1.0) / want)))))
    return float(np.mean(d)), d


if __name__ == "__main__":
    quick = "--quick" in sys.argv
    keys = list(GRID)
    combos = [dict(zip(keys, v)) for v in itertools.product(*[GRID[k] for k in keys])]
    if quick:
        combos = combos[::97]
    t0 = time.time()
    results = []
    for i, combo in enumerate(combos):
        dec = deciles_for(combo, episodes=(24, 200, 200, 200))
        dist, per_row = distance(dec)
        results.append({"combo": combo, "distance": dist, "per_row": per_row, "medians": [float(d[5]) for d in dec]})
        if i % 50 == 0:
            print(f"{i}/{len(combos)} {time.time() - t0:.0f}s best so far {min(r['distance'] for r in results):.3f}", flush=True)
    results.sort(key=lambda r: r["distance"])
    # the leaders and the defaults again with README-sized x10 samples
    def detail(combo):
        dec = deciles_for(combo, episodes=(100, 1000, 1000, 1000), seed=1)
        dist, per_row = distance(dec)
        return {"combo": combo, "distance": dist, "per_row": per_row, "deciles": [[round(float(x), 1) for x in d] for d in dec]}
    defaults = {"inertia_scaling": 3.0, "damping": 0.04, "erp2": 0.08, "warmstart": 0.85, "friction": "2dir-cone", "RS": (2, 5),
                "frame": "WORLD", "cart_friction": 1.0, "pole_width": 0.01}
    blob = {
        "readme": {f"initial_force={f} actions={list(a)}": protocol.README_DECILES[(f, tuple(a))] for f, a in ROWS},
        "distance": "mean over rows of mean_k |log(oracle decile_k / README decile_k)|",
        "defaults_upstream_style": detail(defaults),
        "leaders": [detail(r["combo"]) for r in results[:8]],
        "marginals": {k: {str(v): float(np.mean([r["distance"] for r in results if r["combo"][k] == v])) for v in GRID[k]} for k in keys},
        "best_per_pole_width": {str(w): min((r for r in results if r["combo"]["pole_width"] == w), key=lambda r: r["distance"]) for w in GRID["pole_width"]},
        "n_combinations": len(results),
        "all": [{"combo": r["combo"], "distance": round(r["distance"], 4), "medians": r["medians"]} for r in results],
    }
    path = os.path.join(ROOT, "profiles", "oracle_calibration.json" if not quick else "oracle_calibration_quick.json")
    with open(path, "w") as fh:
        json.dump(blob, fh, indent=1)
    print("written", path, f"{time.time() - t0:.0f}s")
    print("defaults", blob["defaults_upstream_style"]["distance"], blob["defaults_upstream_style"]["deciles"])
    for r in blob["leaders"][:3]:
        print(r["distance"], r["combo"], r["deciles"])
