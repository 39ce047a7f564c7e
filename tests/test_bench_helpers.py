"""CPU tier: the pieces of bench.py that turn the kernels' solve records into the roofline block - the counted flop model
(profiles/flops_model.json, scripts/count_flops.py) and SURVEY 8d's algorithmic bytes - so that a broken model file or key
shows up here and not in the round-end bench."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from cartpoleplusplus_b200 import default_opts, make_config  # noqa: E402


def test_flops_model_is_present_and_evaluates():
    for ov, name in ((dict(), "configs2_fork_defaults"), (dict(cart_friction=1.0, steps_per_repeat=5), "upstream_style_friction")):
        cfg = make_config(default_opts(action_force=100.0, **ov), False, num_envs=8)
        got, model = bench.load_flops_model(cfg)
        assert got == name
        assert abs(sum(model["mean_by_phase"].values()) - model["mean_flops_per_substep"]) < 1.0 and model["model_error"] < 0.01
        cnt = np.array([[4, 4, 0, 50], [0, 0, 4, 7], [0, 0, 0, 1], [3, 1, 2, 9]], dtype=np.int64)   # the last regime is rare: fallback or table
        fixed, pgs = bench.substep_flops(model, cnt)
        assert fixed.shape == pgs.shape == (4,) and (fixed > 500).all() and (pgs >= 0).all()
        r = model["regimes"]["4,4,0"]
        assert fixed[0] == r["fixed_total"] and pgs[0] == 50 * r["pgs_per_iteration"]
        assert pgs[2] <= pgs[1]   # no contacts: nothing to sweep


def test_algorithmic_bytes_follow_survey_8d():
    cfg = make_config(default_opts(), False, num_envs=8)          # R = 2, continuous
    state = (26 + 6 + 3 * 4 * 11) * 4 + 8 * 4
    assert bench.algorithmic_bytes_per_env_step(cfg) == 2 * state + 8 + 2 * 56 + 6
    cfg = make_config(default_opts(action_repeats=3), True, num_envs=8)
    assert bench.algorithmic_bytes_per_env_step(cfg) == 2 * state + 4 + 3 * 56 + 6


def test_traffic_file_has_the_per_step_entry():
    t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    assert t["env_step"]["num_envs"] == 65536 and t["env_step"]["bytes_per_step"] > 5e7
