"""Host logic: flags, URDF parsing, config mirror, spaces (CPU tier)."""
import argparse
import ctypes
import json
import os

import numpy as np
import pytest

from cartpoleplusplus_b200 import add_opts, default_opts, make_config, spaces, urdf
from cartpoleplusplus_b200 import config as C

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_add_opts_matches_reference_defaults():
    """Every flag of bullet_cartpole.add_opts (bullet_cartpole.py:15-40) exists with the same default."""
    gold = json.load(open(os.path.join(GOLD, "add_opts_defaults.json")))
    p = argparse.ArgumentParser()
    add_opts(p)
    ns = p.parse_args([])
    assert len(gold) == 14
    for flag, default in gold.items():
        assert getattr(ns, flag[2:].replace("-", "_")) == default, flag


def test_urdf_values_match_reference_models():
    """models/{ground,cart,pole}.urdf constants (values, not file text) equal the reference's."""
    gold = json.load(open(os.path.join(GOLD, "model_constants.json")))
    for body in urdf.load_scene():
        g = gold[body.name]
        assert body.mass == g["mass"]
        assert list(body.inertia_file) == g["inertia"]
        assert list(body.box_size) == g["box"]
        assert body.lateral_friction == g["contact"].get("lateral_friction", 0.5)
        assert body.inertia_scaling == g["contact"].get("inertia_scaling", 1.0)
        assert body.restitution == 0.0


def test_box_inertia_equals_file_inertia():
    """The URDF inertias are the box formula, so recomputing from the shape agrees."""
    _, cart, pole = urdf.load_scene()
    for b in (cart, pole):
        assert np.allclose(b.box_inertia(), b.inertia_file, rtol=1e-9)
        assert np.allclose(b.effective_inertia(), 3.0 * np.array(b.inertia_file), rtol=1e-9)


def test_config_fields():
    cfg = make_config(default_opts(), True, num_envs=7, dtype="fp64", seed=5)
    assert (cfg.num_envs, cfg.dtype, cfg.seed, cfg.discrete_actions) == (7, C.DTYPE_F64, 5, 1)
    assert (cfg.action_repeats, cfg.steps_per_repeat, cfg.max_episode_len) == (2, 1, 200)
    assert (cfg.initial_force_steps, cfg.settle_steps) == (30, 100)
    assert (cfg.action_force, cfg.initial_force, cfg.pos_threshold, cfg.angle_threshold) == (50.0, 200.0, 3.0, 0.35)
    assert list(cfg.cart.init_pos) == [0, 0, 0.08] and list(cfg.pole.init_pos) == [0, 0, 0.35]
    assert list(cfg.ground.half_extents) == [1.5, 1.5, 0.05]
    assert cfg.cart.friction == 0.0 and cfg.pole.friction == 1.0 and cfg.ground.friction == 0.5
    assert cfg.dt == 1.0 / 240.0 and list(cfg.gravity) == [0, 0, -9.81]
    assert cfg.solver_iterations == 50 and cfg.residual_threshold == 1e-7 and cfg.erp2 == 0.08
    assert cfg.theta_mode == C.THETA_PER_SUBSTEP and cfg.force_frame == C.FRAME_LINK and cfg.termination_enabled == 1
    up = make_config(default_opts(theta_per_reset=True, world_frame_force=True, no_termination=True, cart_friction=1.0), False)
    assert (up.theta_mode, up.force_frame, up.termination_enabled, up.cart.friction) == (1, 1, 0, 1.0)


def test_config_errors_mirror_reference():
    with pytest.raises(ValueError, match="--num-cameras must be 1 or 2"):  # bullet_cartpole.py:113-114
        make_config(default_opts(num_cameras=3), True)
    with pytest.raises(AssertionError):  # :144
        make_config(default_opts(reward_calc="bogus"), True)
    with pytest.raises(NotImplementedError):
        make_config(default_opts(gui=True), True)
    make_config(default_opts(use_raw_pixels=True), True)  # the pixel observation is rendered by the env (render.py)
    with pytest.raises(ValueError):
        make_config(default_opts(action_repeats=0), True)
    with pytest.raises(ValueError):
        make_config(default_opts(), True, dtype="fp16")


def test_urdf_parser_rejects_unsupported(tmp_path):
    bad = tmp_path / "bad.urdf"
    bad.write_text('<robot name="x"><link name="a"><inertial><mass value="1"/><inertia ixx="1" ixy="0.1" ixz="0" '
                   'iyy="1" iyz="0" izz="1"/></inertial><collision><geometry><box size="1 1 1"/></geometry>'
                   '</collision></link></robot>')
    with pytest.raises(ValueError, match="off-diagonal"):
        urdf.parse_box_urdf(str(bad))
    bad.write_text('<robot name="x"><link name="a"><inertial><mass value="1"/><inertia ixx="1" ixy="0" ixz="0" '
                   'iyy="1" iyz="0" izz="1"/></inertial><collision><geometry><sphere radius="1"/></geometry>'
                   '</collision></link></robot>')
    with pytest.raises(ValueError, match="box"):
        urdf.parse_box_urdf(str(bad))


def test_spaces():
    d = spaces.Discrete(5)
    assert d.n == 5 and all(0 <= d.sample() < 5 for _ in range(20)) and d.contains(4) and not d.contains(5)
    b = spaces.Box(-1.0, 1.0, shape=(1, 2))
    assert b.shape == (1, 2) and b.contains(b.sample()) and not b.contains(np.array([[2.0, 0.0]]))
    assert b.shape[1] == 2  # what ddpg/naf read
    fm = np.finfo(np.float32).max
    o = spaces.Box(-fm, fm, (2, 2, 7))
    assert o.shape == (2, 2, 7) and o.high.max() == fm


def test_struct_mirror_size_is_stable():
    assert ctypes.sizeof(C.BcpBody) == 12 * 8
    assert ctypes.sizeof(C.BcpConfig) % 8 == 0
