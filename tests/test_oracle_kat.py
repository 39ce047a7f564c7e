"""Oracle pins (CPU tier).  pybullet golden vectors do not exist (parity
unpinned), so the oracle is pinned by (a) analytic known answers, (b) the reference's own
bookkeeping rules (bullet_cartpole.py, cited per test), (c) the README's exact known-answer
stability protocol (README.md:77-80) and (d) published KATs of the building blocks."""
import dataclasses
import math

import numpy as np
import pytest

from cartpoleplusplus_b200 import PhysicsParams, default_opts, make_config, urdf
from cartpoleplusplus_b200 import config as C
from oracle import Oracle
from oracle.oracle import box_box, elem, philox

G = 9.81
DT = 1.0 / 240.0
KL = C._f32(0.04)


def mk(n=1, discrete=True, dtype="fp64", physics=None, scene=None, **ov):
    cfg = make_config(default_opts(**ov), discrete, num_envs=n, dtype=dtype, seed=3, physics=physics, scene=scene)
    return Oracle(cfg)


def fresh(o, cart=(0, 0, 0.08), pole=(0, 0, 0.35)):
    o.state[:] = 0
    o.state[:, C.S_CART:C.S_CART + 3] = cart
    o.state[:, C.S_POLE:C.S_POLE + 3] = pole
    o.state[:, C.S_CART + 6] = 1
    o.state[:, C.S_POLE + 6] = 1


def test_philox_random123_kat():
    """Random123 kat_vectors: philox4x32-10, ctr=0, key=0 and the all-ones vector."""
    assert [hex(v) for v in philox(0, 0, 0, 0, 0)] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    out = philox(0xFFFFFFFFFFFFFFFF, -1, 0xFFFFFFFF, 0xFF, 0xFFFFFF)
    assert [hex(v) for v in out] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]


def test_elementary_functions_vs_libm():
    rng = np.random.default_rng(0)
    for x in rng.uniform(-math.pi / 4, math.pi / 4, 200):
        e = elem(x)
        assert abs(e[0] - math.sin(x)) < 3e-16 and abs(e[1] - math.cos(x)) < 3e-16
    for k in rng.integers(0, 1 << 24, 200):
        u = k / float(1 << 24)
        e = elem(u)
        assert abs(e[3] - math.sin(2 * math.pi * u)) < 2e-15 and abs(e[4] - math.cos(2 * math.pi * u)) < 2e-15
    for x, y in rng.normal(size=(300, 2)):
        assert abs(elem(x, y)[2] - math.atan2(y, x)) < 2e-15
    assert elem(0.0, 0.0)[2] == 0.0 and abs(elem(-1.0, 0.0)[2] - math.pi) < 1e-15


def test_free_fall_semi_implicit_euler_with_damping():
    """No contact: v_{k+1} = v_k + dt (g - v_k (k + k |v_k|)), z_{k+1} = z_k + dt v_{k+1}."""
    o = mk()
    fresh(o, cart=(0, 0, 1.0), pole=(0.5, 0, 2.0))
    z, v = 1.0, 0.0
    for _ in range(20):
        o.substep()
        v = v + DT * (-G - v * (KL + KL * abs(v)))
        z = z + DT * v
    assert abs(o.state[0, C.S_CART + 2] - z) < 1e-14 and abs(o.state[0, C.S_CART + 9] - v) < 1e-14
    assert np.all(o.state[0, C.S_MANI_CNT:C.S_MANI_CNT + 3] == 0)


def test_force_acts_one_substep_late_and_is_consumed_once():
    """applyExternalForce after stepSimulation (bullet_cartpole.py:199-203): the force drives the NEXT substep only."""
    o = mk()
    fresh(o, cart=(0, 0, 1.0), pole=(0.5, 0, 2.0))  # airborne: no contacts, frictionless anyway
    o.apply_force([50.0, 0.0])
    assert o.state[0, C.S_PEND_F] == 50.0 and o.state[0, C.S_CART + 7] == 0.0
    o.substep()
    assert abs(o.state[0, C.S_CART + 7] - 50.0 * DT) < 1e-15  # a = F/m, m = 1
    assert o.state[0, C.S_PEND_F] == 0.0
    vx, vn = o.state[0, C.S_CART + 7], np.linalg.norm(o.state[0, C.S_CART + 7:C.S_CART + 10])
    o.substep()
    assert abs(o.state[0, C.S_CART + 7] - (vx - DT * vx * (KL + KL * vn))) < 1e-15  # only damping now


def test_link_frame_vs_world_frame_force():
    """LINK_FRAME rotates the force by the cart orientation, zero torque; WORLD_FRAME applies it at the world origin."""
    o = mk()
    fresh(o, cart=(0.3, 0.2, 1.0))
    s = math.sin(math.pi / 4)
    o.state[0, C.S_CART + 3:C.S_CART + 7] = (0, 0, s, s)  # yaw 90 degrees
    o.apply_force([10.0, 0.0])
    assert np.allclose(o.state[0, C.S_PEND_F:C.S_PEND_F + 3], [0, 10, 0], atol=1e-14)
    assert np.allclose(o.state[0, C.S_PEND_T:C.S_PEND_T + 3], 0, atol=1e-14)
    w = mk(world_frame_force=True)
    fresh(w, cart=(0.3, 0.2, 1.0))
    w.apply_force([10.0, 0.0])
    assert np.allclose(w.state[0, C.S_PEND_F:C.S_PEND_F + 3], [10, 0, 0])
    assert np.allclose(w.state[0, C.S_PEND_T:C.S_PEND_T + 3], np.cross([-0.3, -0.2, -1.0], [10, 0, 0]))


def test_box_box_face_contact_known_answer():
    """Unit cube resting 1 mm into a big slab: 4 corner points, normal from slab (B->A = -z ... on B), depth -1e-3."""
    n, c = box_box([0, 0, 0], [0, 0, 0, 1], [2, 2, 0.5], [0.2, -0.1, 0.999], [0, 0, 0, 1], [0.5, 0.5, 0.5])
    assert n == 4
    assert np.allclose(c[:, 0:3], [0, 0, -1]) and np.allclose(c[:, 6], -1e-3, atol=1e-15)
    corners = sorted((round(x, 12), round(y, 12)) for x, y in c[:, 3:5])
    assert corners == [(-0.3, -0.6), (-0.3, 0.4), (0.7, -0.6), (0.7, 0.4)]
    assert np.allclose(c[:, 5], 0.499)  # points lie on B's bottom face
    # separated along z: no contacts at all (the detector only reports penetration)
    assert box_box([0, 0, 0], [0, 0, 0, 1], [2, 2, 0.5], [0, 0, 1.01], [0, 0, 0, 1], [0.5, 0.5, 0.5])[0] == 0


def test_box_box_edge_edge_and_cull():
    # two crossed bars touching edge to edge -> a single contact (code > 6)
    q45x = [math.sin(math.pi / 8), 0, 0, math.cos(math.pi / 8)]
    q45y = [0, math.sin(math.pi / 8), 0, math.cos(math.pi / 8)]
    d = 2 * 0.1 * math.sqrt(2) - 0.01
    n, c = box_box([0, 0, 0], q45x, [1, 0.1, 0.1], [0, 0, d], q45y, [0.1, 1, 0.1])
    # (edge axes carry the detector's 1e-5 fudge on |R|, hence the loose depth tolerance)
    assert n == 1 and abs(c[0, 6] + 0.01) < 1e-5 and np.allclose(np.abs(c[0, 0:3]), [0, 0, 1], atol=1e-12)
    # octagonal overlap (square rotated 45 deg on a square) -> more than 4 clipped points, culled to 4
    qz = [0, 0, math.sin(math.pi / 8), math.cos(math.pi / 8)]
    n, c = box_box([0, 0, 0], [0, 0, 0, 1], [0.5, 0.5, 0.5], [0, 0, 0.99], qz, [0.5, 0.5, 0.5])
    assert n == 4 and np.allclose(c[:, 6], -0.01, atol=1e-12)


def test_cart_settles_on_ground_at_slop_penetration():
    """Resting contact converges to the solver's linear slop (README sample shows z ~ 0.0749)."""
    o = mk()
    fresh(o, pole=(1.0, 0, 2.0))  # pole far away, falling freely for the short test
    for _ in range(60):
        o.substep()
    z = o.state[0, C.S_CART + 2]
    assert 0.0749 < z < 0.07501 and abs(o.state[0, C.S_CART + 9]) < 5e-3
    assert o.state[0, C.S_MANI_CNT + C.S_CART] == 4  # four corner points, ground-cart manifold


def test_reset_runs_130_substeps_and_bootstraps_obs():
    """reset: teleport, 100 settle + 30 push substeps, every repeat slot holds the same pose (:313-346)."""
    o = mk(initial_force=0.0, action_repeats=3)
    obs = o.reset()
    assert obs.shape == (1, 3, 2, 7) and obs.dtype == np.float32
    assert np.array_equal(obs[0, 0], obs[0, 1]) and np.array_equal(obs[0, 0], obs[0, 2])
    m = mk(initial_force=0.0)
    fresh(m)
    for _ in range(130):
        m.substep()
    assert np.array_equal(m.state[0, :26], o.state[0, :26])
    assert np.array_equal(obs[0, 0, 0], m.state[0, 0:7].astype(np.float32))  # pose order x,y,z,qx,qy,qz,qw (:43-45)
    assert (o.state[0, C.S_STEPS], o.state[0, C.S_DONE], o.state[0, C.S_EPISODE]) == (0, 0, 1)


def test_step_bookkeeping():
    """steps counts env.steps (:239), done-by-length uses >= (:255), reward 1 incl. the terminal step (:260),
    step after done returns the old state, reward 0, done (:179-181)."""
    o = mk(initial_force=0.0, max_episode_len=3, no_termination=True)
    o.reset()
    for t in range(3):
        obs, r, d, reason = o.step(0)
        assert r[0] == 1.0 and bool(d[0]) == (t == 2) and reason[0] == (3 if t == 2 else 0)
    st = o.state.copy()
    obs2, r2, d2, reason2 = o.step(0)
    assert r2[0] == 0.0 and d2[0] and np.array_equal(obs2, obs) and np.array_equal(o.state, st)


def test_discrete_action_table_and_continuous_scaling():
    for a, f in [(0, (0, 0)), (1, (50, 0)), (2, (-50, 0)), (3, (0, 50)), (4, (0, -50))]:
        o = mk(initial_force=0.0)
        o.reset()
        o.step(a)
        # LINK_FRAME: the table force rotated by the (slightly yawed) cart orientation
        got = o.state[0, C.S_PEND_F:C.S_PEND_F + 3]
        assert abs(np.linalg.norm(got) - np.hypot(*f)) < 1e-12 and np.allclose(got[:2], f, atol=5e-2)
    o = mk(discrete=False, initial_force=0.0, action_force=100.0)
    o.reset()
    o.step(np.array([[0.25, -0.5]], np.float32))
    assert np.allclose(o.state[0, C.S_PEND_F:C.S_PEND_F + 2], [25, -50], atol=1e-2)


def test_done_rule_matches_pybullet_euler_formulas():
    """Trig-free thresholds == abs(roll), abs(pitch) > 0.35 with pybullet's getEulerFromQuaternion formulas."""
    o = mk()
    rng = np.random.default_rng(1)
    checked = 0
    for _ in range(4000):
        ang = rng.uniform(-0.6, 0.6, 3)
        cr, sr, cp, sp, cy, sy = (f(a / 2) for a in ang for f in (math.cos, math.sin))
        q = [sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy]
        x, y, z, w = q
        sarg = -2 * (x * z - w * y)
        roll = math.atan2(2 * (y * z + w * x), w * w - x * x - y * y + z * z)
        pitch = math.asin(max(-1, min(1, sarg)))
        if min(abs(abs(roll) - 0.35), abs(abs(pitch) - 0.35)) < 1e-9:
            continue
        fresh(o)
        o.state[0, C.S_POLE + 3:C.S_POLE + 7] = q
        expect = 2 if (abs(roll) > 0.35 or abs(pitch) > 0.35) else 0
        assert o.done_reason() == expect
        checked += 1
    assert checked > 3900
    fresh(o)
    o.state[0, C.S_POLE] = 3.0001  # position checked first (:245)
    o.state[0, C.S_POLE + 3:C.S_POLE + 7] = (math.sin(0.3), 0, 0, math.cos(0.3))
    assert o.done_reason() == 1
    o.state[0, C.S_POLE] = 3.0  # strict >
    assert o.done_reason() == 2


@pytest.mark.parametrize("kw", [dict(), dict(cart_friction=1.0, steps_per_repeat=5)])
def test_readme_stability_protocol(kw):
    """README.md:77-80, the reference's one exact known answer: --initial-force=0 --actions="0"
    => every episode lasts max_episode_len = 200 steps.  This pins the contact warm start
    (config.PhysicsParams.warmstart_factor): without it the pole falls after ~49 steps."""
    from cartpoleplusplus_b200 import protocol
    cfg = make_config(default_opts(initial_force=0.0, **kw), True, num_envs=8, dtype="fp64", seed=7)
    ora = Oracle(cfg, dtype="fp64")
    ora.state[:, C.S_DONE] = 1
    ora.rollout_random(450, action_mask=0x1, max_episodes=2)
    assert ora.hist[200] == 16 and ora.hist.sum() == 16
    assert np.array_equal(protocol.deciles_from_histogram(ora.hist), [200] * 11)
    off = C.PhysicsParams(warmstart_factor=0.0)
    cfg0 = make_config(default_opts(initial_force=0.0, **kw), True, num_envs=2, dtype="fp64", seed=7, physics=off)
    ora0 = Oracle(cfg0, dtype="fp64")
    ora0.state[:, C.S_DONE] = 1
    ora0.rollout_random(210, action_mask=0x1, max_episodes=1)
    assert ora0.hist[200] == 0  # the parameter matters: this is what the known answer rules out
