"""oracle/lqr_oracle.py pinned without a GPU: the Euler angles against scipy's extrinsic x-y-z
convention (what pybullet documents for getEulerFromQuaternion), the done rule and the control law
against a line-by-line replay of the fork's scalar code (random_action_agent.py:92-135)."""
import numpy as np
import pytest
from scipy.spatial.transform import Rotation

from oracle import lqr_oracle as L


def test_euler_matches_scipy_extrinsic_xyz():
    rng = np.random.default_rng(0)
    q = rng.normal(size=(2000, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    want = Rotation.from_quat(q).as_euler("xyz")  # x,y,z,w in, extrinsic roll-pitch-yaw out
    got = L.euler_from_quaternion(q)
    keep = np.abs(np.abs(want[:, 1]) - np.pi / 2) > 1e-2  # away from gimbal lock the two agree
    assert keep.sum() > 1900
    assert np.allclose(got[keep], want[keep], atol=1e-9)


@pytest.mark.parametrize("axis,angle", [(0, 0.3), (0, -1.1), (1, 0.2), (1, -0.7), (2, 2.0)])
def test_euler_single_axis(axis, angle):
    v = np.zeros(3)
    v[axis] = 1.0
    q = np.concatenate([v * np.sin(angle / 2), [np.cos(angle / 2)]])
    want = np.zeros(3)
    want[axis] = angle
    assert np.allclose(L.euler_from_quaternion(q[None])[0], want, atol=1e-12)


def test_euler_gimbal_branches():
    for sign in (1.0, -1.0):
        a = sign * np.pi / 2
        q = np.array([0.0, np.sin(a / 2), 0.0, np.cos(a / 2)])  # pitch of +-90 degrees
        r = L.euler_from_quaternion(q[None])[0]
        assert r[0] == 0.0 and r[1] == sign * 0.5 * np.pi


def _records(rng, n):
    rec = np.zeros((n, 176))
    b = rec[:, L.S_POLE:L.S_POLE + 13]
    b[:, 0:3] = rng.normal(scale=2.0, size=(n, 3))
    q = rng.normal(size=(n, 4)) * np.array([0.3, 0.3, 0.05, 1.0])
    b[:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
    b[:, 7:13] = rng.normal(size=(n, 6))
    return rec


def test_state_done_and_control_follow_the_scalar_code():
    rng = np.random.default_rng(1)
    n = 300
    rec = _records(rng, n)
    x0 = rng.normal(scale=0.5, size=n).astype(np.float32)
    st = L.lqr_state(rec, x0)
    K = rng.normal(size=(2, 8)).astype(np.float32)
    zero = rng.normal(scale=0.1, size=8).astype(np.float32)
    done = L.check_done(st, 3.0, np.pi / 4)
    u = L.control_forces(st, K, zero)
    assert set(np.unique(done)) == {0, 1, 2}
    for e in range(n):
        b = rec[e, 13:26]
        rpy = Rotation.from_quat(b[3:7]).as_euler("xyz")
        # _observe_state, :121-135
        ps = np.zeros(8)
        ps[0] = b[0] - float(x0[e]); ps[2] = b[1]; ps[4] = rpy[0]; ps[6] = rpy[1]
        ps[1] = b[7]; ps[3] = b[8]; ps[5] = b[10]; ps[7] = b[11]
        assert np.allclose(st[e], ps, atol=1e-6)
        # _check_done, :108-119, on the float32 state the kernel thresholds
        s32 = st[e]
        if abs(s32[0]) > np.float32(3.0) or abs(s32[2]) > np.float32(3.0):
            want = 1
        elif abs(s32[4]) > np.float32(np.pi / 4) or abs(s32[6]) > np.float32(np.pi / 4):
            want = 2
        else:
            want = 0
        assert done[e] == want
        # lqr_control_forces, :92-95
        assert np.allclose(u[e], -np.dot(K.astype(np.float64), (s32 - zero).astype(np.float64)), rtol=1e-5, atol=1e-5)


def test_actions_add_the_disturbance():
    rng = np.random.default_rng(2)
    st = L.lqr_state(_records(rng, 16))
    K = rng.normal(size=(16, 2, 8)).astype(np.float32)
    d = rng.normal(size=(16, 2)).astype(np.float32)
    assert np.array_equal(L.actions(st, K, None, d), (L.control_forces(st, K) + d).astype(np.float32))
    assert np.array_equal(L.actions(st, None, None, d[0]), np.broadcast_to(d[0], (16, 2)))
