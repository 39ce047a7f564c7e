"""CPU restatement of the fork's 8-state observation, done rule and LQR feedback.

TEST INFRASTRUCTURE ONLY: imported by tests/ (and nothing in the product path).

Follows random_action_agent.py of the reference tree:
  :121-135  _observe_state       -> lqr_state
  :108-119  _check_done          -> check_done
  :92-95    lqr_control_forces   -> control_forces
  :888-903  disturbance + control in the game loop -> actions
It works on canonical state records (include/bcp.h, BCP_STATE_DOUBLES doubles per env), i.e. on
what `BatchedBulletCartpole.get_state()` and the C oracle (`oracle.OracleEnv.state`) both hold.

`euler_from_quaternion` restates pybullet's getEulerFromQuaternion.  pybullet (the reference's
`import pybullet as p`, no version pinned anywhere in the tree) is absent from /root/reference and
from this image, so that one function is "parity unpinned" against pybullet itself; it is pinned
instead against scipy's extrinsic x-y-z Euler angles (tests/test_lqr_oracle.py), which is the
convention pybullet documents for getEulerFromQuaternion.
"""
from __future__ import annotations

import numpy as np

S_POLE = 13  # canonical record: pole body = x[3] q[4] (x,y,z,w) v[3] w[3] at doubles 13..25
X, X_DOT, Y, Y_DOT, ROLL, ROLL_DOT, PITCH, PITCH_DOT = range(8)  # random_action_agent.py:55-62
RUNNING, POSITION, ANGLE = 0, 1, 2


def euler_from_quaternion(q):
    """q (..., 4) as x,y,z,w -> (..., 3) roll, pitch, yaw with pybullet's gimbal-lock branches."""
    q = np.asarray(q, dtype=np.float64)
    x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    sqx, sqy, sqz, squ = x * x, y * y, z * z, w * w
    sarg = -2.0 * (x * z - w * y)
    lo, hi = sarg <= -0.99999, sarg >= 0.99999
    with np.errstate(invalid="ignore"):
        roll = np.arctan2(2.0 * (y * z + w * x), squ - sqx - sqy + sqz)
        pitch = np.arcsin(np.clip(sarg, -1.0, 1.0))
        yaw = np.arctan2(2.0 * (x * y + w * z), squ + sqx - sqy - sqz)
    roll = np.where(lo | hi, 0.0, roll)
    pitch = np.where(lo, -0.5 * np.pi, np.where(hi, 0.5 * np.pi, pitch))
    yaw = np.where(lo, 2.0 * np.arctan2(x, -y), np.where(hi, 2.0 * np.arctan2(-x, y), yaw))
    return np.stack([roll, pitch, yaw], axis=-1)


def lqr_state(records, x0=None):
    """records (B, STATE_DOUBLES) float64 -> (B, 8) float32  (:121-135)."""
    r = np.asarray(records, dtype=np.float64)
    b = r[:, S_POLE:S_POLE + 13]
    rpy = euler_from_quaternion(b[:, 3:7])
    st = np.zeros((r.shape[0], 8), dtype=np.float64)
    st[:, X] = b[:, 0] - (0.0 if x0 is None else np.asarray(x0, dtype=np.float32).astype(np.float64))
    st[:, Y] = b[:, 1]
    st[:, ROLL] = rpy[:, 0]
    st[:, PITCH] = rpy[:, 1]
    st[:, X_DOT] = b[:, 7]
    st[:, Y_DOT] = b[:, 8]
    st[:, ROLL_DOT] = b[:, 10]
    st[:, PITCH_DOT] = b[:, 11]
    return st.astype(np.float32)


def check_done(state, position_threshold, angle_threshold):
    """(B, 8) float32 -> (B,) uint8 codes; position wins over angle like the if/elif of :108-119."""
    s = np.asarray(state, dtype=np.float32)
    pt, at = np.float32(position_threshold), np.float32(angle_threshold)
    pos = (np.abs(s[:, X]) > pt) | (np.abs(s[:, Y]) > pt)
    ang = (np.abs(s[:, ROLL]) > at) | (np.abs(s[:, PITCH]) > at)
    return np.where(pos, POSITION, np.where(ang, ANGLE, RUNNING)).astype(np.uint8)


def control_forces(state, gains, zero_point=None):
    """-(K . (state - zero_point)) per env, float32, summed left to right (:92-95).
    gains (2, 8) or (B, 2, 8)."""
    s = np.asarray(state, dtype=np.float32)
    K = np.asarray(gains, dtype=np.float32)
    if K.ndim == 2:
        K = np.broadcast_to(K, (s.shape[0], 2, 8))
    z = np.zeros(8, np.float32) if zero_point is None else np.asarray(zero_point, dtype=np.float32)
    d = s - z[None, :]
    acc = np.zeros((s.shape[0], 2), dtype=np.float32)
    for j in range(8):
        acc = (acc + K[:, :, j] * d[:, None, j]).astype(np.float32)
    return -acc


def actions(state, gains=None, zero_point=None, disturbance=None):
    """disturbance + control (:888-899): the (fx, fy) handed to step()."""
    s = np.asarray(state, dtype=np.float32)
    u = np.zeros((s.shape[0], 2), np.float32) if gains is None else control_forces(s, gains, zero_point)
    if disturbance is not None:
        u = (u + np.broadcast_to(np.asarray(disturbance, dtype=np.float32), u.shape)).astype(np.float32)
    return u
