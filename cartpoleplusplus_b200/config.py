"""ctypes mirror of `bcp_config` (include/bcp.h) and its construction from opts + URDFs.

Every engine-side number lives in `PhysicsParams` as a *parameter*.  The values are
pybullet's built-in defaults as recalled from the public Bullet sources ([BULLET-RECALL]:
the reference never calls setTimeStep / setPhysicsEngineParameter / changeDynamics,
bullet_cartpole.py:151-166, so defaults are what it runs with); they are not pinned by
anything in the reference tree and a correction is a one-line change here.
"""
from __future__ import annotations

import ctypes
import dataclasses
import math
import struct
from typing import Optional

from . import urdf as _urdf

BCP_ABI_VERSION = 8
DTYPE_F32, DTYPE_F64 = 0, 1
REASON_NAMES = {0: None, 1: "out of position bounds", 2: "out of orientation bounds", 3: "episode length",
                4: "resetting", 5: "fresh"}
REASON_RESETTING, REASON_FRESH = 4, 5
AUTO_RESET_OFF, AUTO_RESET_PIPELINED = 0, 1
FRAME_LINK, FRAME_WORLD = 0, 1
THETA_PER_SUBSTEP, THETA_PER_RESET = 0, 1

# canonical state record (include/bcp.h)
S_CART, S_POLE, S_PEND_F, S_PEND_T, S_MANI = 0, 13, 26, 29, 32
MANI_POINT, MANI_STRIDE, S_MANI_CNT = 11, 44, 164
S_STEPS, S_DONE, S_EPISODE, S_RESET_LEFT, STATE_DOUBLES = 167, 168, 169, 170, 176
STATS_HEADER = 4


def _f32(x: float) -> float:
    """Value of the C float literal `x f` widened to double (Bullet writes 0.04f etc.)."""
    return struct.unpack("f", struct.pack("f", x))[0]


class BcpBody(ctypes.Structure):
    _fields_ = [
        ("mass", ctypes.c_double),
        ("inertia", ctypes.c_double * 3),
        ("half_extents", ctypes.c_double * 3),
        ("friction", ctypes.c_double),
        ("restitution", ctypes.c_double),
        ("init_pos", ctypes.c_double * 3),
    ]


class BcpConfig(ctypes.Structure):
    _fields_ = [
        ("abi_version", ctypes.c_int32),
        ("num_envs", ctypes.c_int32),
        ("env_id_offset", ctypes.c_int64),
        ("seed", ctypes.c_uint64),
        ("dtype", ctypes.c_int32),
        ("device_ordinal", ctypes.c_int32),
        ("discrete_actions", ctypes.c_int32),
        ("action_repeats", ctypes.c_int32),
        ("steps_per_repeat", ctypes.c_int32),
        ("max_episode_len", ctypes.c_int32),
        ("initial_force_steps", ctypes.c_int32),
        ("settle_steps", ctypes.c_int32),
        ("random_theta", ctypes.c_int32),
        ("theta_mode", ctypes.c_int32),
        ("force_frame", ctypes.c_int32),
        ("termination_enabled", ctypes.c_int32),
        ("carry_over_reset", ctypes.c_int32),
        ("auto_reset", ctypes.c_int32),
        ("action_force", ctypes.c_double),
        ("initial_force", ctypes.c_double),
        ("pos_threshold", ctypes.c_double),
        ("angle_threshold", ctypes.c_double),
        ("sin_angle_threshold", ctypes.c_double),
        ("tan_angle_threshold", ctypes.c_double),
        ("ground", BcpBody),
        ("cart", BcpBody),
        ("pole", BcpBody),
        ("dt", ctypes.c_double),
        ("gravity", ctypes.c_double * 3),
        ("solver_iterations", ctypes.c_int32),
        ("friction_dirs", ctypes.c_int32),
        ("cone_friction", ctypes.c_int32),
        ("manifold_order", ctypes.c_int32),
        ("residual_threshold", ctypes.c_double),
        ("erp2", ctypes.c_double),
        ("linear_slop", ctypes.c_double),
        ("global_cfm", ctypes.c_double),
        ("friction_erp", ctypes.c_double),
        ("friction_cfm", ctypes.c_double),
        ("restitution_velocity_threshold", ctypes.c_double),
        ("linear_damping", ctypes.c_double),
        ("angular_damping", ctypes.c_double),
        ("max_coordinate_velocity", ctypes.c_double),
        ("contact_breaking_factor", ctypes.c_double),
        ("max_friction", ctypes.c_double),
        ("warmstart_factor", ctypes.c_double),
        ("gyro_term", ctypes.c_int32),
        ("reset_settle_cache", ctypes.c_int32),
        ("cart_ground_friction", ctypes.c_double),
    ]


@dataclasses.dataclass
class PhysicsParams:
    """pybullet defaults ([BULLET-RECALL]); see SURVEY.md section 8a for the provenance."""

    dt: float = 1.0 / 240.0  # fixedTimeStep; no setTimeStep in bullet_cartpole.py
    gravity: tuple = (0.0, 0.0, -9.81)  # bullet_cartpole.py:152
    solver_iterations: int = 50  # PhysicsServerCommandProcessor::createEmptyDynamicsWorld
    friction_dirs: int = 2  # btMultiBodyDynamicsWorld ctor: SOLVER_USE_2_FRICTION_DIRECTIONS
    cone_friction: int = 1  # SOLVER_DISABLE_IMPLICIT_CONE_FRICTION not set
    manifold_order: int = 0  # island sort reverses equal-key manifolds (btAlignedObjectArray::quickSort)
    residual_threshold: float = 1e-7  # m_leastSquaresResidualThreshold
    erp2: float = 0.08
    linear_slop: float = 0.00001
    global_cfm: float = 0.0
    friction_erp: float = 0.2
    friction_cfm: float = 0.0
    restitution_velocity_threshold: float = _f32(0.2)
    linear_damping: float = _f32(0.04)  # btMultiBody::m_linearDamping(0.04f)
    angular_damping: float = _f32(0.04)
    max_coordinate_velocity: float = 100.0
    contact_breaking_factor: float = 0.02  # gContactBreakingThreshold
    max_friction: float = 10.0  # MAX_FRICTION in btManifoldResult.cpp
    # Contact warm starting: cached normal impulse x factor seeds the next substep's solve.  NOT PINNED by the reference
    # (parity with pybullet is unpinned) - a calibrated choice: Bullet's btContactSolverInfo default
    # (SOLVER_USE_WARMSTARTING, m_warmstartingFactor = 0.85); later Bullet versions compile warm starting out for btMultiBody
    # contacts or scale it by 0.1.  In the joint calibration against all four README rows
    # (profiles/oracle_calibration.json, 1 152 combinations) the three candidates 0 / 0.1 / 0.85 barely differ on the
    # stochastic rows (marginal distance 1.25 / 1.23 / 1.08); 0.85 is the best of them and the one under which the README's
    # exact known answer (README.md:77-80: --initial-force=0 --actions=0 => every episode lasts 200 steps) holds with the
    # shipped 1 cm pole as well as with the wider poles (tests/test_oracle_kat.py::test_readme_stability_protocol).
    warmstart_factor: float = 0.85
    gyro_term: int = 1
    inertia_from_file: bool = False  # loadURDF without URDF_USE_INERTIA_FROM_FILE


# reset / load poses, bullet_cartpole.py:154-156,319-320
INIT_POS = {"ground": (0.0, 0.0, 0.0), "cart": (0.0, 0.0, 0.08), "pole": (0.0, 0.0, 0.35)}
# constants of the ctor, bullet_cartpole.py:58,62,76 and the settle loop :326
POS_THRESHOLD = 3.0
ANGLE_THRESHOLD = 0.35
INITIAL_FORCE_STEPS = 30
SETTLE_STEPS = 100


def _fill_body(dst: BcpBody, body: _urdf.BoxBody, init_pos, inertia_from_file: bool) -> None:
    dst.mass = body.mass
    for i, v in enumerate(body.effective_inertia(inertia_from_file)):
        dst.inertia[i] = v
    for i, v in enumerate(body.half_extents):
        dst.half_extents[i] = v
    dst.friction = body.lateral_friction
    dst.restitution = body.restitution
    for i, v in enumerate(init_pos):
        dst.init_pos[i] = v


def make_config(opts, discrete_actions: bool, num_envs: Optional[int] = None, dtype: Optional[str] = None,
                seed: Optional[int] = None, device: Optional[int] = None, env_id_offset: int = 0,
                physics: Optional[PhysicsParams] = None, scene=None, init_offset=(0.0, 0.0)) -> BcpConfig:
    """Build the POD config from an argparse namespace (`opts.add_opts` flags).
    init_offset: (x, y) added to the cart's and the pole's load / reset position - the fork's second cart-pole is
    the same pair of bodies loaded at x = 1 (bullet_cartpole.py:159-160,322-323)."""
    g = lambda name, default: getattr(opts, name, default)  # noqa: E731 (reference opts lack the extras)
    if g("num_cameras", 1) not in (1, 2):
        raise ValueError("--num-cameras must be 1 or 2")  # bullet_cartpole.py:113-114
    assert g("reward_calc", "fixed") in ["fixed", "angle", "action", "angle_action"]  # :144
    if g("gui", False):
        raise NotImplementedError("--gui needs a display: unavailable on this box")
    physics = physics or PhysicsParams()
    ground, cart, pole = scene if scene is not None else _urdf.load_scene(g("model_dir", None), g("cart_model", "cart"))
    if g("cart_friction", None) is not None:
        cart = dataclasses.replace(cart, lateral_friction=float(opts.cart_friction))

    c = BcpConfig()
    c.abi_version = BCP_ABI_VERSION
    c.num_envs = int(num_envs if num_envs is not None else g("num_envs", 1))
    c.env_id_offset = int(env_id_offset)
    c.seed = int(seed if seed is not None else g("seed", 0)) & 0xFFFFFFFFFFFFFFFF
    dt_name = dtype if dtype is not None else g("dtype", "fp32")
    if dt_name not in ("fp32", "fp64"):
        raise ValueError(f"dtype must be fp32 or fp64, got {dt_name!r}")
    c.dtype = DTYPE_F64 if dt_name == "fp64" else DTYPE_F32
    c.device_ordinal = int(device if device is not None else g("device", 0))
    c.discrete_actions = 1 if discrete_actions else 0
    c.action_repeats = int(opts.action_repeats)
    c.steps_per_repeat = int(opts.steps_per_repeat)
    c.max_episode_len = int(opts.max_episode_len)
    if c.action_repeats < 1 or c.steps_per_repeat < 1 or c.max_episode_len < 1 or c.num_envs < 1:
        raise ValueError("action_repeats, steps_per_repeat, max_episode_len and num_envs must be >= 1")
    c.initial_force_steps = INITIAL_FORCE_STEPS
    c.settle_steps = SETTLE_STEPS
    c.random_theta = 0 if g("no_random_theta", False) else 1
    c.theta_mode = THETA_PER_RESET if g("theta_per_reset", False) else THETA_PER_SUBSTEP
    c.force_frame = FRAME_WORLD if g("world_frame_force", False) else FRAME_LINK
    c.termination_enabled = 0 if g("no_termination", False) else 1
    c.carry_over_reset = 0 if g("no_carry_over_reset", False) else 1
    c.reset_settle_cache = 1 if g("reset_settle_cache", False) else 0
    if c.reset_settle_cache and c.carry_over_reset:
        raise ValueError("--reset-settle-cache needs --no-carry-over-reset: with the pending force and the contact caches carried "
                         "over (pybullet's behaviour) every env settles from a different state")
    ar = g("auto_reset", "off")
    if ar not in ("off", "pipelined"):
        raise ValueError(f"auto_reset must be 'off' or 'pipelined', got {ar!r}")
    c.auto_reset = AUTO_RESET_PIPELINED if ar == "pipelined" else AUTO_RESET_OFF
    c.action_force = float(opts.action_force)
    c.initial_force = float(opts.initial_force)
    c.pos_threshold = POS_THRESHOLD
    c.angle_threshold = ANGLE_THRESHOLD
    c.sin_angle_threshold = math.sin(ANGLE_THRESHOLD)
    c.tan_angle_threshold = math.tan(ANGLE_THRESHOLD)
    c.cart_ground_friction = -1.0 if cart.ground_friction is None else float(cart.ground_friction)
    _fill_body(c.ground, ground, INIT_POS["ground"], physics.inertia_from_file)
    off = (float(init_offset[0]), float(init_offset[1]), 0.0)
    _fill_body(c.cart, cart, tuple(a + b for a, b in zip(INIT_POS["cart"], off)), physics.inertia_from_file)
    _fill_body(c.pole, pole, tuple(a + b for a, b in zip(INIT_POS["pole"], off)), physics.inertia_from_file)
    c.dt = physics.dt
    for i in range(3):
        c.gravity[i] = physics.gravity[i]
    for name in ("solver_iterations", "friction_dirs", "cone_friction", "manifold_order", "residual_threshold",
                 "erp2", "linear_slop", "global_cfm", "friction_erp", "friction_cfm",
                 "restitution_velocity_threshold", "linear_damping", "angular_damping",
                 "max_coordinate_velocity", "contact_breaking_factor", "max_friction", "warmstart_factor",
                 "gyro_term"):
        setattr(c, name, getattr(physics, name))
    if c.friction_dirs not in (1, 2):
        raise ValueError("friction_dirs must be 1 or 2")
    return c


def copy_config(c: BcpConfig) -> BcpConfig:
    out = BcpConfig()
    ctypes.memmove(ctypes.byref(out), ctypes.byref(c), ctypes.sizeof(BcpConfig))
    return out
