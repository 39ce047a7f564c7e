"""ctypes front-end of oracle/libbcp_oracle.so (see bcp_oracle.c for what it restates).

`Oracle` mirrors the env surface of bullet_cartpole.py (reset/step) over numpy arrays so
that parity tests read like driving the reference env, one or many envs at a time.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

from cartpoleplusplus_b200 import config as C

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def oracle_lib_path() -> str:
    return os.path.join(_HERE, "libbcp_oracle.so")


def build_oracle(force: bool = False) -> str:
    """Compile the C restatement (gcc, -ffp-contract=off) if it is missing or stale."""
    so = oracle_lib_path()
    srcs = [os.path.join(_HERE, "bcp_oracle.c"), os.path.join(_HERE, "..", "include", "bcp.h")]
    stale = force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if stale:
        env = dict(os.environ)
        env.pop("CC", None)
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, env=env, capture_output=True)
    return so


def _lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build_oracle())
        for sfx in ("_f64", "_f32"):
            if getattr(_LIB, "bcpo_config_size" + sfx)() != ctypes.sizeof(C.BcpConfig):
                raise RuntimeError("bcp_config mirror out of date")
            getattr(_LIB, "bcpo_config_size" + sfx).restype = ctypes.c_size_t
    return _LIB


def _ptr(a, ctype):
    return a.ctypes.data_as(ctypes.POINTER(ctype)) if a is not None else None


def set_threads(n: int) -> int:
    """OpenMP threads of the oracle's batch loops (the library's own setting: OMP_NUM_THREADS of
    the launcher does not apply afterwards).  Returns the thread count in effect."""
    lib = _lib()
    lib.bcpo_set_threads_f32(int(n))
    return int(lib.bcpo_set_threads_f64(int(n)))


class Oracle:
    """Batched CPU oracle over canonical state records (n, STATE_DOUBLES) float64."""

    def __init__(self, cfg: C.BcpConfig, dtype: str = None):
        self.cfg = C.copy_config(cfg)
        if dtype is not None:
            self.cfg.dtype = C.DTYPE_F64 if dtype == "fp64" else C.DTYPE_F32
        self.sfx = "_f64" if self.cfg.dtype == C.DTYPE_F64 else "_f32"
        self.n = int(cfg.num_envs)
        self.R = int(cfg.action_repeats)
        self.state = np.zeros((self.n, C.STATE_DOUBLES), dtype=np.float64)
        self.state[:, C.S_CART + 6] = 1.0
        self.state[:, C.S_POLE + 6] = 1.0
        self.obs = np.zeros((self.n, self.R, 2, 7), dtype=np.float32)
        self.hist = np.zeros(cfg.max_episode_len + 1, dtype=np.int64)
        self.header = np.zeros(C.STATS_HEADER, dtype=np.int64)
        self.reward_sum = np.zeros(1, dtype=np.float64)

    def _f(self, name):
        return getattr(_lib(), name + self.sfx)

    # --- p.stepSimulation() alone ---
    def substep(self, with_info: bool = False):
        info = np.zeros((self.n, 6), dtype=np.int32) if with_info else None
        self._f("bcpo_substep")(ctypes.byref(self.cfg), _ptr(self.state, ctypes.c_double), self.n,
                                _ptr(info, ctypes.c_int32))
        return info

    def apply_force(self, fxy):
        fxy = np.ascontiguousarray(np.broadcast_to(np.asarray(fxy, dtype=np.float64), (self.n, 2)))
        self._f("bcpo_apply_force")(ctypes.byref(self.cfg), _ptr(self.state, ctypes.c_double), self.n,
                                    _ptr(fxy, ctypes.c_double))

    # --- env surface ---
    def reset(self, mask=None):
        m = None if mask is None else np.ascontiguousarray(np.asarray(mask, dtype=np.uint8))
        self._f("bcpo_reset")(ctypes.byref(self.cfg), _ptr(self.state, ctypes.c_double), self.n,
                              _ptr(m, ctypes.c_uint8), _ptr(self.obs, ctypes.c_float))
        return self.obs.copy()

    def step(self, actions, with_info: bool = False, force_iterations=None):
        """env.step of every env.  A config with auto_reset = pipelined runs the reset of a finished env inside the
        following calls (reason 4 RESETTING / 5 FRESH), like bcp_step.  with_info: also returns int32
        (n, R*S, 10) per substep: iterations (-1: no substep ran), rows, friction rows, contacts[3], SAT code[3], hash of the
        solver's clamp decisions.  force_iterations (n, R*S) int32: run exactly that many PGS iterations where > 0."""
        if self.cfg.discrete_actions:
            a = np.ascontiguousarray(np.broadcast_to(np.asarray(actions, dtype=np.int32), (self.n,)))
        else:
            a = np.ascontiguousarray(np.asarray(actions, dtype=np.float32).reshape(self.n, 2))
        reward = np.zeros(self.n, dtype=np.float32)
        done = np.zeros(self.n, dtype=np.uint8)
        reason = np.zeros(self.n, dtype=np.uint8)
        args = [ctypes.byref(self.cfg), _ptr(self.state, ctypes.c_double), self.n,
                a.ctypes.data_as(ctypes.c_void_p), _ptr(self.obs, ctypes.c_float),
                _ptr(reward, ctypes.c_float), _ptr(done, ctypes.c_uint8), _ptr(reason, ctypes.c_uint8)]
        if with_info or force_iterations is not None:
            rs = self.R * int(self.cfg.steps_per_repeat)
            info = np.zeros((self.n, rs, 10), dtype=np.int32)
            fi = None if force_iterations is None else np.ascontiguousarray(np.asarray(force_iterations, dtype=np.int32).reshape(self.n, rs))
            self._f("bcpo_step_info")(*args, _ptr(info, ctypes.c_int32), _ptr(fi, ctypes.c_int32))
            return self.obs.copy(), reward, done.astype(bool), reason, info
        self._f("bcpo_step")(*args)
        return self.obs.copy(), reward, done.astype(bool), reason

    def rollout_random(self, n_steps: int, action_mask: int = 0x1F, max_episodes: int = 0):
        """n_steps env.steps per env with synchronous resets in place (an env stops early once it
        has finished max_episodes episodes, if > 0)."""
        self._f("bcpo_rollout_random")(ctypes.byref(self.cfg), _ptr(self.state, ctypes.c_double), self.n,
                                       int(n_steps), ctypes.c_uint32(action_mask), int(max_episodes),
                                       _ptr(self.obs, ctypes.c_float),
                                       _ptr(self.hist, ctypes.c_int64), _ptr(self.header, ctypes.c_int64),
                                       _ptr(self.reward_sum, ctypes.c_double))

    def done_reason(self, env: int = 0) -> int:
        row = np.ascontiguousarray(self.state[env])
        return int(self._f("bcpo_done_reason")(ctypes.byref(self.cfg), _ptr(row, ctypes.c_double)))


# ---- free functions used by unit tests ----
def box_box(pA, qA, hA, pB, qB, hB, dtype="fp64"):
    """Bare narrowphase: returns (n, contacts[n,7]) with rows (normalOnB3, pointOnB3, distance)."""
    inp = np.concatenate([pA, qA, hA, pB, qB, hB]).astype(np.float64)
    out = np.zeros(28, dtype=np.float64)
    f = getattr(_lib(), "bcpo_box_box" + ("_f64" if dtype == "fp64" else "_f32"))
    n = f(_ptr(inp, ctypes.c_double), _ptr(out, ctypes.c_double))
    return n, out.reshape(4, 7)[:n].copy()


def elem(x, y=0.0, dtype="fp64"):
    """(sin_core(x), cos_core(x), det_atan2(y,x), sin(2 pi x), cos(2 pi x))."""
    out = np.zeros(5, dtype=np.float64)
    f = getattr(_lib(), "bcpo_elem" + ("_f64" if dtype == "fp64" else "_f32"))
    f.argtypes = [ctypes.c_double, ctypes.c_double, ctypes.POINTER(ctypes.c_double)]
    f(float(x), float(y), _ptr(out, ctypes.c_double))
    return out


def philox(seed, env, episode, purpose, index):
    out = np.zeros(4, dtype=np.uint32)
    f = _lib().bcpo_philox_f64
    f.argtypes = [ctypes.c_uint64, ctypes.c_int64, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32,
                  ctypes.POINTER(ctypes.c_uint32)]
    f(seed, env, episode, purpose, index, _ptr(out, ctypes.c_uint32))
    return out
