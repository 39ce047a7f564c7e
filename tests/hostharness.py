"""ctypes front-end of tests/host_harness (the KERNEL headers compiled by g++ for the host).

TEST TOOL ONLY.  It lets the CPU tier run the exact per-env programs of the CUDA kernels
(parity path and fast path) against the C oracle without a GPU; nothing in the package can
reach it."""
import ctypes
import os
import subprocess

import numpy as np

from cartpoleplusplus_b200 import config as C

_HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_harness")
_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "libhh.so")
    csrc = os.path.join(_ROOT, "cartpoleplusplus_b200", "csrc")
    srcs = [os.path.join(_HERE, "harness.cpp"), os.path.join(_ROOT, "include", "bcp.h")]
    srcs += [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".cuh")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-o", so,
                        os.path.join(_HERE, "harness.cpp")], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.hh_config_size.restype = ctypes.c_size_t
        assert _LIB.hh_config_size() == ctypes.sizeof(C.BcpConfig)
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(ctypes.POINTER(t)) if a is not None else None


class HostKernels:
    """Same surface as oracle.Oracle, running the kernel headers on the host."""

    def __init__(self, cfg, dtype, fast):
        self.cfg = C.copy_config(cfg)
        self.sfx = "_f64" if dtype == "fp64" else "_f32"
        self.fast = int(fast)  # 0 parity path, 1 fast path, 2 split pipeline pieces
        self.n = int(cfg.num_envs)
        self.R = int(cfg.action_repeats)
        self.state = np.zeros((self.n, C.STATE_DOUBLES), dtype=np.float64)
        self.state[:, C.S_CART + 6] = 1.0
        self.state[:, C.S_POLE + 6] = 1.0
        self.obs = np.zeros((self.n, self.R, 2, 7), dtype=np.float32)

    def _f(self, name):
        lib().hh_set_fast(self.fast)
        return getattr(lib(), name + self.sfx)

    def substep(self, with_info=False):
        info = np.zeros((self.n, 6), dtype=np.int32) if with_info else None
        self._f("hh_substep")(ctypes.byref(self.cfg), _p(self.state, ctypes.c_double), self.n, _p(info, ctypes.c_int32))
        return info

    def reset(self, mask=None):
        m = None if mask is None else np.ascontiguousarray(np.asarray(mask, dtype=np.uint8))
        self._f("hh_reset")(ctypes.byref(self.cfg), _p(self.state, ctypes.c_double), self.n, _p(m, ctypes.c_uint8),
                            _p(self.obs, ctypes.c_float))
        return self.obs.copy()

    def step(self, actions):
        if self.cfg.discrete_actions:
            a = np.ascontiguousarray(np.broadcast_to(np.asarray(actions, dtype=np.int32), (self.n,)))
        else:
            a = np.ascontiguousarray(np.asarray(actions, dtype=np.float32).reshape(self.n, 2))
        reward = np.zeros(self.n, dtype=np.float32)
        done = np.zeros(self.n, dtype=np.uint8)
        reason = np.zeros(self.n, dtype=np.uint8)
        self._f("hh_step")(ctypes.byref(self.cfg), _p(self.state, ctypes.c_double), self.n,
                           a.ctypes.data_as(ctypes.c_void_p), _p(self.obs, ctypes.c_float),
                           _p(reward, ctypes.c_float), _p(done, ctypes.c_uint8), _p(reason, ctypes.c_uint8))
        return self.obs.copy(), reward, done.astype(bool), reason
