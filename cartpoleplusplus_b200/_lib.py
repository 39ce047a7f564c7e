"""ctypes binding of libbcp.so (include/bcp.h).  Fails loudly: there is no fallback."""
from __future__ import annotations

import ctypes
import os

from . import config as C

_LIB = None
# BCP_LIB_PATH: a measurement build of the same sources (build.py --precise), never a different implementation
LIB_PATH = os.environ.get("BCP_LIB_PATH") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libbcp.so")

EXPORTS = ["bcp_abi_version", "bcp_config_size", "bcp_last_error", "bcp_create", "bcp_destroy", "bcp_reset",
           "bcp_step", "bcp_rollout_random", "bcp_stats", "bcp_stats_clear", "bcp_get_state", "bcp_set_state",
           "bcp_substep", "bcp_launch_count", "bcp_bench_fma_peak", "bcp_profile_begin", "bcp_profile_end", "bcp_last_solve_info", "bcp_solver_kind", "bcp_step_host", "bcp_step_host_finish", "bcp_solve_trace_enable", "bcp_solve_trace_read",
           # include/bcp_replay.h
           "bcp_replay_config_size", "bcp_replay_create", "bcp_replay_destroy", "bcp_replay_begin", "bcp_replay_observe",
           "bcp_replay_add_episode", "bcp_replay_batch", "bcp_replay_info", "bcp_replay_export", "bcp_replay_launch_count",
           # include/bcp_lqr.h
           "bcp_lqr_args_size", "bcp_lqr_control",
           # include/bcp_render.h
           "bcp_render_args_size", "bcp_render_pixels", "bcp_render_static_scratch_bytes", "bcp_render_pixels_static"]

REPLAY_ABI_VERSION = 1
REPLAY_INFO_WORDS = 10
REPLAY_ERR_UNDERFLOW, REPLAY_ERR_STAGING, REPLAY_ERR_INDEX, REPLAY_ERR_ROUNDS = 1, 2, 4, 8


class BcpReplayConfig(ctypes.Structure):
    """mirror of bcp_replay_config (include/bcp_replay.h)"""
    _fields_ = [("abi_version", ctypes.c_int32), ("device_ordinal", ctypes.c_int32), ("buffer_size", ctypes.c_int64),
                ("state_buffer_size", ctypes.c_int64), ("state_words", ctypes.c_int32), ("action_dim", ctypes.c_int32),
                ("num_envs", ctypes.c_int32), ("max_episode_len", ctypes.c_int32)]


class BcpLqrArgs(ctypes.Structure):
    """mirror of bcp_lqr_args (include/bcp_lqr.h)"""
    _fields_ = [("gains", ctypes.c_void_p), ("gains_per_env", ctypes.c_int32), ("zero_point", ctypes.c_void_p),
                ("x0", ctypes.c_void_p), ("disturbance", ctypes.c_void_p), ("disturbance_per_env", ctypes.c_int32),
                ("position_threshold", ctypes.c_float), ("angle_threshold", ctypes.c_float),
                ("state_out", ctypes.c_void_p), ("action_out", ctypes.c_void_p), ("done_out", ctypes.c_void_p)]


class BcpRenderArgs(ctypes.Structure):
    """mirror of bcp_render_args (include/bcp_render.h)"""
    _f3 = ctypes.c_float * 3
    _fields_ = [("num_envs", ctypes.c_int32), ("repeats", ctypes.c_int32), ("num_cameras", ctypes.c_int32),
                ("width", ctypes.c_int32), ("height", ctypes.c_int32),
                ("half_extents", (ctypes.c_float * 3) * 3), ("rgb", (ctypes.c_float * 3) * 3),
                ("ground_pos", ctypes.c_float * 3), ("camera_pos", (ctypes.c_float * 3) * 2),
                ("target", ctypes.c_float * 3), ("up", ctypes.c_float * 3),
                ("near_val", ctypes.c_float), ("far_val", ctypes.c_float), ("fov_deg", ctypes.c_float), ("clip", ctypes.c_int32),
                ("light_dir", ctypes.c_float * 3), ("ambient", ctypes.c_float), ("diffuse", ctypes.c_float),
                ("background", ctypes.c_float * 3), ("obs", ctypes.c_void_p), ("pixels", ctypes.c_void_p)]


class BcpError(RuntimeError):
    pass


def load(build_if_missing: bool = False):
    """Load the CUDA C-ABI library; raises if it is absent or its ABI does not match."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        if not build_if_missing:
            raise BcpError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(nvcc, sm_100a).  There is no CPU fallback for the env step.")
        from . import build as _build
        _build.build()
    lib = ctypes.CDLL(LIB_PATH)
    vp, i32, u32, u64p = ctypes.c_void_p, ctypes.c_int32, ctypes.c_uint32, ctypes.c_void_p
    lib.bcp_abi_version.restype = ctypes.c_int
    lib.bcp_config_size.restype = ctypes.c_size_t
    lib.bcp_last_error.restype = ctypes.c_char_p
    lib.bcp_create.argtypes = [ctypes.POINTER(C.BcpConfig), ctypes.POINTER(vp)]
    lib.bcp_destroy.argtypes = [vp]
    lib.bcp_reset.argtypes = [vp, vp, vp, vp]
    lib.bcp_step.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.bcp_rollout_random.argtypes = [vp, i32, u32, i32, vp, vp]
    lib.bcp_stats.argtypes = [vp, u64p, u64p, vp, vp]
    lib.bcp_stats_clear.argtypes = [vp, vp]
    lib.bcp_get_state.argtypes = [vp, vp, ctypes.c_size_t, vp]
    lib.bcp_set_state.argtypes = [vp, vp, ctypes.c_size_t, vp]
    lib.bcp_substep.argtypes = [vp, vp]
    lib.bcp_bench_fma_peak.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
    lib.bcp_launch_count.argtypes = [vp]
    lib.bcp_profile_begin.argtypes = [vp]
    lib.bcp_profile_end.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
    lib.bcp_last_solve_info.argtypes = [vp, vp, ctypes.c_size_t, vp]
    lib.bcp_launch_count.restype = ctypes.c_int64
    lib.bcp_solve_trace_enable.argtypes = [vp, ctypes.c_int]
    lib.bcp_solve_trace_read.argtypes = [vp, vp, ctypes.c_size_t, vp]
    lib.bcp_solver_kind.argtypes = [vp]
    lib.bcp_step_host.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.bcp_step_host_finish.argtypes = [vp, vp, vp, vp]
    i64, u64 = ctypes.c_int64, ctypes.c_uint64
    lib.bcp_replay_config_size.restype = ctypes.c_size_t
    lib.bcp_replay_create.argtypes = [ctypes.POINTER(BcpReplayConfig), ctypes.POINTER(vp)]
    lib.bcp_replay_destroy.argtypes = [vp]
    lib.bcp_replay_begin.argtypes = [vp, vp, vp, vp]
    lib.bcp_replay_observe.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp]
    lib.bcp_replay_add_episode.argtypes = [vp, vp, vp, vp, vp, i32, vp]
    lib.bcp_replay_batch.argtypes = [vp, vp, i32, u64, u64, vp, vp, vp, vp, vp, vp, vp]
    lib.bcp_replay_info.argtypes = [vp, vp, vp]
    lib.bcp_replay_export.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.bcp_replay_launch_count.argtypes = [vp]
    lib.bcp_replay_launch_count.restype = i64
    lib.bcp_lqr_args_size.restype = ctypes.c_size_t
    lib.bcp_render_args_size.restype = ctypes.c_size_t
    lib.bcp_render_pixels.argtypes = [ctypes.POINTER(BcpRenderArgs), ctypes.c_int, vp]
    lib.bcp_render_static_scratch_bytes.restype = ctypes.c_size_t
    lib.bcp_render_pixels_static.argtypes = [ctypes.POINTER(BcpRenderArgs), ctypes.c_int, vp]
    lib.bcp_lqr_control.argtypes = [vp, ctypes.POINTER(BcpLqrArgs), vp]
    if lib.bcp_abi_version() != C.BCP_ABI_VERSION:
        raise BcpError(f"libbcp.so ABI {lib.bcp_abi_version()} != python mirror {C.BCP_ABI_VERSION}: rebuild")
    if lib.bcp_config_size() != ctypes.sizeof(C.BcpConfig):
        raise BcpError("sizeof(bcp_config) differs between libbcp.so and the ctypes mirror: rebuild")
    if lib.bcp_replay_config_size() != ctypes.sizeof(BcpReplayConfig):
        raise BcpError("sizeof(bcp_replay_config) differs between libbcp.so and the ctypes mirror: rebuild")
    if lib.bcp_lqr_args_size() != ctypes.sizeof(BcpLqrArgs):
        raise BcpError("sizeof(bcp_lqr_args) differs between libbcp.so and the ctypes mirror: rebuild")
    if lib.bcp_render_args_size() != ctypes.sizeof(BcpRenderArgs):
        raise BcpError("sizeof(bcp_render_args) differs between libbcp.so and the ctypes mirror: rebuild")
    _LIB = lib
    return lib


def check(rc: int, lib=None):
    if rc != 0:
        lib = lib or load()
        raise BcpError((lib.bcp_last_error() or b"unknown error").decode())
