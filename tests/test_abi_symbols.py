"""CPU tier: libbcp.so loads without a GPU and exports every entry point include/*.h
declare; the calls that need no device answer what the headers say."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    inc = os.path.join(ROOT, "include")
    text = "".join(open(os.path.join(inc, f)).read() for f in sorted(os.listdir(inc)) if f.endswith(".h"))
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)   # comments mention entry points too
    text = re.sub(r"//[^\n]*", "", text)
    return sorted(set(re.findall(r"\b(bcp_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_entry_point_is_exported(built):
    from cartpoleplusplus_b200 import _lib
    from cartpoleplusplus_b200 import config as C
    names = _declared()
    assert {"bcp_create", "bcp_destroy", "bcp_reset", "bcp_step", "bcp_rollout_random", "bcp_stats", "bcp_get_state",
            "bcp_set_state", "bcp_substep", "bcp_last_error", "bcp_abi_version", "bcp_config_size",
            "bcp_replay_create", "bcp_replay_destroy", "bcp_replay_begin", "bcp_replay_observe", "bcp_replay_add_episode",
            "bcp_replay_batch", "bcp_replay_info", "bcp_replay_export"} <= set(names)
    assert set(names) == set(_lib.EXPORTS)
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    header = open(os.path.join(ROOT, "include", "bcp.h")).read()
    version = int(re.search(r"#define\s+BCP_ABI_VERSION\s+(\d+)", header).group(1))
    assert lib.bcp_abi_version() == version == C.BCP_ABI_VERSION
    lib.bcp_config_size.restype = ctypes.c_size_t
    assert lib.bcp_config_size() == ctypes.sizeof(C.BcpConfig)


def test_null_handle_is_an_error_not_a_crash(built):
    from cartpoleplusplus_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    lib.bcp_last_error.restype = ctypes.c_char_p
    assert lib.bcp_step(None, None, None, None, None, None, None) < 0
    assert b"null" in lib.bcp_last_error()
    assert lib.bcp_reset(None, None, None, None) < 0
    assert lib.bcp_replay_observe(None, None, 0, None, None, None, None, None) < 0
    assert b"null" in lib.bcp_last_error()


def test_replay_config_mirror_and_validation(built):
    from cartpoleplusplus_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    lib.bcp_last_error.restype = ctypes.c_char_p
    lib.bcp_replay_config_size.restype = ctypes.c_size_t
    assert lib.bcp_replay_config_size() == ctypes.sizeof(_lib.BcpReplayConfig)
    h = ctypes.c_void_p()
    bad = _lib.BcpReplayConfig(_lib.REPLAY_ABI_VERSION, 0, 100, 120, 28, 2, 0, 0)  # load_factor 1.2 < 1.5 (replay_memory.py:13)
    assert lib.bcp_replay_create(ctypes.byref(bad), ctypes.byref(h)) < 0
    assert b"load_factor" in lib.bcp_last_error()
