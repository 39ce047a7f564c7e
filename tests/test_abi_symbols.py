"""CPU tier: libbcp.so loads without a GPU and exports every entry point include/bcp.h
declares; the two calls that need no device answer what the header says."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "bcp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)   # comments mention entry points too
    text = re.sub(r"//[^\n]*", "", text)
    return sorted(set(re.findall(r"\b(bcp_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_entry_point_is_exported(built):
    from cartpoleplusplus_b200 import _lib
    from cartpoleplusplus_b200 import config as C
    names = _declared()
    assert {"bcp_create", "bcp_destroy", "bcp_reset", "bcp_step", "bcp_rollout_random", "bcp_stats", "bcp_get_state",
            "bcp_set_state", "bcp_substep", "bcp_last_error", "bcp_abi_version", "bcp_config_size"} <= set(names)
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
