"""In-tree build of libbcp.so (the C-ABI library of include/bcp.h) with nvcc for sm_100a.

Three translation units: bcp_f32.cu (fast mode, FMA on), bcp_f64.cu (parity mode,
-fmad=false so the FP64 kernel rounds exactly like the CPU oracle) and bcp_abi.cu.
nvcc cross-compiles without a GPU; the resulting .so is git-ignored but travels to the
GPU box with the tree.
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys

CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")
LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libbcp.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]
UNITS = [
    ("bcp_f32.cu", ["-prec-div=false", "-prec-sqrt=false", "-ftz=true"]),
    ("bcp_f64.cu", ["-fmad=false"]),
    ("bcp_abi.cu", []),
    ("bcp_replay.cu", []),
    ("bcp_render.cu", ["-prec-div=false"]),
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _sources():
    inc = os.path.join(CSRC, "..", "..", "include")
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))] + \
        [os.path.join(inc, f) for f in os.listdir(inc) if f.endswith(".h")]


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in _sources())


def build(force: bool = False, verbose: bool = False, precise: bool = False) -> str:
    """precise=True builds libbcp_precise.so: the fp32 unit WITHOUT -prec-div=false -prec-sqrt=false -ftz=true, to
    measure how much of the fp32-vs-oracle tail those flags are (scripts/fp32_tail_report.py; never the product)."""
    lib = LIB.replace("libbcp.so", "libbcp_precise.so") if precise else LIB
    if not force and os.path.exists(lib) and not any(os.path.getmtime(s) > os.path.getmtime(lib) for s in _sources()):
        return lib
    nvcc = _nvcc()
    objdir = os.path.join(CSRC, "build_precise" if precise else "build")
    os.makedirs(objdir, exist_ok=True)
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)

    def compile_one(unit):
        src, extra = unit
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        tune = os.environ.get("BCP_NVCC_EXTRA", "").split()  # tuning experiments (-D...), empty in the product build
        if precise and src == "bcp_f32.cu":
            extra = []
        cmd = [nvcc, *ARCH, *COMMON, *extra, *tune, "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True, env=env)
        with open(obj + ".ptxas.log", "w") as fh:
            fh.write(r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stderr[-4000:]}")
        return obj

    with concurrent.futures.ThreadPoolExecutor(max_workers=len(UNITS)) as ex:
        objs = list(ex.map(compile_one, UNITS))
    cmd = [nvcc, *ARCH, "-shared", "-o", lib, *objs, "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stderr[-4000:]}")
    if verbose:
        print(f"built {lib}", file=sys.stderr)
    return lib


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True, precise="--precise" in sys.argv)
