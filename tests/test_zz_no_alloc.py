"""bcp_render_pixels_static (include/bcp_render.h): the raw-pixel kernel pair with its scratch in a static
device array, for callers that may not allocate.  It has no counterpart in the reference; the checker here is
the existing entry point bcp_render_pixels (itself held to oracle/render_oracle.py by tests/test_gpu_render.py),
and the bar is bit equality: same kernels, same arguments, same scratch layout and alignment.

Collected last on purpose; passes alone in a fresh process and after the others.
"""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _poses(rng, n, R):
    obs = np.zeros((n, R, 2, 7), np.float32)
    cart = rng.uniform(-0.25, 0.25, size=(n, R, 2))
    obs[:, :, 0, :2] = cart
    obs[:, :, 0, 2] = 0.075
    obs[:, :, 0, 6] = 1.0
    q = np.concatenate([rng.normal(scale=0.25, size=(n, R, 2)), rng.normal(scale=0.05, size=(n, R, 1)), np.ones((n, R, 1))], axis=-1)
    obs[:, :, 1, :2] = cart + rng.normal(scale=0.02, size=(n, R, 2))
    obs[:, :, 1, 2] = 0.35
    obs[:, :, 1, 3:] = q / np.linalg.norm(q, axis=-1, keepdims=True)
    return obs


def _scratch_need(C, W, H, R):
    nt = C * W * H
    return (nt * 20 + 15) // 16 * 16 + nt * 3 * R * 2


def test_static_entry_point_has_no_allocating_call():
    """the text of the new entry point and of its set-up helper names no allocating / freeing routine"""
    src = open(os.path.join(ROOT, "cartpoleplusplus_b200", "csrc", "bcp_render.cu")).read()
    tail = src[src.index("bcp_render_pixels_static: the same two kernels"):]
    code = re.sub(r"//[^\n]*", "", tail)
    assert "bcp_render_pixels_static(const bcp_render_args*" in code and "cudaGetSymbolAddress" in code
    for banned in ("cudaMalloc", "cudaFree", "MemPool", "cudaHostAlloc", "malloc(", "new ", "std::vector"):
        assert banned not in code, banned
    assert "__device__ __align__(256) unsigned char g_render_scratch[BCP_RENDER_STATIC_SCRATCH_BYTES]" in code


@pytest.mark.gpu
def test_static_scratch_render_equals_pool_scratch_render_twice(built):
    torch = pytest.importorskip("torch")
    from cartpoleplusplus_b200 import _lib
    from cartpoleplusplus_b200.render import render_args
    lib = _lib.load()
    cap = lib.bcp_render_static_scratch_bytes()
    assert cap == 1 << 20
    dev = torch.device("cuda", 0)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    rng = np.random.default_rng(2024)
    # two calls in a row that lay the shared array out differently, then the first shape again over the second's leftovers
    shapes = [(96, 2, 2, 50, 50), (40, 3, 1, 64, 40), (96, 2, 2, 50, 50)]
    jobs = []
    for n, R, C, W, H in shapes:  # every tensor exists before the first call: the calls themselves get nothing from torch
        assert _scratch_need(C, W, H, R) <= cap
        obs = torch.from_numpy(_poses(rng, n, R)).to(dev)
        want = torch.zeros((n, H, W, 3, C, R), dtype=torch.float16, device=dev)
        got = torch.full((n, H, W, 3, C, R), -1.0, dtype=torch.float16, device=dev)
        a = render_args(num_envs=n, repeats=R, num_cameras=C, width=W, height=H)
        a.obs = ctypes.c_void_p(obs.data_ptr())
        jobs.append((a, obs, want, got))
    for a, obs, want, got in jobs:
        a.pixels = ctypes.c_void_p(want.data_ptr())
        _lib.check(lib.bcp_render_pixels(ctypes.byref(a), 0, stream), lib)
    torch.cuda.synchronize()
    for a, obs, want, got in jobs:  # back to back on one stream, no synchronisation in between
        a.pixels = ctypes.c_void_p(got.data_ptr())
        _lib.check(lib.bcp_render_pixels_static(ctypes.byref(a), 0, stream), lib)
    torch.cuda.synchronize()
    for k, (a, obs, want, got) in enumerate(jobs):
        w = want.cpu().numpy()
        assert float(w.min()) >= 0.0 and float(w.max()) <= 1.0 and len(np.unique(w)) > 8, k  # a picture, not a blank
        assert torch.equal(got, want), (k, int((got != want).sum()))
    assert not torch.equal(jobs[0][2], jobs[2][2])  # the first and third calls drew different poses


@pytest.mark.gpu
def test_static_scratch_limit_and_argument_errors(built):
    torch = pytest.importorskip("torch")
    from cartpoleplusplus_b200 import _lib
    from cartpoleplusplus_b200.render import render_args
    lib = _lib.load()
    cap = lib.bcp_render_static_scratch_bytes()
    n = 2
    obs = torch.zeros((n, 2, 2, 7), device="cuda")
    obs[..., 6] = 1
    # exactly at the limit: 128 x 128, two cameras, two repeats = 1 048 576 bytes
    assert _scratch_need(2, 128, 128, 2) == cap
    a = render_args(num_envs=n, repeats=2, num_cameras=2, width=128, height=128)
    px = torch.zeros((n, 128, 128, 3, 2, 2), dtype=torch.float16, device="cuda")
    ref = torch.zeros_like(px)
    a.obs = ctypes.c_void_p(obs.data_ptr())
    a.pixels = ctypes.c_void_p(px.data_ptr())
    _lib.check(lib.bcp_render_pixels_static(ctypes.byref(a), 0, None), lib)
    a.pixels = ctypes.c_void_p(ref.data_ptr())
    _lib.check(lib.bcp_render_pixels(ctypes.byref(a), 0, None), lib)
    torch.cuda.synchronize()
    assert torch.equal(px, ref)
    # one column more: refused, nothing launched, the message names both sizes
    b = render_args(num_envs=n, repeats=2, num_cameras=2, width=129, height=128)
    big = torch.full((n, 128, 129, 3, 2, 2), 0.25, dtype=torch.float16, device="cuda")
    b.obs = ctypes.c_void_p(obs.data_ptr())
    b.pixels = ctypes.c_void_p(big.data_ptr())
    assert lib.bcp_render_pixels_static(ctypes.byref(b), 0, None) == -1
    msg = lib.bcp_last_error().decode()
    assert "static array holds 1048576" in msg and str(_scratch_need(2, 129, 128, 2)) in msg, msg
    torch.cuda.synchronize()
    assert bool((big == 0.25).all())
    _lib.check(lib.bcp_render_pixels(ctypes.byref(b), 0, None), lib)  # the pool-backed entry point takes that size
    torch.cuda.synchronize()
    assert not bool((big == 0.25).all())
    # the other argument checks mirror bcp_render_pixels
    c = render_args(num_envs=n)
    assert lib.bcp_render_pixels_static(ctypes.byref(c), 0, None) == -1 and b"null" in lib.bcp_last_error()
    c.obs = ctypes.c_void_p(obs.data_ptr())
    c.pixels = ctypes.c_void_p(px.data_ptr())
    c.repeats = 9
    assert lib.bcp_render_pixels_static(ctypes.byref(c), 0, None) == -1 and b"repeats" in lib.bcp_last_error()
    c.repeats = 2
    assert lib.bcp_render_pixels_static(ctypes.byref(c), 99, None) == -1
