"""The raw-pixel observation (`--use-raw-pixels`, bullet_cartpole.py:277-306) for a batch of envs.

`PixelRenderer.render(obs)` turns the (B, R, 2, 7) cart / pole poses of a call into the reference's
(H, W, 3, C, R) float16 image block per env - (B, H, W, 3, C, R) - with one kernel launch through
include/bcp_render.h.  The cameras, target, clip planes and field of view are the reference's
(:278-284); the boxes and their colours come from the URDFs.  pybullet's software renderer is not
available here, so the shading model and its constants are this module's: the
picture has the reference's geometry, not provably its exact pixel values.  There is no CPU path.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib
from . import urdf as _urdf

CAMERAS = ((0.0, 0.75, 0.75), (0.75, 0.0, 0.75))  # :278-280
TARGET, UP = (0.0, 0.0, 0.3), (0.0, 0.0, 1.0)     # :281-282
NEAR, FAR, FOV = 1.0, 20.0, 30.0                  # :283-284
LIGHT_DIR = (0.4, 0.3, 0.866)   # towards the light; a choice, like the three below
AMBIENT, DIFFUSE = 0.6, 0.35
BACKGROUND = (1.0, 1.0, 1.0)


def render_args(scene=None, num_envs=1, repeats=2, num_cameras=1, width=50, height=50, light_dir=LIGHT_DIR,
                ambient=AMBIENT, diffuse=DIFFUSE, background=BACKGROUND, clip=False) -> "_lib.BcpRenderArgs":
    """The scene / camera part of bcp_render_args (pointers left NULL)."""
    ground, cart, pole = scene if scene is not None else _urdf.load_scene(None)
    a = _lib.BcpRenderArgs()
    a.num_envs, a.repeats, a.num_cameras, a.width, a.height = int(num_envs), int(repeats), int(num_cameras), int(width), int(height)
    for b, body in enumerate((ground, cart, pole)):
        for i in range(3):
            a.half_extents[b][i] = body.half_extents[i]
            a.rgb[b][i] = body.rgba[i]
    for i in range(3):
        a.ground_pos[i] = 0.0  # loaded at the origin, bullet_cartpole.py:151
        a.target[i], a.up[i] = TARGET[i], UP[i]
        a.light_dir[i], a.background[i] = light_dir[i], background[i]
        for c in range(2):
            a.camera_pos[c][i] = CAMERAS[c][i]
    a.near_val, a.far_val, a.fov_deg = NEAR, FAR, FOV
    a.clip = 1 if clip else 0  # pybullet's software rasteriser does not clip (include/bcp_render.h)
    a.ambient, a.diffuse = ambient, diffuse
    return a


class PixelRenderer:
    def __init__(self, num_envs: int, repeats: int = 2, num_cameras: int = 1, width: int = 50, height: int = 50,
                 scene=None, device=0, **shading):
        if not torch.cuda.is_available():
            raise _lib.BcpError("PixelRenderer needs a CUDA device: there is no CPU path")
        if num_cameras not in (1, 2):
            raise ValueError("--num-cameras must be 1 or 2")  # bullet_cartpole.py:113-114
        self._lib = _lib.load()
        self.device = torch.device("cuda", int(device) if not isinstance(device, torch.device) else device.index or 0)
        self.num_envs, self.repeats, self.num_cameras = int(num_envs), int(repeats), int(num_cameras)
        self.width, self.height = int(width), int(height)
        self.args = render_args(scene, num_envs, repeats, num_cameras, width, height, **shading)
        # (H, W, 3, C, R) per env, bullet_cartpole.py:96-101
        self.state_shape = (self.height, self.width, 3, self.num_cameras, self.repeats)
        self.pixels = torch.zeros((self.num_envs,) + self.state_shape, dtype=torch.float16, device=self.device)
        self.launches = 0

    def render(self, obs: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """obs (B, R, 2, 7) float32 on the device -> (B, H, W, 3, C, R) float16 (the renderer's own buffer unless `out`)."""
        o = obs if (type(obs) is torch.Tensor and obs.is_cuda and obs.dtype == torch.float32 and obs.is_contiguous()) \
            else torch.as_tensor(obs, dtype=torch.float32, device=self.device).contiguous()
        if o.numel() != self.num_envs * self.repeats * 14:
            raise ValueError(f"obs must be ({self.num_envs}, {self.repeats}, 2, 7), got {tuple(o.shape)}")
        dst = self.pixels if out is None else out
        if dst.dtype != torch.float16 or not dst.is_cuda or not dst.is_contiguous() or dst.numel() != self.pixels.numel():
            raise ValueError("out must be a contiguous CUDA float16 tensor of the renderer's shape")
        self.args.obs = ctypes.c_void_p(o.data_ptr())
        self.args.pixels = ctypes.c_void_p(dst.data_ptr())
        st = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(self._lib.bcp_render_pixels(ctypes.byref(self.args), self.device.index, st), self._lib)
        self.launches += 1
        return dst
