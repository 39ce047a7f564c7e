"""CPU restatement of the raw-pixel observation (bullet_cartpole.py:277-306) - numpy, one env at a time.

TEST INFRASTRUCTURE ONLY: imported by tests/ (and nothing in the product path).

PARITY UNPINNED against the reference's pixels: they come out of pybullet's software renderer
(`p.renderImage`), which is unavailable on this box and absent from the reference tree.  What is
restated here, and what the CUDA kernel is checked against, is
  * pybullet's camera model as published: computeViewMatrix (forward = target - eye, side = forward x
    up, up' = side x forward) and computeProjectionMatrixFOV (y scale 1 / tan(fov / 2), x scale = y
    scale / aspect, aspect = width / height), used as a per-pixel ray through the pixel centre;
  * the three boxes of the scene (ground at the origin, cart and pole at the observed poses), nearest
    hit (between the near and far planes only if `clip`: the reference's near plane, 1 m, lies behind its
    cart and pole, and pybullet's software rasteriser does not clip against it);
  * flat Lambert shading (ambient + diffuse * max(0, n.l)), uint8 truncation, float16(u8 / 255)
    exactly as :289-294 converts the renderer's bytes.
The geometry is pinned by closed-form cases in tests/test_render_oracle.py (projection of known
points, silhouette of an axis-aligned box).
"""
from __future__ import annotations

import numpy as np

f32 = np.float32


def camera_basis(eye, target, up):
    eye, target, up = (np.asarray(v, dtype=f32) for v in (eye, target, up))
    f = target - eye
    f = f / np.linalg.norm(f)
    up = up / np.linalg.norm(up)
    s = np.cross(f, up)
    s = s / np.linalg.norm(s)
    u = np.cross(s, f)
    return eye, s.astype(f32), u.astype(f32), f.astype(f32)


def pixel_rays(width, height, fov_deg, s, u, f):
    """-> (H, W, 3) float32 ray directions whose component along f is 1 (the ray parameter is depth)."""
    ty = f32(np.tan(f32(0.5) * f32(fov_deg) * f32(np.pi) / f32(180.0)))
    tx = f32(ty * f32(width) / f32(height))
    px = np.arange(width, dtype=f32)
    py = np.arange(height, dtype=f32)
    nx = (f32(2.0) * (px + f32(0.5))) / f32(width) - f32(1.0)
    ny = f32(1.0) - (f32(2.0) * (py + f32(0.5))) / f32(height)
    dx, dy = (nx * tx)[None, :, None], (ny * ty)[:, None, None]
    return (s[None, None, :] * dx + u[None, None, :] * dy + f[None, None, :]).astype(f32)


def quat_axes(q):
    x, y, z, w = (f32(v) for v in q)
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y + w * z), 2 * (x * z - w * y)],
                     [2 * (x * y - w * z), 1 - 2 * (x * x + z * z), 2 * (y * z + w * x)],
                     [2 * (x * z + w * y), 2 * (y * z - w * x), 1 - 2 * (x * x + y * y)]], dtype=f32)  # rows = axes


def ray_box(o, d, centre, axes, h, tlo, thi):
    """o (3,), d (H, W, 3) -> entry depth (H, W) (inf where missed) and entry normal (H, W, 3)."""
    rel = (np.asarray(o, f32) - np.asarray(centre, f32)).astype(f32)
    t0 = np.full(d.shape[:2], f32(tlo), dtype=f32)
    t1 = np.full(d.shape[:2], f32(thi), dtype=f32)
    axis = np.full(d.shape[:2], -1, dtype=np.int32)
    sign = np.zeros(d.shape[:2], dtype=f32)
    miss = np.zeros(d.shape[:2], dtype=bool)
    for a in range(3):
        e = axes[a]
        lo = f32(rel @ e)
        ld = (d @ e).astype(f32)
        par = np.abs(ld) < f32(1e-12)
        miss |= par & (abs(lo) > f32(h[a]))
        with np.errstate(divide="ignore", invalid="ignore"):
            inv = f32(1.0) / ld
            ta, tb = (f32(-h[a]) - lo) * inv, (f32(h[a]) - lo) * inv
        swap = ta > tb
        ta, tb = np.where(swap, tb, ta), np.where(swap, ta, tb)
        sg = np.where(swap, f32(1.0), f32(-1.0))
        live = ~par & ~miss
        upd = live & (ta > t0)
        t0 = np.where(upd, ta, t0)
        axis = np.where(upd, a, axis)
        sign = np.where(upd, sg, sign)
        t1 = np.where(live & (tb < t1), tb, t1)
        miss |= live & (t0 > t1)
    hit = ~miss & (axis >= 0)
    n = sign[..., None] * axes[np.clip(axis, 0, 2)]
    return np.where(hit, t0, np.inf).astype(f32), n.astype(f32)


def render_env(obs_env, half_extents, rgb, cameras, target, up, near, far, fov_deg, width, height, light_dir, ambient,
               diffuse, background, ground_pos=(0.0, 0.0, 0.0), clip=False):
    """obs_env (R, 2, 7) -> (H, W, 3, C, R) float16: the reference's state block of one env."""
    R, C = obs_env.shape[0], len(cameras)
    out = np.zeros((height, width, 3, C, R), dtype=np.float16)
    light = np.asarray(light_dir, f32)
    light = light / np.linalg.norm(light)
    eye3 = np.eye(3, dtype=f32)
    if not clip:  # pybullet's software rasteriser z-tests only: nothing is dropped at the near / far planes
        near, far = 1e-6, 1e30
    for c in range(C):
        eye, s, u, f = camera_basis(cameras[c], target, up)
        d = pixel_rays(width, height, fov_deg, s, u, f)
        tg, ng = ray_box(eye, d, ground_pos, eye3, half_extents[0], near, far)
        for r in range(R):
            best = np.where(np.isfinite(tg), tg, f32(far))
            body = np.where(np.isfinite(tg), 0, -1)
            n = ng.copy()
            for b in (1, 2):
                pose = obs_env[r, b - 1]
                t, nb = ray_box(eye, d, pose[:3], quat_axes(pose[3:7]), half_extents[b], near, far)
                nearer = t < best
                best = np.where(nearer, t, best)
                body = np.where(nearer, b, body)
                n = np.where(nearer[..., None], nb, n)
            ndl = np.maximum(f32(0.0), (n @ light).astype(f32))
            shade = (f32(ambient) + f32(diffuse) * ndl).astype(f32)
            col = np.asarray(rgb, f32)[np.clip(body, 0, 2)] * shade[..., None]
            col = np.where((body < 0)[..., None], np.asarray(background, f32)[None, None, :], col).astype(f32)
            u8 = np.clip((col * f32(255.0)).astype(np.int32), 0, 255)
            out[:, :, :, c, r] = (u8.astype(f32) / f32(255.0)).astype(np.float16)
    return out
