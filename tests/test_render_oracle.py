"""oracle/render_oracle.py pinned without a GPU: pybullet's published camera matrices (computeViewMatrix,
computeProjectionMatrixFOV) written out independently in float64, closed-form silhouettes, the
reference's uint8 -> float16 / 255 conversion (bullet_cartpole.py:289-294)."""
import numpy as np
import pytest

from oracle import render_oracle as RO

CAMS = ((0.0, 0.75, 0.75), (0.75, 0.0, 0.75))
TARGET, UP = (0.0, 0.0, 0.3), (0.0, 0.0, 1.0)


def _view_proj(eye, target, up, fov, aspect, near, far):
    """pybullet's column-major matrices, as row-major numpy (float64)."""
    eye, target, up = (np.asarray(v, float) for v in (eye, target, up))
    f = (target - eye) / np.linalg.norm(target - eye)
    u = up / np.linalg.norm(up)
    s = np.cross(f, u)
    s /= np.linalg.norm(s)
    u = np.cross(s, f)
    V = np.eye(4)
    V[0, :3], V[1, :3], V[2, :3] = s, u, -f
    V[:3, 3] = -V[:3, :3] @ eye
    ys = 1.0 / np.tan(np.radians(fov) / 2)
    P = np.zeros((4, 4))
    P[0, 0], P[1, 1] = ys / aspect, ys
    P[2, 2], P[2, 3] = (near + far) / (near - far), 2 * near * far / (near - far)
    P[3, 2] = -1.0
    return V, P


@pytest.mark.parametrize("cam", [0, 1])
@pytest.mark.parametrize("wh", [(50, 50), (64, 40)])
def test_rays_invert_pybullets_projection(cam, wh):
    W, H = wh
    V, P = _view_proj(CAMS[cam], TARGET, UP, 30.0, W / H, 1.0, 20.0)
    eye, s, u, f = RO.camera_basis(CAMS[cam], TARGET, UP)
    d = RO.pixel_rays(W, H, 30.0, s, u, f)
    rng = np.random.default_rng(cam)
    for _ in range(200):
        py, px = rng.integers(0, H), rng.integers(0, W)
        depth = rng.uniform(1.0, 5.0)
        world = np.asarray(eye, float) + depth * d[py, px].astype(float)
        clip = P @ V @ np.append(world, 1.0)
        ndc = clip[:3] / clip[3]
        # pixel centre of the window transform (row 0 at the top)
        assert abs((ndc[0] + 1) / 2 * W - (px + 0.5)) < 1e-3
        assert abs((1 - ndc[1]) / 2 * H - (py + 0.5)) < 1e-3
        assert abs(-(V @ np.append(world, 1.0))[2] - depth) < 1e-4  # the ray parameter is camera-space depth


def test_camera_zero_looks_down_minus_y_with_world_minus_x_to_the_right():
    eye, s, u, f = RO.camera_basis(CAMS[0], TARGET, UP)
    assert np.allclose(s, [-1, 0, 0], atol=1e-6) and f[1] < 0 and f[2] < 0 and u[2] > 0


def _scene_kw(W=51, H=51, cams=1):
    return dict(half_extents=[(1.5, 1.5, 0.05), (0.1, 0.1, 0.025), (0.005, 0.005, 0.25)],
                rgb=[(0.3, 0.3, 0.0), (0.9, 0.2, 0.1), (0.2, 0.7, 0.1)], cameras=CAMS[:cams], target=TARGET, up=UP,
                near=1.0, far=20.0, fov_deg=30.0, width=W, height=H, light_dir=(0.4, 0.3, 0.866), ambient=0.6, diffuse=0.35,
                background=(1.0, 1.0, 1.0))


def test_pole_on_the_optical_axis_is_a_vertical_green_bar():
    obs = np.zeros((1, 2, 7), np.float32)
    obs[0, 0] = [0, 0, 0.075, 0, 0, 0, 1]   # cart on the ground
    obs[0, 1] = [0, 0, 0.35, 0, 0, 0, 1]    # pole standing on it, centred on the target's axis
    img = RO.render_env(obs, **_scene_kw()).astype(np.float32)[:, :, :, 0, 0]
    H, W = img.shape[:2]
    green = (img[:, :, 1] > img[:, :, 0] + 0.1) & (img[:, :, 1] > img[:, :, 2] + 0.1)
    red = (img[:, :, 0] > img[:, :, 1] + 0.2) & (img[:, :, 0] > img[:, :, 2] + 0.2)
    cols = np.where(green.any(axis=0))[0]
    assert len(cols) >= 1 and abs(cols.mean() - W // 2) <= 0.5   # a bar one or two pixels wide in the middle column
    assert cols.max() - cols.min() <= 1
    rows = np.where(green.any(axis=1))[0]
    assert rows.max() - rows.min() >= H // 3                      # 0.5 m tall at ~1 m: a good part of a 30 degree view
    assert red.any() and np.where(red.any(axis=1))[0].min() > rows.min()  # the cart lies below the pole's top
    # the u8 grid: every value is k / 255 rounded to float16
    k = np.round(img * 255)
    assert np.array_equal((k / 255).astype(np.float16), RO.render_env(obs, **_scene_kw())[:, :, :, 0, 0])


def test_far_and_near_planes_clip():
    obs = np.zeros((1, 2, 7), np.float32)
    obs[0, :, 6] = 1
    obs[0, :, 2] = 100.0  # both boxes out of sight
    kw = _scene_kw()
    kw.update(near=0.01, far=0.3, clip=True)  # the ground lies beyond the far plane along every ray
    img = RO.render_env(obs, **kw)
    assert np.all(img == np.float16(1.0))
    kw.update(near=1.0, far=20.0, clip=False)
    img = RO.render_env(obs, **kw).astype(np.float32)
    assert (img[-1, :, 2, 0, 0] < 0.01).all()  # bottom row: ground colour (no blue), shaded
