"""GPU tier: bcp_render_pixels (the raw-pixel observation, include/bcp_render.h) against oracle/render_oracle.py,
and the env's --use-raw-pixels mode (bullet_cartpole.py:96-101,277-306)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _oracle_kw(r):
    a = r.args
    return dict(half_extents=[[a.half_extents[b][i] for i in range(3)] for b in range(3)],
                rgb=[[a.rgb[b][i] for i in range(3)] for b in range(3)],
                cameras=[[a.camera_pos[c][i] for i in range(3)] for c in range(a.num_cameras)],
                target=list(a.target), up=list(a.up), near=a.near_val, far=a.far_val, fov_deg=a.fov_deg, width=a.width,
                height=a.height, light_dir=list(a.light_dir), ambient=a.ambient, diffuse=a.diffuse, background=list(a.background),
                clip=bool(a.clip))


def _poses(rng, n, R):
    obs = np.zeros((n, R, 2, 7), np.float32)
    for e in range(n):
        for r in range(R):
            cart = rng.uniform(-0.25, 0.25, size=2)
            obs[e, r, 0, :3] = [cart[0], cart[1], 0.075]
            obs[e, r, 0, 3:] = [0, 0, 0, 1]
            q = np.array([rng.normal(scale=0.25), rng.normal(scale=0.25), rng.normal(scale=0.05), 1.0])
            obs[e, r, 1, :3] = [cart[0] + rng.normal(scale=0.02), cart[1] + rng.normal(scale=0.02), 0.35]
            obs[e, r, 1, 3:] = q / np.linalg.norm(q)
    return obs


@pytest.mark.parametrize("R,C,W,H,clip", [(2, 1, 50, 50, False), (2, 2, 50, 50, False), (3, 2, 64, 40, False), (1, 1, 33, 47, True)])
def test_kernel_matches_the_oracle(built, R, C, W, H, clip):
    from cartpoleplusplus_b200.render import PixelRenderer
    from oracle import render_oracle as RO
    n = 6
    r = PixelRenderer(n, repeats=R, num_cameras=C, width=W, height=H, clip=clip)
    obs = _poses(np.random.default_rng(R * 10 + C), n, R)
    got = r.render(torch.from_numpy(obs).cuda()).cpu().numpy()
    assert got.shape == (n, H, W, 3, C, R) and got.dtype == np.float16
    kw = _oracle_kw(r)
    seen_pole = 0
    for e in range(n):
        want = RO.render_env(obs[e], **kw)
        diff = np.abs(got[e].astype(np.float32) - want.astype(np.float32))
        # same arithmetic up to FMA contraction: a silhouette pixel or a uint8 truncation may fall the other way
        assert (diff > 0).mean() < 0.004, (e, (diff > 0).mean())
        assert (diff > 1.5 / 255).mean() < 0.002, (e, (diff > 1.5 / 255).mean())
        g = want.astype(np.float32)
        seen_pole += int(((g[:, :, 1] > g[:, :, 0] + 0.1) & (g[:, :, 1] > g[:, :, 2] + 0.1)).any())
    assert seen_pole == (0 if clip else n)  # the reference's near plane would clip its own cart and pole away


def test_argument_checks(built):
    import ctypes
    from cartpoleplusplus_b200 import _lib
    from cartpoleplusplus_b200.render import PixelRenderer, render_args
    with pytest.raises(ValueError, match="--num-cameras must be 1 or 2"):
        PixelRenderer(4, num_cameras=3)
    lib = _lib.load()
    a = render_args(num_envs=4)
    assert lib.bcp_render_pixels(ctypes.byref(a), 0, None) != 0 and b"null" in lib.bcp_last_error()
    r = PixelRenderer(4)
    with pytest.raises(ValueError):
        r.render(torch.zeros((3, 2, 2, 7), device="cuda"))


def test_env_raw_pixel_mode(built):
    """--use-raw-pixels: reset() / step() return the (H, W, 3, C, R) block per env, rendered from the same call's poses."""
    from cartpoleplusplus_b200 import default_opts
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    from cartpoleplusplus_b200.render import PixelRenderer
    n = 64
    opts = default_opts(use_raw_pixels=True, num_cameras=2, render_width=40, render_height=30, action_force=100.0)
    env = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=n, device=0, seed=1)
    assert env.observation_space.shape == (30, 40, 3, 2, 2)
    s0 = env.reset()
    assert s0.shape == (n, 30, 40, 3, 2, 2) and s0.dtype == torch.float16
    ref = PixelRenderer(n, 2, 2, 40, 30)
    assert torch.equal(s0, ref.render(env.obs))
    g = torch.Generator(device="cuda").manual_seed(0)
    moved = False
    for _ in range(25):
        s, rew, done, info = env.step(torch.rand((n, 2), device="cuda", generator=g) * 2 - 1)
        assert torch.equal(s, ref.render(env.obs))
        moved |= not torch.equal(s, s0)
    assert moved and float(s.min()) >= 0.0 and float(s.max()) <= 1.0
    with pytest.raises(Exception):
        env.step_host(torch.zeros((n, 2)))
    env.close()


def test_full_batch_throughput_shape(built):
    """65 536 envs (BASELINE configs[2]) at the reference's 50 x 50: the grid spans more than 65 535 blocks in x."""
    from cartpoleplusplus_b200.render import PixelRenderer
    n = 65536
    r = PixelRenderer(n)
    obs = torch.zeros((n, 2, 2, 7), device="cuda")
    obs[:, :, :, 6] = 1
    obs[:, :, 0, 2] = 0.075
    obs[:, :, 1, 2] = 0.35
    obs[:, :, :, 0] = torch.linspace(-0.2, 0.2, n, device="cuda")[:, None, None]
    px = r.render(obs)
    torch.cuda.synchronize()
    assert torch.equal(px[0], px[0]) and not torch.equal(px[0], px[-1])
    first = PixelRenderer(1).render(obs[:1])
    last = PixelRenderer(1).render(obs[-1:])
    assert torch.equal(px[:1], first) and torch.equal(px[-1:], last)
