"""-m gpu: the README's random-agent protocols on the CUDA path (README.md:76-93)."""
import numpy as np
import pytest

from cartpoleplusplus_b200 import default_opts, protocol
from cartpoleplusplus_b200 import config as C

pytestmark = pytest.mark.gpu


def _env(n, dtype, discrete=True, offset=0, seed=7, **ov):
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    return BatchedBulletCartpole(default_opts(**ov), discrete_actions=discrete, num_envs=n, device=0, dtype=dtype,
                                 seed=seed, env_id_offset=offset)


@pytest.mark.parametrize("dtype", ["fp64", "fp32"])
def test_stability_protocol_all_episodes_time_out(built, dtype):
    """BASELINE.json configs[1] / README.md:77-80: 4096 envs, --initial-force=0 --actions=0:
    every episode must last exactly 200 steps (the one exact known answer of the reference)."""
    env = _env(4096, dtype, initial_force=0.0)
    hist, header = protocol.run_random_agent(env, 4096, actions=(0,))
    assert header["episodes"] == 4096 and hist[200] == 4096 and hist.sum() == 4096
    assert np.array_equal(protocol.deciles_from_histogram(hist), [200] * 11)
    env.close()


def test_fp64_rollout_histogram_equals_oracle(built):
    """The in-kernel protocol (Philox actions, pipelined resets) against the oracle's rollout
    (synchronous resets): same episodes, bit-exact lengths."""
    from oracle import Oracle
    n = 192
    env = _env(n, "fp64", initial_force=55.0, max_episode_len=60)
    ora = Oracle(env.cfg, dtype="fp64")
    ora.state[:, C.S_DONE] = 1
    ora.rollout_random(400, action_mask=0x1F, max_episodes=3)
    hist, header = protocol.run_random_agent(env, 3 * n, actions=(0, 1, 2, 3, 4))
    assert np.array_equal(hist, ora.hist)
    assert header["episodes"] == 3 * n == int(ora.header[0]) and header["env_steps"] == int(ora.header[1])
    env.close()


def test_sharded_histograms_merge_to_the_unsharded_one(built):
    """RNG keyed by global env id: two half shards == one full batch, exactly (fp32 fast path)."""
    n = 2048
    full = _env(n, "fp32", initial_force=55.0)
    h_full, hd_full = protocol.run_random_agent(full, 2 * n)
    full.close()
    merged = np.zeros_like(h_full)
    for off in (0, n // 2):
        part = _env(n // 2, "fp32", offset=off, initial_force=55.0)
        h, _ = protocol.run_random_agent(part, n)
        merged += h
        part.close()
    assert np.array_equal(merged, h_full) and h_full.sum() == 2 * n


def test_fp32_vs_fp64_episode_lengths_ks(built):
    """north_star correctness part 3, with the fp64 parity mode standing in for pybullet
    (unavailable on this box): two-sample KS over ~10k episodes per mode, p > 0.05.  Different
    seeds per mode: the samples must be independent for the test to mean anything."""
    n = 2048
    hists = []
    # upstream-style switches: with the fork's frictionless cart.urdf the pushes never reach the
    # pole and every episode times out at 200 - a degenerate sample
    up = dict(cart_friction=1.0, steps_per_repeat=5, theta_per_reset=True, world_frame_force=True)
    for dtype, seed in (("fp64", 101), ("fp32", 202)):
        env = _env(n, dtype, seed=seed, initial_force=55.0, **up)
        h, hd = protocol.run_random_agent(env, 5 * n)
        assert hd["episodes"] == 5 * n
        hists.append(h)
        env.close()
    d, p = protocol.ks_two_sample(hists[0], hists[1])
    assert p > 0.05, (d, p, protocol.deciles_from_histogram(hists[0]), protocol.deciles_from_histogram(hists[1]))
    # README.md:90-92 for this protocol: median 9 (upstream's models and pybullet, n = 100); the
    # restated physics gives ~5 - same regime; a sanity window, not a pin
    med = protocol.deciles_from_histogram(hists[1])[5]
    assert 3 <= med <= 27, med
