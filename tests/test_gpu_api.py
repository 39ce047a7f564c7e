"""-m gpu: the reference-facing surface end to end - the num_envs = 1 drop-in through a reference-style agent loop
(ddpg_cartpole.py:309-326), --event-log-out written from reset()/step() exactly where the reference does
(bullet_cartpole.py:97-102,272-273,342-344), the random-agent protocol called twice on one env."""
import numpy as np
import pytest

from cartpoleplusplus_b200 import default_opts, protocol
from cartpoleplusplus_b200 import config as C
from cartpoleplusplus_b200.event_log import EventLogReader, read_state_from_event

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def test_single_env_dropin_in_a_reference_style_agent_loop(built, tmp_path, capsys):
    from cartpoleplusplus_b200.env import BulletCartpole
    from oracle import Oracle
    path = str(tmp_path / "one.log")
    opts = default_opts(initial_force=55.0, action_force=100.0, max_episode_len=6, event_log_out=path)
    env = BulletCartpole(opts, discrete_actions=False, dtype="fp64", seed=9)
    ora = Oracle(env.cfg, dtype="fp64")
    assert env.observation_space.shape == (2, 2, 7) and env.action_space.shape == (1, 2)
    rng = np.random.default_rng(0)
    episodes = []
    for ep in range(3):
        state = env.reset()                                    # ddpg_cartpole.py:309
        assert isinstance(state, np.ndarray) and state.shape == (2, 2, 7) and state.dtype == np.float32
        assert np.array_equal(state, ora.reset()[0])
        done, trace = False, [(state, None, None)]
        while not done:                                        # :313-326
            action = rng.uniform(-1, 1, size=(1, 2)).astype(np.float32)
            state, reward, done, info = env.step(action)
            o = ora.step(action)
            assert isinstance(state, np.ndarray) and state.dtype == np.float32 and np.array_equal(state, o[0][0])
            assert reward == 1.0 and isinstance(reward, float) and isinstance(done, bool) and isinstance(info, dict)
            assert done == bool(o[2][0])
            if done:
                assert info["done_reason"] == C.REASON_NAMES[int(o[3][0])]
            trace.append((state, action, reward))
        episodes.append(trace)
        again = env.step(action)                               # bullet_cartpole.py:179-181
        assert np.array_equal(again[0], state) and again[1] == 0 and again[2] is True and again[3] == {}
        assert "calling step after done????" in capsys.readouterr().err
    env.close()                                                # flushes the open episode
    logged = list(EventLogReader(path).entries())
    assert len(logged) == 3
    for ep, trace in zip(logged, episodes):
        assert len(ep.event) == len(trace)
        for ev, (state, action, reward) in zip(ep.event, trace):
            assert np.array_equal(read_state_from_event(ev), state)
            if action is None:
                assert len(ev.action) == 0 and not ev.HasField("reward")
            else:
                assert np.array_equal(np.asarray(ev.action, dtype=np.float32), action[0]) and ev.reward == reward


def test_batched_event_log_out(built, tmp_path):
    """--event-log-out on the batched env: every finished episode is in the file - reset observation, then one event per
    live step with the action, the observation and the reward that step() returned."""
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    path = str(tmp_path / "batch.log")
    n = 5
    opts = default_opts(initial_force=55.0, max_episode_len=4, event_log_out=path)
    env = BatchedBulletCartpole(opts, discrete_actions=True, num_envs=n, device=0, dtype="fp32", seed=2)
    obs = env.reset()
    open_eps = {e: [(obs[e].cpu().numpy(), None)] for e in range(n)}
    finished = []
    rng = np.random.default_rng(1)
    for t in range(11):
        a = rng.integers(0, 5, size=n).astype(np.int32)
        obs, rew, done, info = env.step(torch.from_numpy(a).cuda())
        for e in range(n):
            if e in open_eps:
                open_eps[e].append((obs[e].cpu().numpy(), float(a[e])))
                if bool(done[e]):
                    finished.append(open_eps.pop(e))
        if t in (3, 8):   # masked reset of the finished envs
            m = done.to(torch.uint8)
            o2 = env.reset(m)
            for e in range(n):
                if bool(m[e]):
                    open_eps[e] = [(o2[e].cpu().numpy(), None)]
    env.close()
    logged = list(EventLogReader(path).entries())
    assert len(logged) == len(finished) == 2 * n
    for ep, want in zip(logged, finished):
        assert len(ep.event) == len(want) == 5
        for ev, (state, action) in zip(ep.event, want):
            assert np.array_equal(read_state_from_event(ev), state)
            assert (list(ev.action) == [] and not ev.HasField("reward")) if action is None else (list(ev.action) == [action] and ev.reward == 1.0)


def test_random_agent_protocol_twice_on_one_env(built):
    """the episode target is relative to where the env stands: a second run delivers its full share, with new episodes"""
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    n = 512
    env = BatchedBulletCartpole(default_opts(initial_force=55.0), discrete_actions=True, num_envs=n, device=0, dtype="fp32", seed=3)
    h1, hd1 = protocol.run_random_agent(env, 2 * n)
    h2, hd2 = protocol.run_random_agent(env, 3 * n)
    assert h1.sum() == hd1["episodes"] == 2 * n and h2.sum() == hd2["episodes"] == 3 * n
    assert not np.array_equal(h1 * 3, h2 * 2)
    # and after ordinary stepping (envs mid-episode, unequal reset counts)
    env.reset()
    for _ in range(5):
        env.step(torch.zeros(n, dtype=torch.int32, device="cuda"))
    h3, hd3 = protocol.run_random_agent(env, n)
    assert h3.sum() == hd3["episodes"] == n
    env.close()
