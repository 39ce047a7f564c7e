"""GPU tier: the device replay memory (csrc/bcp_replay.cu through include/bcp_replay.h) against
the oracle restatement, the reference's own known answers (replay_memory_test.py) and the arrays
the reference class produced (tests/golden/replay_golden.npz).  Integer / index work and the
float16 pool are compared bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from test_replay_oracle import CASES, assert_matches_golden, golden_case, s_for  # noqa: E402


def _mem(**kw):
    from cartpoleplusplus_b200.replay import DeviceReplayMemory
    return DeviceReplayMemory(device=0, **kw)


def assert_same(dev, ora):
    a, b = dev.export(), ora.export()
    assert (a["insert"], a["full"]) == (b["insert"], b["full"])
    sz = dev.size()
    assert sz == ora.size()
    for k in ("state_1_idx", "state_2_idx", "action", "reward", "terminal_mask"):
        np.testing.assert_array_equal(a[k][:sz], b[k][:sz], err_msg=k)
    np.testing.assert_array_equal(a["free_slots"], b["free_slots"])
    used = np.unique(np.concatenate([a["state_1_idx"][:sz], a["state_2_idx"][:sz]]))
    np.testing.assert_array_equal(a["state"][used].view(np.uint16), b["state"][used].view(np.uint16))
    sa, sb = dev.current_stats(), ora.current_stats()
    for k in (">add_episode", ">add", "cache_evicted_s2", "free_slots"):
        assert sa.get(k, 0) == sb.get(k, 0), k


def test_reference_known_answers(built):  # replay_memory_test.py:19-87
    rm = _mem(buffer_size=3, state_shape=(2, 3), action_dim=2, load_factor=2)
    assert rm.size() == 0 and rm.insert == 0 and rm.full is False and rm.random_indexes() == []
    b = rm.batch(4)
    assert len(b) == 5 and all(len(x) == 0 for x in b)
    rm.add_episode([[11, 12, 13], [14, 15, 16]], [((17, 17), 18, [[21, 22, 23], [24, 25, 26]]), ((27, 27), 28, [[31, 32, 33], [34, 35, 36]]),
                                                  ((37, 37), 38, [[41, 42, 43], [44, 45, 46]])])
    assert rm.size() == 3 and rm.insert == 0 and rm.full is True
    idxs = rm.random_indexes(n=100).cpu().numpy()
    assert len(idxs) == 100 and sorted(set(idxs.tolist())) == [0, 1, 2]
    state = rm.export()["state"]
    assert [state[i][0][0] for i in range(4)] == [11, 21, 31, 41]
    rm.close()
    rm = _mem(buffer_size=3, state_shape=(2, 3), action_dim=2, load_factor=2)
    rm.add_episode(s_for(0), [(((i * 10) + 7,) * 2, (i * 10) + 8, s_for(i)) for i in range(1, 5)])
    rm.add_episode(s_for(5), [(((i * 10) + 7,) * 2, (i * 10) + 8, s_for(i)) for i in range(6, 9)])
    assert rm.size() == 3
    b = rm.batch(idxs=[0, 1, 2])
    np.testing.assert_array_equal(b.reward.cpu().numpy(), [[88], [68], [78]])
    np.testing.assert_array_equal(b.terminal_mask.cpu().numpy(), [[0], [1], [1]])
    np.testing.assert_array_equal(b.state_2.cpu().numpy()[:, 0, 0], [81, 61, 71])
    rm.close()


@pytest.mark.parametrize("name", CASES)
def test_reference_class_arrays(built, name):
    g, kw, eps = golden_case(name)
    rm = _mem(**kw)
    for init, acts, rews, st2 in eps:
        rm.add_episode_arrays(init, acts, rews, st2)
    assert_matches_golden(g, name, rm.export(), rm.current_stats())
    rm.close()


def test_pool_exhaustion_is_reported_like_the_reference(built):
    rm = _mem(buffer_size=8, state_shape=(2,), action_dim=1)
    for i in range(8):
        rm.add_episode([i, i], [((0,), 1.0, [i, -i])])
    with pytest.raises(IndexError):
        rm.size()
    rm.close()


@pytest.mark.parametrize("seed,buffer_size,load_factor", [(0, 50, 2.0), (1, 7, 3.0), (2, 4000, 1.5)])
def test_batched_feeding_matches_oracle(built, seed, buffer_size, load_factor):
    """random done / reason streams over 70 envs (a ragged last warp); every call compared."""
    from oracle.replay_oracle import BatchedFeeder, ReplayMemoryOracle
    rng = np.random.default_rng(seed)
    B, L, shape, A = 70, 12, (2, 2, 7), 2
    dev = _mem(buffer_size=buffer_size, state_shape=shape, action_dim=A, load_factor=load_factor, num_envs=B, max_episode_len=L)
    ora = ReplayMemoryOracle(buffer_size, shape, A, load_factor)
    fd = BatchedFeeder(ora, B)
    obs = rng.normal(size=(B,) + shape).astype(np.float32)
    dev.begin(torch.from_numpy(obs).cuda())
    fd.begin(obs)
    age = np.zeros(B, dtype=np.int64)
    phase = np.zeros(B, dtype=np.int64)  # 0 live, > 0: calls of RESETTING left, then FRESH
    for call in range(60):
        acts = rng.uniform(-1, 1, size=(B, A)).astype(np.float32)
        obs = (rng.normal(size=(B,) + shape) * 2).astype(np.float32)
        rew = rng.uniform(0, 1, size=B).astype(np.float32)
        done = np.zeros(B, dtype=np.uint8)
        reason = np.zeros(B, dtype=np.uint8)
        for e in range(B):
            if phase[e] > 1:
                phase[e] -= 1; reason[e] = 4; done[e] = 1
            elif phase[e] == 1:
                phase[e] = 0; reason[e] = 5; age[e] = 0
            else:
                age[e] += 1
                if age[e] >= L or rng.random() < 0.15:
                    done[e] = 1; reason[e] = 3; phase[e] = int(rng.integers(1, 4))
        args = (acts, obs, rew, done, reason)
        dev.observe(*(torch.from_numpy(x).cuda() for x in args))
        fd.observe(*args)
        if call % 7 == 0 or call == 59:
            assert_same(dev, ora)
    idxs = rng.integers(0, ora.size(), size=33)
    b, bo = dev.batch(idxs=idxs), ora.batch(idxs)
    for x, y in zip(b, bo):
        np.testing.assert_array_equal(x.cpu().numpy().reshape(y.shape), y)
    dev.close()


def test_discrete_actions_and_manual_reset(built):
    """int32 actions (action_dim 1), auto_reset off style: begin(mask) after reset(mask); step after done is skipped."""
    from oracle.replay_oracle import BatchedFeeder, ReplayMemoryOracle
    rng = np.random.default_rng(5)
    B, L, shape = 33, 6, (3,)
    dev = _mem(buffer_size=64, state_shape=shape, action_dim=1, num_envs=B, max_episode_len=L)
    ora = ReplayMemoryOracle(64, shape, 1)
    fd = BatchedFeeder(ora, B)
    obs = rng.normal(size=(B,) + shape).astype(np.float32)
    dev.begin(obs)
    fd.begin(obs)
    closed = np.zeros(B, dtype=bool)
    age = np.zeros(B, dtype=np.int64)
    for call in range(25):
        acts = rng.integers(0, 5, size=B).astype(np.int32)
        obs = rng.normal(size=(B,) + shape).astype(np.float32)
        rew = np.where(closed, 0, 1).astype(np.float32)
        age += 1
        done = (closed | (rng.random(B) < 0.2) | (age >= L)).astype(np.uint8)
        dev.observe(torch.from_numpy(acts).cuda(), obs, rew, done)
        fd.observe(acts.reshape(B, 1), obs, rew, done)
        closed |= done.astype(bool)
        if call % 5 == 4:  # reset the finished envs
            obs = rng.normal(size=(B,) + shape).astype(np.float32)
            dev.begin(obs, mask=closed.astype(np.uint8))
            fd.begin(obs, mask=closed)
            age[closed] = 0
            closed[:] = False
    assert_same(dev, ora)
    dev.close()


def test_staging_overflow_is_an_error(built):
    from cartpoleplusplus_b200._lib import BcpError
    dev = _mem(buffer_size=64, state_shape=(3,), action_dim=1, num_envs=4, max_episode_len=2)
    z = np.zeros((4, 3), dtype=np.float32)
    dev.begin(z)
    for _ in range(3):  # a third transition of an episode that may hold two
        dev.observe(np.zeros((4, 1), dtype=np.float32), z, np.ones(4, dtype=np.float32), np.zeros(4, dtype=np.uint8))
    with pytest.raises(BcpError):
        dev.size()
    dev.close()


def test_env_rollout_into_replay(built):
    """the real thing: BatchedBulletCartpole (pipelined auto-reset) feeding the ring; the same
    outputs fed to the oracle give the same memory."""
    from cartpoleplusplus_b200 import default_opts
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    from oracle.replay_oracle import BatchedFeeder, ReplayMemoryOracle
    B = 256
    opts = default_opts(action_force=100.0, max_episode_len=12, auto_reset="pipelined")
    env = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=B, device=0, dtype="fp32", seed=1)
    dev = _mem(buffer_size=3000, state_shape=(env.repeats, 2, 7), action_dim=2, num_envs=B, max_episode_len=12)
    ora = ReplayMemoryOracle(3000, (env.repeats, 2, 7), 2)
    fd = BatchedFeeder(ora, B)
    obs = env.reset()
    dev.begin(obs)
    fd.begin(obs.cpu().numpy())
    g = torch.Generator(device="cuda").manual_seed(0)
    for _ in range(100):
        a = torch.rand((B, 2), device="cuda", generator=g) * 2 - 1
        obs, rew, done, info = env.step(a)
        dev.observe(a, obs, rew, done, info["done_reason"])
        fd.observe(a.cpu().numpy(), obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy(), info["done_reason"].cpu().numpy())
    assert ora.stats[">add_episode"] > B and ora.full
    assert_same(dev, ora)
    b = dev.batch(128)
    bo = ora.batch(b.idxs.cpu().numpy())
    for x, y in zip(b, bo):
        np.testing.assert_array_equal(x.cpu().numpy().reshape(y.shape), y)
    env.close()
    dev.close()


def test_event_log_round_trip_into_replay(built, tmp_path):
    """env rollout -> BatchedEventLog file -> DeviceReplayMemory.reset_from_event_log (replay_memory.py:43-62)
    == the oracle fed the same episodes; the file also parses as the reference's format."""
    from cartpoleplusplus_b200 import default_opts
    from cartpoleplusplus_b200 import event_log as el
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    from oracle.replay_oracle import ReplayMemoryOracle
    B = 64
    opts = default_opts(action_force=100.0, max_episode_len=9, auto_reset="pipelined")
    env = BatchedBulletCartpole(opts, discrete_actions=False, num_envs=B, device=0, dtype="fp32", seed=2)
    path = str(tmp_path / "events.bin")
    log = el.BatchedEventLog(path, B)
    obs = env.reset()
    log.begin(obs.cpu().numpy())
    g = torch.Generator(device="cuda").manual_seed(1)
    for _ in range(90):
        a = torch.rand((B, 2), device="cuda", generator=g) * 2 - 1
        obs, rew, done, info = env.step(a)
        log.observe(a.cpu().numpy(), obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy(), info["done_reason"].cpu().numpy())
    log.close()
    env.close()
    assert log.episodes_written >= B
    dev = _mem(buffer_size=400, state_shape=(2, 2, 7), action_dim=2)
    n_eps, n_events = dev.reset_from_event_log(path)
    ora = ReplayMemoryOracle(400, (2, 2, 7), 2)
    k = 0
    for ep in el.EventLogReader(path).entries():
        seq = [(np.float32(ev.action), ev.reward, el.read_state_from_event(ev)) for ev in ep.event[1:]]
        ora.add_episode(el.read_state_from_event(ep.event[0]), seq)
        k += 1
        if ora.full:
            break
    assert k == n_eps and ora.full and dev.full
    assert_same(dev, ora)
    dev.close()
