"""CPU tier: the event-log writer/reader (cartpoleplusplus_b200/event_log.py) against the bytes the
reference's own EventLog wrote through the real protobuf serialiser
(tests/golden/eventlog_golden.bin, made by tests/golden/gen_eventlog_golden.py)."""
import os
import struct

import numpy as np

from cartpoleplusplus_b200 import event_log as el

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD_BIN = os.path.join(HERE, "golden", "eventlog_golden.bin")
GOLD = os.path.join(HERE, "golden", "eventlog_golden.npz")


def scenario():
    g = np.load(GOLD)
    eps = []
    for k in range(int(g["n_episodes"][0])):
        n = int(g[f"ep{k}/n"][0])
        seq = []
        for i in range(n):
            a = g[f"ep{k}/actions"][i]
            a = int(a[0]) if int(g[f"ep{k}/discrete"][0]) else a.reshape(1, -1)
            seq.append((a, float(g[f"ep{k}/rewards"][i]), g[f"ep{k}/states"][i]))
        eps.append((g[f"ep{k}/init"], seq))
    return g, eps


def test_writer_reproduces_the_reference_bytes(tmp_path):
    g, eps = scenario()
    path = str(tmp_path / "log.bin")
    log = el.EventLog(path, False)
    for init, seq in eps:  # bullet_cartpole.py:342-344, :272-273
        log.reset()
        log.add_just_state(init)
        for a, r, s in seq:
            log.add(s, a, r)
    log.reset()
    log.log_file.close()
    assert open(path, "rb").read() == open(GOLD_BIN, "rb").read()


def test_reader_on_the_reference_file():
    g, eps = scenario()
    got = list(el.EventLogReader(GOLD_BIN).entries())
    assert [len(e.event) for e in got] == list(g["events_per_episode"])
    np.testing.assert_array_equal(el.read_state_from_event(got[0].event[1]), g["first_step_state"])
    for ep, (init, seq) in zip(got, eps):
        assert len(ep.event[0].action) == 0 and not ep.event[0].HasField("reward")  # replay_memory.py:50-52
        np.testing.assert_array_equal(el.read_state_from_event(ep.event[0]).astype(np.float32), init)
        for ev, (a, r, s) in zip(ep.event[1:], seq):
            assert ev.HasField("reward") and ev.reward == r
            np.testing.assert_array_equal(np.float32(ev.action), np.float32([a] if isinstance(a, int) else a[0]))
            np.testing.assert_array_equal(el.read_state_from_event(ev).astype(np.float32), s)


def test_array_encoder_equals_event_encoder():
    rng = np.random.default_rng(3)
    for R, A, n in [(2, 2, 5), (1, 1, 1), (3, 2, 40), (2, 2, 0)]:
        init = rng.normal(size=(R, 2, 7)).astype(np.float32)
        acts = rng.normal(size=(n, A)).astype(np.float32)
        rews = rng.normal(size=n).astype(np.float32)
        sts = rng.normal(size=(n, R, 2, 7)).astype(np.float32)
        ref = el.frame_episode([el.encode_event(init)] + [el.encode_event(sts[i], acts[i], rews[i]) for i in range(n)])
        assert el.encode_episode_arrays(init, acts, rews, sts) == ref


def test_batched_log_orders_by_finish_then_env_id(tmp_path):
    path = str(tmp_path / "b.bin")
    log = el.BatchedEventLog(path, 3)
    z = lambda v: np.full((3, 2, 2, 7), v, dtype=np.float32)  # noqa: E731
    log.begin(z(0))
    log.observe(np.zeros((3, 2), np.float32), z(1), np.ones(3), np.array([0, 0, 0]))
    log.observe(np.ones((3, 2), np.float32), z(2), np.ones(3), np.array([0, 1, 1]), reason=np.array([0, 3, 3]))
    log.observe(np.ones((3, 2), np.float32), z(3), np.ones(3), np.array([1, 1, 0]), reason=np.array([3, 4, 5]))  # env 2 restarts
    log.observe(np.ones((3, 2), np.float32), z(4), np.ones(3), np.array([1, 1, 1]), reason=np.array([4, 4, 2]))
    log.close()
    eps = list(el.EventLogReader(path).entries())
    assert [len(e.event) for e in eps] == [3, 3, 4, 2]  # envs 1, 2 (2 steps), env 0 (3 steps), env 2's second episode
    assert eps[2].event[3].state[0].cart_pose[0] == 3.0 and eps[3].event[0].state[0].cart_pose[0] == 3.0
    assert struct.unpack("=l", open(path, "rb").read(4))[0] == 147 + 2 * 162


def test_empty_episode_is_not_written(tmp_path):  # event_log.py:51-53
    path = str(tmp_path / "e.bin")
    log = el.EventLog(path)
    log.reset()
    log.reset()
    log.close()
    assert os.path.getsize(path) == 0
