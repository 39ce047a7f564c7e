"""CPU tier: the replay-memory oracle (oracle/replay_oracle.py) against the reference's own
known answers (replay_memory_test.py) and against arrays the reference class produced
(tests/golden/replay_golden.npz, made by tests/golden/gen_replay_golden.py)."""
import os

import numpy as np
import pytest

from oracle.replay_oracle import BatchedFeeder, ReplayMemoryOracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "replay_golden.npz")
CASES = ("wrap", "tiny", "notfull")


def golden_case(name):
    g = np.load(GOLD)
    bs, ad, sbs, seed, n_eps, max_len = (int(v) for v in g[name + "/cfg"])
    shape = tuple(int(v) for v in g[name + "/state_shape"])
    lens = g[name + "/ep_len"]
    offs = np.concatenate([[0], np.cumsum(lens)])
    eps = [(g[name + "/init"][k], g[name + "/actions"][offs[k]:offs[k + 1]], g[name + "/rewards"][offs[k]:offs[k + 1]],
            g[name + "/states2"][offs[k]:offs[k + 1]]) for k in range(len(lens))]
    return g, dict(buffer_size=bs, state_shape=shape, action_dim=ad, load_factor=float(g[name + "/load_factor"][0])), eps


def assert_matches_golden(g, name, ex, stats):
    assert [ex["insert"], ex["full"]] == list(g[name + "/insert"][:2])
    for k in ("state_1_idx", "action", "reward", "terminal_mask", "state_2_idx", "free_slots"):
        np.testing.assert_array_equal(ex[k].reshape(g[name + "/" + k].shape), g[name + "/" + k], err_msg=k)
    assert ex["state"].dtype == np.float16
    np.testing.assert_array_equal(ex["state"].view(np.uint16).reshape(-1), g[name + "/state"].view(np.uint16).reshape(-1))
    assert [stats.get(">add_episode", 0), stats.get(">add", 0), stats.get("cache_evicted_s2", 0), stats["free_slots"]] == \
        list(g[name + "/stats"])


def s_for(i):  # replay_memory_test.py:60-61
    return (np.array(range(1, 7)) + (10 * i)).reshape(2, 3)


def test_empty_memory():  # replay_memory_test.py:19-31
    rm = ReplayMemoryOracle(buffer_size=3, state_shape=(2, 3), action_dim=2, load_factor=2)
    assert rm.size() == 0 and rm.insert == 0 and rm.full is False
    b = rm.batch([])
    assert len(b) == 5 and all(len(x) == 0 for x in b)


def test_adds_to_full():  # replay_memory_test.py:33-57
    rm = ReplayMemoryOracle(buffer_size=3, state_shape=(2, 3), action_dim=2, load_factor=2)
    rm.add_episode([[11, 12, 13], [14, 15, 16]], [(17, 18, [[21, 22, 23], [24, 25, 26]]), (27, 28, [[31, 32, 33], [34, 35, 36]]),
                                                  (37, 38, [[41, 42, 43], [44, 45, 46]])])
    assert rm.size() == 3 and rm.insert == 0 and rm.full is True
    assert [rm.state[i][0][0] for i in range(4)] == [11, 21, 31, 41]


def test_adds_over_full():  # replay_memory_test.py:59-87
    rm = ReplayMemoryOracle(buffer_size=3, state_shape=(2, 3), action_dim=2, load_factor=2)
    rm.add_episode(s_for(0), [((i * 10) + 7, (i * 10) + 8, s_for(i)) for i in range(1, 5)])
    rm.add_episode(s_for(5), [((i * 10) + 7, (i * 10) + 8, s_for(i)) for i in range(6, 9)])
    assert rm.size() == 3
    b = rm.batch([0, 1, 2])
    np.testing.assert_array_equal(b.reward, [[88], [68], [78]])
    np.testing.assert_array_equal(b.terminal_mask, [[0], [1], [1]])


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_the_reference_class(name):
    g, kw, eps = golden_case(name)
    rm = ReplayMemoryOracle(**kw)
    for init, acts, rews, st2 in eps:
        rm.add_episode(init, [(acts[i], rews[i], st2[i]) for i in range(len(rews))])
    assert_matches_golden(g, name, rm.export(), rm.current_stats())


def test_short_episodes_exhaust_the_state_pool():
    # 1-step episodes hold two states per row: 2N > int(1.5 N) slots => the reference pops an empty list
    rm = ReplayMemoryOracle(buffer_size=8, state_shape=(2,), action_dim=1)
    with pytest.raises(IndexError):
        for i in range(8):
            rm.add_episode([i, i], [((0,), 1.0, [i, -i])])
    assert rm.underflow


def test_batched_feeder_orders_same_step_finishes_by_env_id():
    rm = ReplayMemoryOracle(buffer_size=16, state_shape=(1,), action_dim=1)
    fd = BatchedFeeder(rm, 3)
    fd.begin(np.array([[10.], [20.], [30.]]))
    done = np.array([0, 0, 0])
    fd.observe(np.zeros((3, 1)), np.array([[11.], [21.], [31.]]), np.ones(3), done)
    fd.observe(np.zeros((3, 1)), np.array([[12.], [22.], [32.]]), np.ones(3), np.array([0, 1, 1]))
    # env 1's episode (2 rows) lands before env 2's
    assert rm.insert == 4
    assert [float(rm.state[i][0]) for i in rm.state_1_idx[:4]] == [20., 21., 30., 31.]
    np.testing.assert_array_equal(rm.terminal_mask[:4, 0], [1, 0, 1, 0])
    fd.observe(np.zeros((3, 1)), np.array([[13.], [0.], [0.]]), np.ones(3), np.array([1, 1, 1]))  # envs 1, 2: step after done
    assert rm.insert == 7 and rm.stats[">add_episode"] == 3
