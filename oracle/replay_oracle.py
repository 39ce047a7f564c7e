"""CPU restatement of the reference's replay memory — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this
module; nothing under cartpoleplusplus_b200/ does (the product path is the CUDA ring of
csrc/bcp_replay.cu and fails loudly without it).

Follows /root/reference/replay_memory.py (cited per function).  Pin status: PINNED —
 * the known answers of the reference's own replay_memory_test.py (test_empty_memory :19-31,
   test_adds_to_full :33-57, test_adds_over_full :59-87) hold for this restatement
   (tests/test_replay_oracle.py);
 * tests/golden/replay_golden.npz holds the arrays the reference class itself produced here
   (tests/golden/gen_replay_golden.py executes the reference's own method sources), and this
   restatement reproduces them bit for bit.

Plain numpy / Python loops: sized for the small cases of the parity tests.
"""
from __future__ import annotations

import collections

import numpy as np

Batch = collections.namedtuple("Batch", "state_1 action reward terminal_mask state_2")  # replay_memory.py:9


class ReplayMemoryOracle:
    """replay_memory.py:11-160: ring of (s1_idx, action, reward, terminal_mask, s2_idx) rows over a
    shared float16 state pool with a FIFO free-slot list."""

    def __init__(self, buffer_size, state_shape, action_dim, load_factor=1.5):  # :12-41
        assert load_factor >= 1.5, "load_factor has to be at least 1.5"  # :13
        self.buffer_size = int(buffer_size)
        self.state_shape = tuple(state_shape)
        self.insert = 0
        self.full = False
        self.state_1_idx = np.zeros(buffer_size, dtype=np.int32)  # np.empty in the reference; 0 keeps exports comparable
        self.action = np.zeros((buffer_size, action_dim), dtype=np.float32)
        self.reward = np.zeros((buffer_size, 1), dtype=np.float32)
        self.terminal_mask = np.zeros((buffer_size, 1), dtype=np.float32)
        self.state_2_idx = np.zeros(buffer_size, dtype=np.int32)
        self.state_buffer_size = int(buffer_size * load_factor)  # :31
        self.state = np.zeros([self.state_buffer_size] + list(self.state_shape), dtype=np.float16)  # :32-33
        self.state_free_slots = collections.deque(range(self.state_buffer_size))  # :36 (list.pop(0) == popleft)
        self.stats = collections.Counter()
        self.underflow = False  # the reference raises IndexError (pop from empty list) here

    def add_episode(self, initial_state, action_reward_state_sequence):  # :64-72
        self.stats[">add_episode"] += 1
        assert len(action_reward_state_sequence) > 0
        state_1_idx = self._pop()
        self.state[state_1_idx] = initial_state
        for n, (action, reward, state_2) in enumerate(action_reward_state_sequence):
            terminal = n == len(action_reward_state_sequence) - 1
            state_1_idx = self._add(state_1_idx, action, reward, terminal, state_2)

    def _pop(self):
        if not self.state_free_slots:
            self.underflow = True
            raise IndexError("pop from empty list")
        return self.state_free_slots.popleft()

    def _add(self, s1_idx, a, r, t, s2):  # :74-118
        self.stats[">add"] += 1
        if self.full:  # :83-93 the row about to be clobbered frees its state_1 slot, and its
            self.state_free_slots.append(int(self.state_1_idx[self.insert]))  # state_2 slot if terminal
            if self.terminal_mask[self.insert] == 0:
                self.state_free_slots.append(int(self.state_2_idx[self.insert]))
                self.stats["cache_evicted_s2"] += 1
        self.state_1_idx[self.insert] = s1_idx  # :96-98
        self.action[self.insert] = a
        self.reward[self.insert] = r
        self.terminal_mask[self.insert] = 0.0 if t else 1.0  # :102
        s2_idx = self._pop()  # :105-107
        self.state_2_idx[self.insert] = s2_idx
        self.state[s2_idx] = s2
        self.insert += 1  # :110-113
        if self.insert >= self.buffer_size:
            self.insert = 0
            self.full = True
        return s2_idx

    def size(self):  # :120-121
        return self.buffer_size if self.full else self.insert

    def batch(self, idxs):  # :131-138 with the indexes given (random_indexes :123-129 draws them)
        idxs = np.asarray(idxs, dtype=np.int64)
        return Batch(np.copy(self.state[self.state_1_idx[idxs]]), np.copy(self.action[idxs]), np.copy(self.reward[idxs]),
                     np.copy(self.terminal_mask[idxs]), np.copy(self.state[self.state_2_idx[idxs]]))

    def current_stats(self):  # :162-165
        out = dict(self.stats)
        out["free_slots"] = len(self.state_free_slots)
        return out

    def export(self):
        return {"insert": self.insert, "full": int(self.full), "state_1_idx": self.state_1_idx.copy(),
                "action": self.action.copy(), "reward": self.reward.copy(), "terminal_mask": self.terminal_mask.copy(),
                "state_2_idx": self.state_2_idx.copy(), "state": self.state.copy(),
                "free_slots": np.asarray(self.state_free_slots, dtype=np.int32)}


class BatchedFeeder:
    """How a batch of envs feeds the reference's per-episode API (ddpg_cartpole.py:308-326,
    naf_cartpole.py:340-361: the agent caches (action, reward, copy(state_2)) per step and calls
    add_episode once the episode is done).  With B envs stepping together, episodes that finish
    in the same env.step call are added in env-id order."""

    def __init__(self, memory, num_envs):
        self.rm = memory
        self.initial = [None] * num_envs
        self.seq = [None] * num_envs

    def begin(self, obs, mask=None):  # env.reset(): obs[e] is the initial state of a new episode
        for e in range(len(self.seq)):
            if mask is None or mask[e]:
                self.initial[e] = np.array(obs[e], dtype=np.float32)
                self.seq[e] = []

    def observe(self, actions, obs, reward, done, reason=None):
        for e in range(len(self.seq)):
            code = 0 if reason is None else int(reason[e])
            if code == 4:  # BCP_REASON_RESETTING: not a transition
                continue
            if code == 5:  # BCP_REASON_FRESH: obs[e] is the reset observation
                self.initial[e] = np.array(obs[e], dtype=np.float32)
                self.seq[e] = []
                continue
            if self.seq[e] is None:  # step after done (:179-181): nothing to record
                continue
            self.seq[e].append((np.array(actions[e], dtype=np.float32), float(reward[e]), np.array(obs[e], dtype=np.float32)))
            if done[e]:
                self.rm.add_episode(self.initial[e], self.seq[e])
                self.seq[e] = None
