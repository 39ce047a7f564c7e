"""DeviceReplayMemory — the reference's `ReplayMemory` (replay_memory.py:11-165) resident in HBM,
fed by the batched env step through the C ABI of include/bcp_replay.h.

Same constructor arguments and method names as the reference class (`add_episode`, `size`,
`batch`, `random_indexes`, `current_stats`, `reset_from_event_log`); new are `begin` /
`observe`, which do for a batch of envs what the agents' loops do for one
(ddpg_cartpole.py:308-326): cache every step's (action, reward, state_2) and add the episode
when it is done.  Ring, state pool and free-slot list are bit-identical to the reference's
after the same episodes in the same order.  There is no CPU path: the library must load and a
CUDA device must be present.
"""
from __future__ import annotations

import collections
import ctypes
from typing import Optional

import numpy as np
import torch

from . import _lib

Batch = collections.namedtuple("Batch", "state_1 action reward terminal_mask state_2")  # replay_memory.py:9


class _Batch(Batch):
    """Batch (unpacks like the reference's 5-tuple) that also remembers the row indexes it used."""
    idxs = None


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


class DeviceReplayMemory:
    def __init__(self, buffer_size, state_shape, action_dim, load_factor=1.5, num_envs: int = 0,
                 max_episode_len: int = 200, device=0, seed: int = 0):
        assert load_factor >= 1.5, "load_factor has to be at least 1.5"  # replay_memory.py:13
        if not torch.cuda.is_available():
            raise _lib.BcpError("DeviceReplayMemory needs a CUDA device: there is no CPU path")
        self._lib = _lib.load()
        self.device = torch.device("cuda", int(device) if not isinstance(device, torch.device) else device.index or 0)
        self.buffer_size = int(buffer_size)
        self.state_shape = tuple(int(s) for s in state_shape)
        self.action_dim = int(action_dim)
        self.state_buffer_size = int(buffer_size * load_factor)  # :31
        self.num_envs = int(num_envs)
        self.max_episode_len = int(max_episode_len)
        self._words = int(np.prod(self.state_shape))
        cfg = _lib.BcpReplayConfig(_lib.REPLAY_ABI_VERSION, self.device.index, self.buffer_size, self.state_buffer_size,
                                   self._words, self.action_dim, self.num_envs, self.max_episode_len)
        h = ctypes.c_void_p()
        _lib.check(self._lib.bcp_replay_create(ctypes.byref(cfg), ctypes.byref(h)), self._lib)
        self._h = h
        self._seed = int(seed)
        self._draws = 0
        self._nonempty = False
        self.stats = collections.Counter()  # '>batch' is counted here; the rest lives on the device

    def close(self):
        if getattr(self, "_h", None):
            self._lib.bcp_replay_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _u8(self, x):
        if type(x) is torch.Tensor and x.is_cuda and x.is_contiguous():
            if x.dtype == torch.uint8:
                return x
            if x.dtype == torch.bool:
                return x.view(torch.uint8)  # 0/1 bytes: a view, no kernel
        return torch.as_tensor(x, device=self.device).to(torch.uint8).contiguous()

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _f32(self, x, shape):
        if type(x) is torch.Tensor and x.dtype == torch.float32 and x.is_cuda and x.is_contiguous():
            t = x  # the env's own output buffers: nothing to convert
        else:
            t = torch.as_tensor(x, dtype=torch.float32, device=self.device).contiguous()
        if t.numel() != int(np.prod(shape)):
            raise ValueError(f"expected {shape}, got {tuple(t.shape)}")
        return t

    # ---- batched feeding ----
    def begin(self, obs, mask=None):
        """After env.reset(): obs[e] is the initial state of a new episode of env e (every env, or where mask)."""
        o = self._f32(obs, (self.num_envs,) + self.state_shape)
        m = None if mask is None else self._u8(mask)
        _lib.check(self._lib.bcp_replay_begin(self._h, _p(o), _p(m), self._stream()), self._lib)

    def observe(self, actions, obs, reward, done, reason=None):
        """After env.step(actions) -> (obs, reward, done, info): stage one transition per open episode
        and add every episode that finished (env-id order).  `reason` = info['done_reason'] of a
        pipelined auto-reset env (codes 4 / 5 are not transitions)."""
        a = actions if type(actions) is torch.Tensor and actions.is_cuda else torch.as_tensor(actions, device=self.device)
        is_int = not a.dtype.is_floating_point
        want = torch.int32 if is_int else torch.float32
        if a.dtype != want or not a.is_contiguous():
            a = a.to(want).contiguous()
        if a.numel() != self.num_envs * (1 if is_int else self.action_dim):
            raise ValueError(f"actions: got {tuple(a.shape)} for {self.num_envs} envs, action_dim {self.action_dim}")
        o = self._f32(obs, (self.num_envs,) + self.state_shape)
        r = self._f32(reward, (self.num_envs,))
        d = self._u8(done)
        rs = None if reason is None else self._u8(reason)
        _lib.check(self._lib.bcp_replay_observe(self._h, _p(a), 1 if is_int else 0, _p(o), _p(r), _p(d), _p(rs), self._stream()),
                   self._lib)

    # ---- the reference's API ----
    def add_episode(self, initial_state, action_reward_state_sequence):  # replay_memory.py:64-72
        assert len(action_reward_state_sequence) > 0
        acts = np.stack([np.asarray(a, dtype=np.float32).reshape(self.action_dim) for a, _, _ in action_reward_state_sequence])
        rews = np.asarray([r for _, r, _ in action_reward_state_sequence], dtype=np.float32)
        st2 = np.stack([np.asarray(s, dtype=np.float32).reshape(self._words) for _, _, s in action_reward_state_sequence])
        self.add_episode_arrays(np.asarray(initial_state, dtype=np.float32), acts, rews, st2)

    def add_episode_arrays(self, initial_state, actions, rewards, states2):
        """add_episode with the sequence as arrays (numpy or torch, host or device)."""
        n = int(len(rewards))
        i0 = self._f32(initial_state, self.state_shape)
        a = self._f32(actions, (n, self.action_dim))
        r = self._f32(rewards, (n,))
        s2 = self._f32(states2, (n,) + self.state_shape)
        _lib.check(self._lib.bcp_replay_add_episode(self._h, _p(i0), _p(a), _p(r), _p(s2), n, self._stream()), self._lib)

    def info(self):
        out = np.zeros(_lib.REPLAY_INFO_WORDS, dtype=np.int64)
        _lib.check(self._lib.bcp_replay_info(self._h, out.ctypes.data_as(ctypes.c_void_p), self._stream()), self._lib)
        names = ("insert", "full", "size", "free_slots", ">add_episode", ">add", "cache_evicted_s2", "error",
                 "last_commit_episodes", "last_commit_transitions")
        d = dict(zip(names, (int(v) for v in out)))
        if d["error"] & _lib.REPLAY_ERR_UNDERFLOW:
            raise IndexError("pop from empty list")  # what the reference raises, replay_memory.py:66,105
        if d["error"]:
            raise _lib.BcpError(f"replay memory error bits {d['error']} (see include/bcp_replay.h)")
        return d

    @property
    def insert(self):
        return self.info()["insert"]

    @property
    def full(self):
        return bool(self.info()["full"])

    def size(self):  # :120-121
        return self.info()["size"]

    def random_indexes(self, n=1):  # :123-129, a counter-based draw on the device
        if self.size() == 0:
            return []
        return self.batch(n).idxs

    def batch(self, batch_size=None, idxs=None):  # :131-138
        """-> Batch of CUDA tensors (states float16 like the reference's pool); `idxs` as in the
        reference's own test (`rm.batch(idxs=[0,1,2])`, replay_memory_test.py:85) fixes the rows."""
        self.stats[">batch"] += 1
        it = None
        if idxs is not None:
            it = torch.as_tensor(idxs, dtype=torch.int64, device=self.device).contiguous()
            n = int(it.numel())
        else:
            n = int(batch_size or 1)
        if n > 0 and not self._nonempty:  # one sync until the memory has rows, none afterwards
            self._nonempty = self.size() > 0
            if not self._nonempty:
                n = 0  # :126-127: an empty memory yields empty arrays
        dev = self.device
        s1 = torch.empty((n,) + self.state_shape, dtype=torch.float16, device=dev)
        s2 = torch.empty((n,) + self.state_shape, dtype=torch.float16, device=dev)
        a = torch.empty((n, self.action_dim), dtype=torch.float32, device=dev)
        r = torch.empty((n, 1), dtype=torch.float32, device=dev)
        m = torch.empty((n, 1), dtype=torch.float32, device=dev)
        used = torch.empty((n,), dtype=torch.int64, device=dev)
        if n:
            self._draws += 1
            _lib.check(self._lib.bcp_replay_batch(self._h, _p(it), n, self._seed, self._draws, _p(s1), _p(a), _p(r), _p(m),
                                                  _p(s2), _p(used), self._stream()), self._lib)
        b = _Batch(s1, a, r, m, s2)
        b.idxs = used
        return b

    def current_stats(self):  # :162-165
        d = self.info()
        out = {k: d[k] for k in (">add_episode", ">add", "cache_evicted_s2") if d[k]}
        out.update(self.stats)
        out["free_slots"] = d["free_slots"]
        return out

    def export(self):
        """The internals as numpy arrays (parity tests): what the reference holds in its attributes."""
        d = self.info()
        N, M, A = self.buffer_size, self.state_buffer_size, self.action_dim
        out = {"insert": d["insert"], "full": d["full"], "state_1_idx": np.zeros(N, np.int32), "action": np.zeros((N, A), np.float32),
               "reward": np.zeros((N, 1), np.float32), "terminal_mask": np.zeros((N, 1), np.float32),
               "state_2_idx": np.zeros(N, np.int32), "state": np.zeros((M,) + self.state_shape, np.float16),
               "free_slots": np.zeros(M, np.int32)}
        vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)  # noqa: E731
        _lib.check(self._lib.bcp_replay_export(self._h, vp(out["state_1_idx"]), vp(out["action"]), vp(out["reward"]),
                                               vp(out["terminal_mask"]), vp(out["state_2_idx"]), vp(out["state"]),
                                               vp(out["free_slots"]), self._stream()), self._lib)
        out["free_slots"] = out["free_slots"][:d["free_slots"]]
        return out

    def reset_from_event_log(self, log_file):  # :43-62
        from . import event_log
        num_episodes = num_events = 0
        for episode in event_log.EventLogReader(log_file).entries():
            initial_state, seq = None, []
            for event_id, event in enumerate(episode.event):
                if event_id == 0:
                    assert len(event.action) == 0 and event.reward is None
                    initial_state = event_log.read_state_from_event(event)
                else:
                    seq.append((np.asarray(event.action, dtype=np.float32), event.reward, event_log.read_state_from_event(event)))
                num_events += 1
            num_episodes += 1
            self.add_episode(initial_state, seq)
            if self.full:
                break
        return num_episodes, num_events

    @property
    def launch_count(self) -> int:
        return int(self._lib.bcp_replay_launch_count(self._h))
