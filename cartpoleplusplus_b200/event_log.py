"""Event log of the env path: the reference's on-disk format (event.proto, event_log.py),
written and read without protoc / google.protobuf.

File = sequence of episodes, each framed by a native-endian int32 length (event_log.py:48-58,
105-111: `struct.pack('=l', len(buff))` + `Episode.SerializeToString()`).  Messages
(event.proto:1-35, proto2 — repeated floats are NOT packed, fields go out in field-number
order, which is what the protobuf serialiser the reference calls does):

    Episode { repeated Event event = 1; }                                     tag 0x0a
    Event   { repeated float action = 1; repeated State state = 2;            tags 0x0d 0x12
              optional float reward = 3; }                                    tag 0x1d
    State   { repeated float cart_pose = 1; repeated float pole_pose = 2;     tags 0x0d 0x15
              repeated Render render = 3; }                                   tag 0x1a (pixel mode only)

`EventLog` / `EventLogReader` / `read_state_from_event` mirror the reference classes
(event_log.py:22-111) for one env; `BatchedEventLog` does for a batch of envs what
bullet_cartpole.py:97-102,272-273,342-344 do for one (reset -> add_just_state, step -> add):
every env's episode is buffered and written when it ends, episodes that end in the same call in
env-id order.  Low-dim states only: the pixel mode needs pybullet's renderer (unavailable here).
Byte-for-byte parity with the reference writer is pinned by tests/golden/eventlog_golden.bin.
"""
from __future__ import annotations

import gzip
import struct
from typing import List, Optional

import numpy as np

_F32 = struct.Struct("<f")


def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _floats(tag: int, values) -> bytes:
    """repeated float, unpacked: tag + 4 little-endian bytes per element."""
    a = np.asarray(values, dtype="<f4").reshape(-1)
    rec = np.empty(a.size, dtype=np.dtype([("t", "u1"), ("v", "<f4")]))
    rec["t"] = tag
    rec["v"] = a
    return rec.tobytes()


def encode_state(cart_pose, pole_pose) -> bytes:
    return _floats(0x0D, cart_pose) + _floats(0x15, pole_pose)


def encode_event(state, action=None, reward=None) -> bytes:
    """state (R,2,7); action None | python int | sequence of floats; reward None | float."""
    out = bytearray()
    if action is not None:
        out += _floats(0x0D, [float(action)] if np.isscalar(action) else action)
    st = np.asarray(state, dtype=np.float32)
    for r in range(st.shape[0]):  # add_state_to_event, event_log.py:73-78
        body = encode_state(st[r][0], st[r][1])
        out += b"\x12" + _varint(len(body)) + body
    if reward is not None:
        out += b"\x1d" + _F32.pack(float(reward))
    return bytes(out)


def frame_episode(event_bodies: List[bytes]) -> bytes:
    """Episode bytes with the int32 length header; b'' for an episode without events (:51-53)."""
    buff = b"".join(b"\x0a" + _varint(len(e)) + e for e in event_bodies)
    if not buff:
        return b""
    return struct.pack("=l", len(buff)) + buff


def encode_episode_arrays(initial_state, actions, rewards, states) -> bytes:
    """A whole low-dim episode at once (numpy, no per-event Python): initial_state (R,2,7),
    actions (n,A), rewards (n,), states (n,R,2,7).  Same bytes as frame_episode of the events."""
    init = np.asarray(initial_state, dtype="<f4")
    R = init.shape[0]
    n = len(rewards)
    state_dt = np.dtype([("tag", "u1"), ("len", "u1"), ("f", [("t", "u1"), ("v", "<f4")], (14,))])  # 2 + 70 bytes

    def states_block(st):  # (m,R,2,7) -> (m,R) records
        st = np.asarray(st, dtype="<f4").reshape(-1, R, 14)
        rec = np.empty(st.shape[:2], dtype=state_dt)
        rec["tag"], rec["len"] = 0x12, 70
        rec["f"]["t"][..., :7] = 0x0D
        rec["f"]["t"][..., 7:] = 0x15
        rec["f"]["v"] = st
        return rec

    first = states_block(init[None])[0].tobytes()
    out = [b"\x0a" + _varint(len(first)) + first]
    if n:
        A = np.asarray(actions).reshape(n, -1).shape[1]
        body_len = 5 * A + 72 * R + 5
        head = b"\x0a" + _varint(body_len)
        ev_dt = np.dtype([("head", "u1", (len(head),)), ("a", [("t", "u1"), ("v", "<f4")], (A,)), ("s", state_dt, (R,)),
                          ("rt", "u1"), ("rv", "<f4")])
        ev = np.empty(n, dtype=ev_dt)
        ev["head"] = np.frombuffer(head, dtype="u1")
        ev["a"]["t"] = 0x0D
        ev["a"]["v"] = np.asarray(actions, dtype="<f4").reshape(n, A)
        ev["s"] = states_block(states)
        ev["rt"] = 0x1D
        ev["rv"] = np.asarray(rewards, dtype="<f4")
        out.append(ev.tobytes())
    buff = b"".join(out)
    return struct.pack("=l", len(buff)) + buff


# ---- reading ----
class State:
    def __init__(self):
        self.cart_pose: List[float] = []
        self.pole_pose: List[float] = []
        self.render: list = []


class Event:
    def __init__(self):
        self.action: List[float] = []
        self.state: List[State] = []
        self.reward: Optional[float] = None

    def HasField(self, name):
        assert name == "reward"
        return self.reward is not None


class Episode:
    def __init__(self):
        self.event: List[Event] = []


def _fields(buf: bytes):
    at, n = 0, len(buf)
    while at < n:
        key = shift = 0
        while True:
            b = buf[at]; at += 1
            key |= (b & 0x7F) << shift
            shift += 7
            if not b & 0x80:
                break
        field, wt = key >> 3, key & 7
        if wt == 5:
            yield field, _F32.unpack_from(buf, at)[0]
            at += 4
        elif wt == 2:
            ln = shift = 0
            while True:
                b = buf[at]; at += 1
                ln |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
            yield field, buf[at:at + ln]
            at += ln
        elif wt == 0:
            v = shift = 0
            while True:
                b = buf[at]; at += 1
                v |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
            yield field, v
        else:
            raise ValueError(f"unsupported wire type {wt}")


def parse_episode(buf: bytes) -> Episode:
    ep = Episode()
    for f, v in _fields(buf):
        if f != 1:
            continue
        ev = Event()
        for f2, v2 in _fields(v):
            if f2 == 1:
                ev.action.append(v2)
            elif f2 == 3:
                ev.reward = v2
            elif f2 == 2:
                st = State()
                for f3, v3 in _fields(v2):
                    if f3 == 1:
                        st.cart_pose.append(v3)
                    elif f3 == 2:
                        st.pole_pose.append(v3)
                    elif f3 == 3:
                        st.render.append(v3)
                ev.state.append(st)
        ep.event.append(ev)
    return ep


def read_state_from_event(event) -> np.ndarray:
    """event_log.py:22-41, low-dim branch: (len(event.state), 2, 7) float64."""
    if len(event.state[0].render) > 0:
        raise NotImplementedError("pixel states: pybullet's renderer is unavailable on this box")
    state = np.empty((len(event.state), 2, 7))
    for i, s in enumerate(event.state):
        state[i][0] = s.cart_pose
        state[i][1] = s.pole_pose
    return state


class EventLogReader:  # event_log.py:96-111
    def __init__(self, path):
        self.log_file = gzip.open(path, "rb") if path.endswith(".gz") else open(path, "rb")

    def entries(self):
        while True:
            buff_len_bytes = self.log_file.read(4)
            if len(buff_len_bytes) == 0:
                return
            buff_len = struct.unpack("=l", buff_len_bytes)[0]
            yield parse_episode(self.log_file.read(buff_len))


class EventLog:  # event_log.py:43-93, one env
    def __init__(self, path, use_raw_pixels=False):
        if use_raw_pixels:
            raise NotImplementedError("pixel states: pybullet's renderer is unavailable on this box")
        self.log_file = open(path, "ab")
        self.episode_entry: Optional[List[bytes]] = None
        self.use_raw_pixels = use_raw_pixels

    def reset(self):  # :48-58: the previous episode is written when the next one starts
        if self.episode_entry is not None:
            framed = frame_episode(self.episode_entry)
            if framed:
                self.log_file.write(framed)
                self.log_file.flush()
        self.episode_entry = []

    def add(self, state, action, reward):  # :80-88
        if isinstance(action, int):
            a = [float(action)]
        else:
            action = np.asarray(action)
            assert action.shape[0] == 1  # never log batch operations
            a = action[0]
        self.episode_entry.append(encode_event(state, a, reward))

    def add_just_state(self, state):  # :90-92
        self.episode_entry.append(encode_event(state))

    def close(self):
        """flush the open episode (the reference leaves it unwritten) and close the file"""
        self.reset()
        self.episode_entry = None
        self.log_file.close()


class BatchedEventLog:
    """One log file fed by a batch of envs.  `begin(obs[, mask])` after reset(), `observe(actions,
    obs, reward, done[, reason])` after step() — host arrays (numpy or CPU tensors, e.g. the pinned
    buffers of `BatchedBulletCartpole.step_host`).  An episode is written when it ends; episodes
    ending in the same call go out in env-id order."""

    def __init__(self, path, num_envs: int, max_episodes: Optional[int] = None):
        self.log_file = open(path, "ab")
        self.num_envs = int(num_envs)
        self.max_episodes = max_episodes
        self.episodes_written = 0
        self._init = [None] * self.num_envs
        self._acts = [None] * self.num_envs
        self._rews = [None] * self.num_envs
        self._sts = [None] * self.num_envs

    def _open(self, e, obs_e):
        self._init[e] = np.array(obs_e, dtype=np.float32)
        self._acts[e], self._rews[e], self._sts[e] = [], [], []

    def begin(self, obs, mask=None):
        obs = np.asarray(obs)
        for e in range(self.num_envs):
            if mask is None or mask[e]:
                self._open(e, obs[e])

    def observe(self, actions, obs, reward, done, reason=None):
        actions, obs, reward, done = np.asarray(actions), np.asarray(obs), np.asarray(reward), np.asarray(done)
        reason = None if reason is None else np.asarray(reason)
        chunks = []
        for e in range(self.num_envs):
            code = 0 if reason is None else int(reason[e])
            if code == 4:
                continue
            if code == 5:
                self._open(e, obs[e])
                continue
            if self._acts[e] is None:
                continue
            self._acts[e].append(np.array(actions[e], dtype=np.float32).reshape(-1))
            self._rews[e].append(float(reward[e]))
            self._sts[e].append(np.array(obs[e], dtype=np.float32))
            if done[e]:
                if self.max_episodes is None or self.episodes_written < self.max_episodes:
                    chunks.append(encode_episode_arrays(self._init[e], np.stack(self._acts[e]), self._rews[e], np.stack(self._sts[e])))
                    self.episodes_written += 1
                self._acts[e] = None
        if chunks:
            self.log_file.write(b"".join(chunks))
            self.log_file.flush()

    def close(self):
        self.log_file.close()
