"""BatchedBulletCartpole — the reference's gym surface over the CUDA C-ABI.

Mirrors `bullet_cartpole.BulletCartpole` (bullet_cartpole.py:48-359): same constructor
arguments `(opts, discrete_actions)`, same `reset()` / `step(action)` / `action_space` /
`observation_space` / no-op `configure`, `seed`, `render`; the batch dimension is new.
Tensors are PyTorch CUDA tensors; the extension only ever sees their raw pointers.

`BulletCartpole` below is the `num_envs == 1` drop-in returning NumPy values exactly like
the reference (obs `(R,2,7)` float32 copy, python float reward, bool done, info dict).
"""
from __future__ import annotations

import ctypes
import sys
from typing import Optional

import numpy as np
import torch

from . import _lib, spaces
from . import config as C


class BatchedBulletCartpole:
    def __init__(self, opts, discrete_actions: bool, num_envs: Optional[int] = None, device=None,
                 dtype: Optional[str] = None, seed: Optional[int] = None, env_id_offset: int = 0,
                 physics: Optional[C.PhysicsParams] = None, scene=None, init_offset=(0.0, 0.0)):
        if not torch.cuda.is_available():
            raise _lib.BcpError("BatchedBulletCartpole needs a CUDA device: the env step has no CPU path")
        self._lib = _lib.load()
        if device is None:
            device = getattr(opts, "device", 0)
        self.device = torch.device("cuda", int(device) if not isinstance(device, torch.device) else device.index or 0)
        self.cfg = C.make_config(opts, discrete_actions, num_envs=num_envs, dtype=dtype, seed=seed,
                                 device=self.device.index, env_id_offset=env_id_offset, physics=physics, scene=scene,
                                 init_offset=init_offset)
        self.num_envs = int(self.cfg.num_envs)
        self.discrete_actions = bool(discrete_actions)
        self.repeats = int(self.cfg.action_repeats)
        self.steps_per_repeat = int(self.cfg.steps_per_repeat)
        self.max_episode_len = int(self.cfg.max_episode_len)
        # spaces, bullet_cartpole.py:91-94,138-141
        self.action_space = spaces.Discrete(5) if discrete_actions else spaces.Box(-1.0, 1.0, shape=(1, 2))
        float_max = np.finfo(np.float32).max
        self.observation_space = spaces.Box(-float_max, float_max, (self.repeats, 2, 7))
        # --use-raw-pixels: the state is the (H, W, 3, C, R) image block (bullet_cartpole.py:96-101,299-306),
        # ray-cast on the device from the poses of the call (render.py); env.obs keeps the poses
        self.use_raw_pixels = bool(getattr(opts, "use_raw_pixels", False))
        self._renderer = None
        if self.use_raw_pixels:
            from .render import PixelRenderer
            from . import urdf as _urdf
            scene = scene if scene is not None else _urdf.load_scene(getattr(opts, "model_dir", None))
            self._renderer = PixelRenderer(self.cfg.num_envs, self.repeats, getattr(opts, "num_cameras", 1),
                                           getattr(opts, "render_width", 50), getattr(opts, "render_height", 50),
                                           scene=scene, device=self.device)
            self.observation_space = spaces.Box(-float_max, float_max, self._renderer.state_shape)  # :138-141, pixel mode too
        h = ctypes.c_void_p()
        _lib.check(self._lib.bcp_create(ctypes.byref(self.cfg), ctypes.byref(h)), self._lib)
        self._h = h
        # open event log, bullet_cartpole.py:97-102: one file for the batch, an episode is written when it ends
        self.event_log = None
        if getattr(opts, "event_log_out", None):
            if self.use_raw_pixels:
                raise NotImplementedError("--event-log-out with --use-raw-pixels: the PNG render records of event.proto are not built")
            self.event_log = self._open_event_log(opts.event_log_out)
        n, dev = self.num_envs, self.device
        # the four outputs of a call are views of ONE device buffer, so that a host-side caller
        # fetches them with a single device->host copy (step_host)
        ob, rb = n * self.repeats * 14 * 4, n * 4
        self._out = torch.zeros(ob + rb + 2 * n, dtype=torch.uint8, device=dev)
        self.obs = self._out[:ob].view(torch.float32).view(n, self.repeats, 2, 7)
        self.reward = self._out[ob:ob + rb].view(torch.float32)
        self.done = self._out[ob + rb:ob + rb + n]
        self.done.fill_(1)  # no state until reset (:147-148)
        self.reason = self._out[ob + rb + n:]
        self._was_reset = False

    def _open_event_log(self, path):
        from .event_log import BatchedEventLog
        return BatchedEventLog(path, self.num_envs)

    # ---- gym no-ops, bullet_cartpole.py:169-176 ----
    def configure(self, display=None):
        pass

    def seed(self, seed=None):
        pass

    def render(self, mode=None, close=None):
        pass

    def close(self):
        if getattr(self, "event_log", None) is not None:
            self.event_log.close()
            self.event_log = None
        if getattr(self, "_h", None):
            self._lib.bcp_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ---- reset, bullet_cartpole.py:313-346 ----
    def reset(self, mask: Optional[torch.Tensor] = None, copy: bool = True) -> torch.Tensor:
        """Reset every env (mask None) or the envs where `mask` is non-zero.  Returns obs (B,R,2,7)."""
        mp = None
        if mask is not None:
            if not self._was_reset:
                raise _lib.BcpError("first reset() must reset every env")
            mask = mask.to(device=self.device, dtype=torch.uint8).contiguous()
            if mask.shape != (self.num_envs,):
                raise ValueError(f"mask must have shape ({self.num_envs},)")
            mp = ctypes.c_void_p(mask.data_ptr())
        _lib.check(self._lib.bcp_reset(self._h, mp, ctypes.c_void_p(self.obs.data_ptr()), self._stream()), self._lib)
        if mask is None:
            self.done.zero_()
        else:
            self.done.masked_fill_(mask.bool(), 0)
        self._was_reset = True
        if self.event_log is not None:  # :342-344: event_log.reset(); add_just_state(state)
            self._log_reset(None if mask is None else mask.cpu().numpy())
        return self._state(copy)

    def _log_reset(self, mask):
        self.event_log.begin(self.obs.cpu().numpy(), mask)

    def _log_step(self, actions):
        a = actions.detach().cpu().numpy().reshape(self.num_envs, -1).astype(np.float32)
        self.event_log.observe(a, self.obs.cpu().numpy(), self.reward.cpu().numpy(), self.done.cpu().numpy(), self.reason.cpu().numpy())

    def _state(self, copy: bool) -> torch.Tensor:
        if self._renderer is not None:
            px = self._renderer.render(self.obs)
            return px.clone() if copy else px
        return self.obs.clone() if copy else self.obs

    # ---- step, bullet_cartpole.py:178-275 ----
    def _actions_tensor(self, actions) -> torch.Tensor:
        n = self.num_envs
        if self.discrete_actions:
            a = torch.as_tensor(actions, device=self.device)
            if a.dtype.is_floating_point:
                raise TypeError("discrete actions must be integers in 0..4")
            a = a.to(torch.int32).reshape(-1)
            if a.numel() == 1 and n > 1:
                a = a.expand(n)
            if a.numel() != n:
                raise ValueError(f"expected {n} discrete actions, got {a.numel()}")
            return a.contiguous()
        a = torch.as_tensor(actions, dtype=torch.float32, device=self.device)
        if a.numel() != 2 * n:
            raise ValueError(f"expected continuous actions of shape ({n},2) or ({n},1,2), got {tuple(a.shape)}")
        return a.reshape(n, 2).contiguous()

    def step(self, actions, copy: bool = True):
        """-> (obs (B,R,2,7) f32, reward (B,) f32, done (B,) bool, info {'done_reason': (B,) u8 codes})."""
        if not self._was_reset:
            raise _lib.BcpError("step() before reset()")
        a = self._actions_tensor(actions)
        _lib.check(self._lib.bcp_step(self._h, ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(self.obs.data_ptr()),
                                      ctypes.c_void_p(self.reward.data_ptr()), ctypes.c_void_p(self.done.data_ptr()),
                                      ctypes.c_void_p(self.reason.data_ptr()), self._stream()), self._lib)
        if self.event_log is not None:  # :272-273: event_log.add(state, action, reward)
            self._log_step(a)
        info = {"done_reason": self.reason.clone() if copy else self.reason}
        if copy:
            return self._state(True), self.reward.clone(), self.done.bool(), info
        return self._state(False), self.reward, self.done.view(torch.bool), info  # 0/1 bytes: a view, no kernel

    # ---- host-buffer call: what a CPU-side agent loop uses (pinned buffers, one sync) ----
    def step_host(self, actions_host: torch.Tensor, auto_reset: bool = False):
        """env.step with HOST tensors through bcp_step_host: H2D of the actions, the step kernels, D2H of
        obs/reward/done into pinned buffers - the observation slab of a repeat leaves as soon as that
        repeat is complete, beside the remaining substeps - one stream sync.  Returns CPU tensors
        (views of pinned buffers, overwritten by the next call): obs (B,R,2,7) as a permuted view of
        the repeat-major buffer the copies fill, reward (B,), done (B,) uint8, info {'done_reason': (B,) uint8} - with
        auto_reset='pipelined' codes 4 / 5 tell reset progress and reset observations from transitions.
        The call's outputs live in the library's own device block: the DEVICE tensors env.obs / env.reward / env.done /
        env.reason are those of the last step() and are NOT updated by this call (use the returned host tensors).
        auto_reset=True (reset(mask=done) after the step, so that the next observation of a finished
        env is its reset observation) takes the plain path: step, one D2H, reset."""
        if self._renderer is not None:
            raise _lib.BcpError("step_host moves the low-dimensional observation; with --use-raw-pixels use step()")
        self._host_buffers()
        if auto_reset:
            self._d_act.copy_(actions_host.reshape(self._d_act.shape), non_blocking=True)
            self.step(self._d_act, copy=False)
            self._h_out.copy_(self._out, non_blocking=True)  # obs, reward, done, reason in one transfer
            self.reset(mask=self.done, copy=False)
            torch.cuda.current_stream(self.device).synchronize()
            return self._h_obs, self._h_reward, self._h_done, {"done_reason": self._h_reason}
        self.step_host_begin(actions_host)
        return self.step_host_end()

    def step_host_begin(self, actions_host: torch.Tensor):
        """First half of step_host: enqueues the H2D copy of the actions, the step and the D2H copies of its outputs on the
        env's own host-step stream and returns at once.  A host loop that drives two envs alternately
        (`a.step_host_begin(x); b.step_host_end(); b.step_host_begin(y); a.step_host_end(); ...`) has one env's outputs
        crossing PCIe - and being read on the host - while the other env steps."""
        if not self._was_reset:
            raise _lib.BcpError("step() before reset()")
        if self._renderer is not None:
            raise _lib.BcpError("step_host moves the low-dimensional observation; with --use-raw-pixels use step()")
        self._host_buffers()
        if getattr(self, "_host_pending", False):
            raise _lib.BcpError("step_host_begin() twice without step_host_end()")
        a = actions_host
        if not (type(a) is torch.Tensor and a.dtype == self._h_act.dtype and a.is_pinned() and a.is_contiguous()
                and a.numel() == self._h_act.numel()):
            self._h_act.copy_(torch.as_tensor(actions_host).reshape(self._h_act.shape))
            a = self._h_act
        vp = ctypes.c_void_p
        if self._host_stream is None:  # ordered after whatever the caller's stream did to the env (reset, set_state)
            self._host_stream = torch.cuda.Stream(device=self.device)
        self._host_stream.wait_stream(torch.cuda.current_stream(self.device))
        st = vp(self._host_stream.cuda_stream)
        _lib.check(self._lib.bcp_step_host(self._h, vp(a.data_ptr()), vp(self._h_obs_rm.data_ptr()), vp(self._h_reward.data_ptr()),
                                           vp(self._h_done.data_ptr()), vp(self._h_reason.data_ptr()), st), self._lib)
        self._host_actions, self._host_pending = a, True

    def step_host_end(self):
        """Second half of step_host: waits for the call begun by step_host_begin and returns its host tensors."""
        if not getattr(self, "_host_pending", False):
            raise _lib.BcpError("step_host_end() without step_host_begin()")
        vp = ctypes.c_void_p
        _lib.check(self._lib.bcp_step_host_finish(self._h, vp(self._h_obs_rm.data_ptr()), vp(self._h_reason.data_ptr()),
                                                  vp(self._host_stream.cuda_stream)), self._lib)
        self._host_pending = False
        torch.cuda.current_stream(self.device).wait_stream(self._host_stream)
        if self.event_log is not None:
            self.event_log.observe(self._host_actions.numpy().reshape(self.num_envs, -1), self._h_obs_rm.permute(1, 0, 2, 3).numpy(),
                                   self._h_reward.numpy(), self._h_done.numpy(), self._h_reason.numpy())
        return self._h_obs_rm.permute(1, 0, 2, 3), self._h_reward, self._h_done, {"done_reason": self._h_reason}

    def _host_buffers(self):
        if hasattr(self, "_h_out"):
            return
        n, R = self.num_envs, self.repeats
        ob, rb = self.obs.numel() * 4, n * 4
        self._h_out = torch.empty(self._out.numel(), dtype=torch.uint8).pin_memory()
        self._h_obs = self._h_out[:ob].view(torch.float32).view(self.obs.shape)
        self._h_obs_rm = self._h_out[:ob].view(torch.float32).view(R, n, 2, 7)  # the same bytes, repeat-major
        self._h_reward = self._h_out[ob:ob + rb].view(torch.float32)
        self._h_done = self._h_out[ob + rb:ob + rb + n]
        self._h_reason = self._h_out[ob + rb + n:]
        adt = torch.int32 if self.discrete_actions else torch.float32
        ashape = (n,) if self.discrete_actions else (n, 2)
        self._d_act = torch.empty(ashape, dtype=adt, device=self.device)
        self._h_act = torch.empty(ashape, dtype=adt).pin_memory()
        self._host_stream = None

    @property
    def h2d_bytes_per_step(self) -> int:
        return self.num_envs * (4 if self.discrete_actions else 8)

    @property
    def d2h_bytes_per_step(self) -> int:
        return self.num_envs * (self.repeats * 14 * 4 + 4 + 1 + 1)

    # ---- device-side random-agent protocol (README.md:76-93) ----
    def rollout_random(self, n_calls: int, actions=(0, 1, 2, 3, 4), max_episodes: int = 0):
        """n_calls rounds of R*S substeps per env in one launch, actions drawn in-kernel; finished
        episodes go to the histogram and reset in place (pipelined).  max_episodes > 0: an env
        idles after that many episodes (protocol runs of a fixed number of episodes)."""
        if not self._was_reset:
            self.reset(copy=False)
        mask = 0
        for a in actions:
            if not 0 <= int(a) <= 4:
                raise ValueError("discrete actions are 0..4")
            mask |= 1 << int(a)
        _lib.check(self._lib.bcp_rollout_random(self._h, int(n_calls), mask, int(max_episodes),
                                                ctypes.c_void_p(self.obs.data_ptr()),
                                                self._stream()), self._lib)

    def stats(self):
        """(hist int64[max_len+1], dict(episodes, env_steps, substeps, resets), reward_sum).  The histogram, episodes and
        env_steps count every finished episode / live step of step() and rollout_random() alike; substeps, resets and
        reward_sum are accumulated by rollout_random() only (step() returns the rewards to its caller)."""
        hist = np.zeros(self.max_episode_len + 1, dtype=np.int64)
        header = np.zeros(C.STATS_HEADER, dtype=np.int64)
        rsum = np.zeros(1, dtype=np.float64)
        _lib.check(self._lib.bcp_stats(self._h, hist.ctypes.data_as(ctypes.c_void_p),
                                       header.ctypes.data_as(ctypes.c_void_p),
                                       rsum.ctypes.data_as(ctypes.c_void_p), self._stream()), self._lib)
        names = ("episodes", "env_steps", "substeps", "resets")
        return hist, dict(zip(names, (int(v) for v in header))), float(rsum[0])

    def clear_stats(self):
        _lib.check(self._lib.bcp_stats_clear(self._h, self._stream()), self._lib)

    def last_solve_info(self) -> np.ndarray:
        """(B, 4) int32: contact points of the three pairs (ground-cart, cart-pole, ground-pole)
        and PGS iterations of each env's last substep (split pipeline only; bench.py's census)."""
        out = np.zeros((self.num_envs, 4), dtype=np.int32)
        _lib.check(self._lib.bcp_last_solve_info(self._h, out.ctypes.data_as(ctypes.c_void_p), out.size, self._stream()), self._lib)
        return out

    def solve_trace(self, enable: Optional[bool] = None) -> Optional[np.ndarray]:
        """solve_trace(True/False) switches the per-substep trace of the split pipeline on / off; solve_trace()
        returns (R*S, B, 4) int32 of the last step(): contact points of the three pairs and PGS iterations of every
        substep (test instrumentation, include/bcp.h bcp_solve_trace_*)."""
        if enable is not None:
            _lib.check(self._lib.bcp_solve_trace_enable(self._h, 1 if enable else 0), self._lib)
            return None
        out = np.zeros((self.repeats * self.steps_per_repeat, self.num_envs, 4), dtype=np.int32)
        _lib.check(self._lib.bcp_solve_trace_read(self._h, out.ctypes.data_as(ctypes.c_void_p), out.size, self._stream()), self._lib)
        return out

    # ---- raw physics access used by parity tests ----
    def substep(self):
        _lib.check(self._lib.bcp_substep(self._h, self._stream()), self._lib)

    def get_state(self) -> np.ndarray:
        buf = np.zeros((self.num_envs, C.STATE_DOUBLES), dtype=np.float64)
        _lib.check(self._lib.bcp_get_state(self._h, buf.ctypes.data_as(ctypes.c_void_p), buf.size, self._stream()),
                   self._lib)
        return buf

    def set_state(self, state: np.ndarray):
        buf = np.ascontiguousarray(state, dtype=np.float64)
        if buf.shape != (self.num_envs, C.STATE_DOUBLES):
            raise ValueError(f"state must have shape ({self.num_envs},{C.STATE_DOUBLES})")
        _lib.check(self._lib.bcp_set_state(self._h, buf.ctypes.data_as(ctypes.c_void_p), buf.size, self._stream()),
                   self._lib)
        self.done.copy_(torch.from_numpy(buf[:, C.S_DONE].astype(np.uint8)))
        self._was_reset = True

    state_dict = get_state
    load_state_dict = set_state

    def profile_begin(self):
        _lib.check(self._lib.bcp_profile_begin(self._h), self._lib)

    def profile_end(self):
        """-> dict(ms_prep, ms_solve, n_prep, n_solve): device time per kernel class since profile_begin."""
        out = (ctypes.c_double * 4)()
        _lib.check(self._lib.bcp_profile_end(self._h, out), self._lib)
        return {"ms_prep": out[0], "ms_solve": out[1], "n_prep": int(out[2]), "n_solve": int(out[3])}

    @property
    def solver_kind(self) -> str:
        """which pipeline step() runs: 'one-launch', 'split/row-space' or 'split/impulse-space'"""
        return ("one-launch", "split/row-space", "split/impulse-space")[int(self._lib.bcp_solver_kind(self._h))]

    @property
    def launch_count(self) -> int:
        return int(self._lib.bcp_launch_count(self._h))


class BulletCartpole(BatchedBulletCartpole):
    """Single-env drop-in with the reference's exact return types (bullet_cartpole.py:48)."""

    def __init__(self, opts, discrete_actions: bool, **kw):
        kw["num_envs"] = 1
        super().__init__(opts, discrete_actions, **kw)

    def _open_event_log(self, path):  # the reference's own writer semantics: an episode goes out at the next reset() (:48-58)
        from .event_log import EventLog
        return EventLog(path, self.use_raw_pixels)

    def _log_reset(self, mask):
        self.event_log.reset()
        self.event_log.add_just_state(self.obs[0].cpu().numpy())

    def _log_step(self, actions):  # :272-273 logs the caller's action object: an int, or the (1, 2) array
        a = self._last_action
        self.event_log.add(self.obs[0].cpu().numpy(), int(a) if self.discrete_actions else np.asarray(a, dtype=np.float32).reshape(1, 2), 1.0)

    def reset(self):
        return super().reset(copy=False)[0].cpu().numpy().copy()

    def step(self, action):
        if bool(self.done[0].item()):  # :179-181
            print("calling step after done????", file=sys.stderr)
            return self._state(False)[0].cpu().numpy().copy(), 0, True, {}
        if self.discrete_actions:
            a = np.asarray(action).reshape(-1)[:1].astype(np.int64)
            self._last_action = int(a[0])
        else:
            a = np.asarray(action, dtype=np.float32).reshape(1, 2)
            self._last_action = a
        obs, _reward, done, info = super().step(a, copy=False)
        out = {}
        code = int(info["done_reason"][0].item())
        if code:
            out["done_reason"] = C.REASON_NAMES[code]
        return obs[0].cpu().numpy().copy(), 1.0, bool(done[0].item()), out
