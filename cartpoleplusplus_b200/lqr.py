"""The fork's LQR cart-pole (random_action_agent.py:48-140) for a batch of envs on the GPU.

The fork's script drives two cart-poles with a feedback force `-(K . (state - zero_point))` on top
of a disturbance and asks a human which of two gain matrices did better.  The part of it that is
computation - the 8-state observation read off the pole, the 3.0 m / pi/4 done rule, the control
law and the one-substep step - is what `BatchedLqrCartpole` does for every env of a
`BatchedBulletCartpole` in one kernel launch per call (include/bcp_lqr.h), with a gain matrix per
env if wanted, so that a whole population of candidate gains is rolled out side by side.  The
window, the keyboard loop and the plots of that script are not rebuilt.

Differences from the fork's world that remain: one cart per env built from cart.urdf (the fork's
double_cart_*.urdf is a two-link multibody glued by a fixed joint, which the box solver here does
not model), and every env starts from the reference env's reset rather than the fork's drop from
12.5 mm.  There is no CPU path.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np
import torch

from . import _lib
from .opts import default_opts

DONE_REASONS = {0: None, 1: "cart position bounds exceeded", 2: "pole angle bounds exceeded"}  # :108-119


def lqr_opts(**overrides):
    """Env options of the fork's loop: (fx, fy) newtons applied as given, one substep per step
    (random_action_agent.py:97-106), termination left to `_check_done`, no initial push."""
    kw = dict(action_force=1.0, action_repeats=1, steps_per_repeat=1, no_termination=True, initial_force=0.0,
              max_episode_len=30 * 240 + 1)  # duration_second * steps_per_second (:226-227): never the reason to stop
    kw.update(overrides)
    return default_opts(**kw)


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


class BatchedLqrCartpole:
    X, X_DOT, Y, Y_DOT, ROLL, ROLL_DOT, PITCH, PITCH_DOT = range(8)  # :55-62

    def __init__(self, env, position_threshold: float = 3.0, angle_threshold: float = np.pi / 4,
                 initial_pole_state=None, lqr_zero_point=None):
        """env: a BatchedBulletCartpole with continuous actions (see lqr_opts).
        initial_pole_state (8,) or (B, 8): only its X entry is used, as in the fork (:127-128)."""
        if env.discrete_actions:
            raise ValueError("the LQR loop applies continuous (fx, fy) forces: build the env with discrete_actions=False")
        self.env, self._lib, self.device = env, env._lib, env.device
        n = self.num_envs = env.num_envs
        self.position_threshold, self.angle_threshold = float(position_threshold), float(angle_threshold)
        f32 = dict(dtype=torch.float32, device=self.device)
        self._x0 = None
        if initial_pole_state is not None:
            ips = torch.as_tensor(np.asarray(initial_pole_state, dtype=np.float32), **f32).reshape(-1, 8)
            self._x0 = ips[:, 0].expand(n).contiguous()
        self._zero = None if lqr_zero_point is None else torch.as_tensor(np.asarray(lqr_zero_point, dtype=np.float32), **f32).reshape(8).contiguous()
        self.pole_state = torch.zeros((n, 8), **f32)
        self.control = torch.zeros((n, 2), **f32)
        self.done = torch.zeros(n, dtype=torch.uint8, device=self.device)

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _gains(self, controller):
        K = controller if type(controller) is torch.Tensor and controller.is_cuda and controller.dtype == torch.float32 \
            else torch.as_tensor(np.asarray(controller, dtype=np.float32), device=self.device)
        K = K.contiguous()
        if tuple(K.shape) == (2, 8):
            return K, 0
        if tuple(K.shape) == (self.num_envs, 2, 8):
            return K, 1
        raise ValueError(f"controller must be (2, 8) or ({self.num_envs}, 2, 8), got {tuple(K.shape)}")

    def _launch(self, gains=None, per_env=0, disturbance=None, state=True, action=False, done=True):
        d, d_per_env = None, 0
        if disturbance is not None:
            d = torch.as_tensor(disturbance, dtype=torch.float32, device=self.device).contiguous()
            if d.numel() == 2:
                d_per_env = 0
            elif d.numel() == 2 * self.num_envs:
                d_per_env = 1
            else:
                raise ValueError(f"disturbance must be (2,) or ({self.num_envs}, 2), got {tuple(d.shape)}")
        args = _lib.BcpLqrArgs(_p(gains), per_env, _p(self._zero), _p(self._x0), _p(d), d_per_env,
                               self.position_threshold, self.angle_threshold,
                               _p(self.pole_state) if state else None, _p(self.control) if action else None,
                               _p(self.done) if done else None)
        _lib.check(self._lib.bcp_lqr_control(self.env._h, ctypes.byref(args), self._stream()), self._lib)

    # ---- the fork's methods ----
    def observe_state(self) -> torch.Tensor:  # _observe_state, :121-135
        self._launch()
        return self.pole_state

    def check_done(self):  # _check_done, :108-119 (on the state as it is now)
        self._launch()
        return self.done.bool(), {"done_reason": self.done}

    def lqr_control_forces(self, controller, disturbance=None) -> torch.Tensor:
        """-(controller . (pole_state - lqr_zero_point)) [+ disturbance] for every env -> (B, 2) (:92-95, :888-899)."""
        K, per_env = self._gains(controller)
        self._launch(K, per_env, disturbance, action=True)
        return self.control

    def step(self, action):
        """One substep with (fx, fy) applied to the cart (:97-106) -> (state (B, 8), reward (B,) zeros, done (B,) bool, info)."""
        self.env.step(action, copy=False)
        self._launch()
        return self.pole_state, torch.zeros_like(self.env.reward), self.done.bool(), {"done_reason": self.done}

    def reset(self) -> torch.Tensor:
        """The batched env can start over (the fork's class cannot, :137-140)."""
        self.env.reset(copy=False)
        return self.observe_state()

    # ---- the fork's loop body (:888-903), for a whole population ----
    def closed_loop_step(self, controller, disturbance=None):
        """control from the current state, then step(disturbance + control)."""
        return self.step(self.lqr_control_forces(controller, disturbance))

    def rollout(self, controller, n_steps: int, disturbance_fn=None, stop_when_all_done: bool = True):
        """n_steps of closed_loop_step; -> dict(steps_survived (B,) int32, first_done_reason (B,) u8, final_state (B, 8)).
        disturbance_fn(step) -> (2,) or (B, 2) force, or None."""
        n = self.num_envs
        alive = torch.ones(n, dtype=torch.bool, device=self.device)
        survived = torch.zeros(n, dtype=torch.int32, device=self.device)
        reason = torch.zeros(n, dtype=torch.uint8, device=self.device)
        K, per_env = self._gains(controller)
        for t in range(int(n_steps)):
            _, _, done, info = self.closed_loop_step(K, None if disturbance_fn is None else disturbance_fn(t))
            newly = alive & done
            reason = torch.where(newly, info["done_reason"], reason)
            alive = alive & ~done
            survived += alive.to(torch.int32)
            if stop_when_all_done and (t & 63) == 63 and not bool(alive.any()):
                break
        return {"steps_survived": survived, "first_done_reason": reason, "final_state": self.pole_state.clone()}
