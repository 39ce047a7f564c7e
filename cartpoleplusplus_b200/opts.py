"""Command-line surface of the env: the reference's flags, verbatim, plus batching flags.

`add_opts(parser)` registers exactly the flags of the reference's
`bullet_cartpole.add_opts` (bullet_cartpole.py:15-40) with the same defaults, so an agent
script that did `bullet_cartpole.add_opts(parser)` keeps working unchanged.  The extra
flags (all prefixed by what they control) select batching, precision and the semantic
switches SURVEY.md section 0.4 lists between this fork and upstream.
"""
from __future__ import annotations

import argparse


def add_opts(parser: argparse.ArgumentParser) -> None:
    # --- reference flags, bullet_cartpole.py:16-40 (same names, types, defaults) ---
    parser.add_argument("--gui", action="store_true")
    parser.add_argument("--delay", type=float, default=0.0)
    parser.add_argument("--action-force", type=float, default=50.0,
                        help="magnitude of action force applied per step")
    parser.add_argument("--initial-force", type=float, default=200.0,
                        help="magnitude of initial push, in random direction")
    parser.add_argument("--no-random-theta", action="store_true")
    parser.add_argument("--action-repeats", type=int, default=2, help="number of action repeats")
    parser.add_argument("--steps-per-repeat", type=int, default=1, help="number of sim steps per repeat")
    parser.add_argument("--num-cameras", type=int, default=1, help="how many camera points to render; 1 or 2")
    parser.add_argument("--event-log-out", type=str, default=None, help="path to record event log.")
    parser.add_argument("--max-episode-len", type=int, default=200, help="maximum episode len for cartpole")
    parser.add_argument("--use-raw-pixels", action="store_true",
                        help="use raw pixels as state instead of cart/pole poses")
    parser.add_argument("--render-width", type=int, default=50)
    parser.add_argument("--render-height", type=int, default=50)
    parser.add_argument("--reward-calc", type=str, default="fixed",
                        help="'fixed': 1 per step. (angle/action shaping is commented out in the reference)")
    # --- batching / device ---
    parser.add_argument("--num-envs", type=int, default=1, help="environments simulated by this process")
    parser.add_argument("--dtype", type=str, default="fp32", choices=["fp32", "fp64"],
                        help="physics arithmetic: fp32 fast mode, fp64 parity mode")
    parser.add_argument("--seed", type=int, default=0)
    parser.add_argument("--device", type=int, default=0, help="CUDA device ordinal")
    parser.add_argument("--auto-reset", type=str, default="off", choices=["off", "pipelined"],
                        help="off: the reference - a done env stays done until reset(); pipelined: a done env "
                             "runs its 130-substep reset R*S substeps per step() call (info['done_reason'] 4/5)")
    # --- semantic switches (fork vs upstream) ---
    parser.add_argument("--no-termination", action="store_true",
                        help="fork behaviour: only the episode-length rule ends an episode "
                             "(bullet_cartpole.py:243-253 is commented out there)")
    parser.add_argument("--theta-per-reset", action="store_true",
                        help="upstream behaviour: one push direction per reset instead of one per push substep")
    parser.add_argument("--world-frame-force", action="store_true",
                        help="apply forces in WORLD_FRAME (upstream) instead of LINK_FRAME (fork)")
    parser.add_argument("--cart-friction", type=float, default=None,
                        help="override cart.urdf lateral_friction (fork: 0.0)")
    parser.add_argument("--no-carry-over-reset", action="store_true",
                        help="clear the pending force and contact caches on reset (deviation from pybullet)")
    parser.add_argument("--reset-settle-cache", action="store_true",
                        help="with --no-carry-over-reset: compute the 100 settle substeps of reset() once and replay their result "
                             "(exact in that mode: every env settles from the same state)")
    parser.add_argument("--model-dir", type=str, default=None, help="directory with ground/cart/pole .urdf")
    parser.add_argument("--cart-model", type=str, default="cart", choices=["cart", "double_cart"],
                        help="cart: models/cart.urdf (bullet_cartpole.py:155); double_cart: the fork's compound cart of "
                             "random_action_agent.py:781-794 - upper half friction 1.0 under the pole, lower half 0.0 on the ground")


def default_opts(**overrides) -> argparse.Namespace:
    """Namespace with every flag at its default (what an agent gets with no CLI args)."""
    parser = argparse.ArgumentParser()
    add_opts(parser)
    opts = parser.parse_args([])
    for k, v in overrides.items():
        if not hasattr(opts, k):
            raise AttributeError(f"unknown option {k!r}")
        setattr(opts, k, v)
    return opts
