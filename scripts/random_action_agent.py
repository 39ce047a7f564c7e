#!/usr/bin/env python
"""Upstream's random_action_agent CLI (README.md:76-93), restored over the batched B200 env.

  python scripts/random_action_agent.py --initial-force=55 --actions="0,1,2,3,4" --num-eval=100 | python deciles.py

prints one episode length per line (sorted: the device keeps a histogram, not the order);
--deciles prints the deciles.py summary directly.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cartpoleplusplus_b200 import add_opts  # noqa: E402
from cartpoleplusplus_b200 import protocol  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--actions", type=str, default="0,1,2,3,4", help="comma separated discrete actions to pick from")
    ap.add_argument("--num-eval", type=int, default=100)
    ap.add_argument("--action-type", type=str, default="discrete", choices=["discrete", "continuous"])
    ap.add_argument("--deciles", action="store_true", help="print the deciles instead of the episode lengths")
    add_opts(ap)
    opts = ap.parse_args()
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    if opts.num_envs == 1:
        opts.num_envs = min(opts.num_eval, 65536)
    env = BatchedBulletCartpole(opts, discrete_actions=opts.action_type == "discrete")
    actions = tuple(int(a) for a in opts.actions.split(","))
    hist, header = protocol.run_random_agent(env, opts.num_eval, actions)
    if opts.deciles:
        print(protocol.deciles_from_histogram(hist))
    else:
        for length in protocol.lengths_from_histogram(hist):
            print(int(length))
    print(f"episodes {header['episodes']} env_steps {header['env_steps']} resets {header['resets']}", file=sys.stderr)


if __name__ == "__main__":
    main()
