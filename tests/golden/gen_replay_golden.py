"""Regenerates tests/golden/replay_golden.npz from the reference's own ReplayMemory
(run in the build container only; /root/reference does not exist on the GPU box).

replay_memory.py cannot be imported as a module here: it imports tensorflow and its
dump()/reset_from_event_log use Python-2 print statements.  The methods on the data path —
__init__, add_episode, _add, size, batch's gather — are Python-3 clean, so this script
takes THEIR SOURCE TEXT out of /root/reference/replay_memory.py at run time (nothing is
copied into the repo), executes it as a class body and drives the resulting reference
class with a seeded scenario.  The npz holds the scenario's inputs and the reference's
final arrays; tests compare the oracle restatement (CPU tier) and the CUDA ring (GPU
tier) against it bit for bit.
"""
import collections
import os
import re

import numpy as np

REF = "/root/reference/replay_memory.py"
HERE = os.path.dirname(os.path.abspath(__file__))
KEEP = ("__init__", "add_episode", "_add", "size", "current_stats")


def reference_class():
    text = open(REF).read()
    body = text[text.index("class ReplayMemory(object):"):text.index('if __name__ == "__main__":')]
    # split the class body into its methods and keep the Python-3-clean ones on the data path
    parts = re.split(r"\n(?=  def )", body)
    methods = [p for p in parts[1:] if re.match(r"  def (\w+)", p).group(1) in KEEP]
    src = "class ReplayMemory(object):\n" + "\n".join(methods) + "\n"
    ns = {"np": np, "collections": collections}
    exec(compile(src, REF, "exec"), ns)
    return ns["ReplayMemory"]


def scenario(seed, n_episodes, state_shape, action_dim, max_len):
    rng = np.random.default_rng(seed)
    eps = []
    for _ in range(n_episodes):
        n = int(rng.integers(1, max_len + 1))
        init = rng.normal(size=state_shape).astype(np.float32)
        acts = rng.uniform(-1, 1, size=(n, action_dim)).astype(np.float32)
        rews = rng.uniform(0, 2, size=n).astype(np.float32)
        st2 = (rng.normal(size=(n,) + tuple(state_shape)) * 3).astype(np.float32)
        eps.append((init, acts, rews, st2))
    return eps


def run(cls, buffer_size, state_shape, action_dim, load_factor, eps):
    rm = cls(buffer_size=buffer_size, state_shape=state_shape, action_dim=action_dim, load_factor=load_factor)
    # np.empty in the reference: zero what it has not written so that exports are comparable
    for a in (rm.state_1_idx, rm.action, rm.reward, rm.terminal_mask, rm.state_2_idx, rm.state):
        a[...] = 0
    for init, acts, rews, st2 in eps:
        rm.add_episode(init, [(acts[i], rews[i], st2[i]) for i in range(len(rews))])
    return rm


if __name__ == "__main__":
    cls = reference_class()
    out = {}
    cases = {  # name: (buffer_size, state_shape, action_dim, load_factor, seed, episodes, max_len)
        "wrap": (37, (2, 2, 7), 2, 1.5, 11, 40, 9),          # wraps several times, episodes 1..9 (cartpole state shape)
        "tiny": (3, (2, 3), 2, 2.0, 12, 7, 6),               # episodes longer than the ring (replay_memory_test.py:13)
        "notfull": (64, (2, 2, 7), 2, 1.5, 13, 5, 8),        # never wraps
    }
    for name, (bs, shape, ad, lf, seed, n_eps, max_len) in cases.items():
        eps = scenario(seed, n_eps, shape, ad, max_len)
        rm = run(cls, bs, shape, ad, lf, eps)
        out[name + "/cfg"] = np.array([bs, ad, int(bs * lf), seed, n_eps, max_len], dtype=np.int64)
        out[name + "/load_factor"] = np.array([lf])
        out[name + "/state_shape"] = np.array(shape, dtype=np.int64)
        out[name + "/ep_len"] = np.array([len(e[2]) for e in eps], dtype=np.int64)
        out[name + "/init"] = np.stack([e[0] for e in eps])
        out[name + "/actions"] = np.concatenate([e[1] for e in eps])
        out[name + "/rewards"] = np.concatenate([e[2] for e in eps])
        out[name + "/states2"] = np.concatenate([e[3] for e in eps])
        out[name + "/insert"] = np.array([rm.insert, int(rm.full), rm.size()], dtype=np.int64)
        out[name + "/state_1_idx"] = rm.state_1_idx
        out[name + "/action"] = rm.action
        out[name + "/reward"] = rm.reward
        out[name + "/terminal_mask"] = rm.terminal_mask
        out[name + "/state_2_idx"] = rm.state_2_idx
        out[name + "/state"] = rm.state
        out[name + "/free_slots"] = np.array(rm.state_free_slots, dtype=np.int32)
        st = rm.current_stats()
        out[name + "/stats"] = np.array([st.get(">add_episode", 0), st.get(">add", 0), st.get("cache_evicted_s2", 0),
                                         st["free_slots"]], dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "replay_golden.npz"), **out)
    print("written", {k: v.shape for k, v in out.items() if k.endswith("/state")})
