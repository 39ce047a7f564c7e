"""profiles/flops_model.json: the ALGORITHMIC flops of one p.stepSimulation() by phase and contact regime.

Counted, not estimated: tests/host_harness/flopcount.cpp instantiates the parity substep (the operation sequence that is
bit-identical to the fp64 oracle's) with a scalar type that counts every add / subtract / multiply / divide / square root
it performs (1 flop each, an FMA is 2; comparisons, negations, abs, selects are 0) and attributes it to a phase.  States
come from the oracle's stationary mix of the workload (pipelined resets: live, resetting and fresh envs).  The census is
folded into a model bench.py can evaluate from the kernel's own solve records:
    flops(substep) = fixed[gc, cp, gp] + iterations * pgs_per_iteration[gc, cp, gp]
with gc / cp / gp the contact points of ground-cart, cart-pole, ground-pole.  TEST / MEASUREMENT TOOL (runs oracle/)."""
import ctypes
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cartpoleplusplus_b200 import config as C  # noqa: E402
from cartpoleplusplus_b200 import default_opts, make_config  # noqa: E402
from oracle import Oracle  # noqa: E402

PHASES = ["narrowphase", "manifold", "dynamics", "rows", "pgs", "integrate"]
HH = os.path.join(ROOT, "tests", "host_harness")

WORKLOADS = {
    # bench.py's headline workload, BASELINE.json configs[2] (reference models: cart friction 0)
    "configs2_fork_defaults": dict(initial_force=200.0, action_force=100.0, auto_reset="pipelined"),
    # the upstream-style workload bench.py reports beside it: pole stays on the cart, R*S = 10
    "upstream_style_friction": dict(initial_force=55.0, action_force=100.0, cart_friction=1.0, steps_per_repeat=5, auto_reset="pipelined"),
}


def lib():
    so = os.path.join(HH, "libfc.so")
    src = os.path.join(HH, "flopcount.cpp")
    csrc = os.path.join(ROOT, "cartpoleplusplus_b200", "csrc")
    deps = [src] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.run(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-o", so, src], check=True)
    return ctypes.CDLL(so)


def census(cfg, state):
    n = state.shape[0]
    out = np.zeros((n, len(PHASES) + 4), dtype=np.int64)
    st_out = np.zeros_like(state)
    L = lib()
    got = L.fc_substep(ctypes.byref(cfg), state.ctypes.data_as(ctypes.c_void_p), n, st_out.ctypes.data_as(ctypes.c_void_p),
                       out.ctypes.data_as(ctypes.c_void_p))
    assert got == len(PHASES)
    return out, st_out


def model_for(name, ov, n=4096, mix_steps=300, seed=0):
    cfg = make_config(default_opts(**ov), False, num_envs=n, dtype="fp64", seed=seed)
    ora = Oracle(cfg, dtype="fp64")
    ora.reset()
    rng = np.random.default_rng(seed)
    for _ in range(mix_steps):
        ora.step(rng.uniform(-1, 1, size=(n, 2)).astype(np.float32))
    state = ora.state.copy()
    counts, st_out = census(cfg, state)
    # the counted program IS the oracle's substep: same state afterwards, bit for bit
    ora.state[:] = state
    info = ora.substep(with_info=True)
    assert np.array_equal(st_out[:, :C.S_MANI_CNT], ora.state[:, :C.S_MANI_CNT]), "the counting instantiation left the oracle"
    assert np.array_equal(counts[:, -1], info[:, 0]) and np.array_equal(counts[:, -4:-1], info[:, 3:6])
    gc, cp, gp, it = (counts[:, len(PHASES) + k] for k in range(4))
    pgs = counts[:, PHASES.index("pgs")].astype(np.float64)
    fixed = counts[:, :len(PHASES)].astype(np.float64)
    fixed[:, PHASES.index("pgs")] = 0
    regimes = {}
    pred = np.zeros(n)
    for key in sorted(set(zip(gc.tolist(), cp.tolist(), gp.tolist()))):
        m = (gc == key[0]) & (cp == key[1]) & (gp == key[2])
        per_it = float((pgs[m] / np.maximum(it[m], 1)).mean())
        assert np.abs(pgs[m] / np.maximum(it[m], 1) - per_it).max() <= 0.25 * max(per_it, 1.0) + 1e-9, key  # linear in the iterations up to the cone-clip branch
        f = fixed[m].mean(axis=0)
        regimes["%d,%d,%d" % key] = {"envs": int(m.sum()), "fixed": {p: float(v) for p, v in zip(PHASES, f) if p != "pgs"},
                                     "fixed_total": float(f.sum()), "pgs_per_iteration": per_it, "mean_iterations": float(it[m].mean())}
        pred[m] = f.sum() + it[m] * per_it
    total = counts[:, :len(PHASES)].sum(axis=1).astype(np.float64)
    err = abs(pred.sum() - total.sum()) / total.sum()
    assert err < 0.01, err
    by_phase = counts[:, :len(PHASES)].mean(axis=0)
    return {"workload": ov, "envs": n, "mean_flops_per_substep": float(total.mean()),
            "mean_by_phase": {p: float(v) for p, v in zip(PHASES, by_phase)}, "mean_iterations": float(it.mean()),
            "model_error": float(err), "regimes": regimes,
            "fallback": {"fixed_total": float(fixed.sum(axis=1).mean()), "pgs_per_iteration": float((pgs / np.maximum(it, 1)).mean())}}


if __name__ == "__main__":
    blob = {"convention": "add, subtract, multiply, divide, sqrt = 1 flop (FMA = 2); comparisons, negation, abs, select = 0; counted on the "
                          "parity substep (tests/host_harness/flopcount.cpp), the oracle's own operation sequence",
            "model": "flops(substep) = fixed_total[gc,cp,gp] + iterations * pgs_per_iteration[gc,cp,gp]",
            "workloads": {name: model_for(name, ov) for name, ov in WORKLOADS.items()}}
    path = os.path.join(ROOT, "profiles", "flops_model.json")
    with open(path, "w") as fh:
        json.dump(blob, fh, indent=1)
    for name, m in blob["workloads"].items():
        print(name, "mean flops/substep %.0f" % m["mean_flops_per_substep"], {k: round(v) for k, v in m["mean_by_phase"].items()},
              "iterations %.1f" % m["mean_iterations"], "regimes", len(m["regimes"]), "model error %.2e" % m["model_error"])
