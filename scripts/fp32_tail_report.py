"""How much of the fp32-vs-fp64-oracle tail is the approximate division / square root / flush-to-zero of the fast unit?

Runs tests/parity_util.classify_step (one env.step of the timed pipeline from the stationary mix at configs[2]) once
with the product library (bcp_f32.cu compiled with -prec-div=false -prec-sqrt=false -ftz=true) and once, in a second
process, with `python cartpoleplusplus_b200/build.py --precise` (IEEE division and square root, denormals kept), from the
same states and actions.  Writes profiles/r2_fp32_tail.json.  TEST TOOL (imports oracle/)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def one(n):
    import numpy as np
    import parity_util as PU
    from cartpoleplusplus_b200 import default_opts
    from cartpoleplusplus_b200.env import BatchedBulletCartpole
    from oracle import Oracle
    ov = dict(initial_force=200.0, action_force=100.0, auto_reset="pipelined")
    env = BatchedBulletCartpole(default_opts(**ov), discrete_actions=False, num_envs=n, device=0, dtype="fp32", seed=17)
    ora = Oracle(env.cfg, dtype="fp64")
    cache = os.path.join("/tmp", f"bcp_tail_state_{n}.npz")
    if os.path.exists(cache):  # the second process starts from the first one's states
        z = np.load(cache)
        state, acts = z["state"], z["acts"]
    else:
        state = PU.stationary_mix(env, 400)
        acts = np.random.default_rng(23).uniform(-1, 1, size=(n, 2)).astype(np.float32)
        os.makedirs(os.path.dirname(cache), exist_ok=True)
        np.savez(cache, state=state, acts=acts)
    rep = PU.classify_step(env, ora, state, acts, rng=np.random.default_rng(1))
    dev = rep.pop("dev")
    rep.pop("spread")
    rep["quantiles"] = {q: float(np.quantile(dev, float(q))) for q in ("0.5", "0.9", "0.99", "0.999", "0.9999")}
    rep["above_1e-5"] = int((dev > 1e-5).sum())
    return rep


if __name__ == "__main__":
    n = int(sys.argv[sys.argv.index("--envs") + 1]) if "--envs" in sys.argv else 65536
    if "--child" in sys.argv:
        print("REPORT " + json.dumps(one(n)))
        sys.exit(0)
    from cartpoleplusplus_b200 import build
    out = {}
    for name, lib in (("product (-prec-div=false -prec-sqrt=false -ftz=true)", build.build()), ("precise", build.build(precise=True))):
        env = dict(os.environ, BCP_LIB_PATH=lib)
        r = subprocess.run([sys.executable, __file__, "--child", "--envs", str(n)], env=env, capture_output=True, text=True)
        line = [l for l in r.stdout.splitlines() if l.startswith("REPORT ")]
        if not line:
            raise SystemExit(r.stdout + r.stderr)
        out[name] = json.loads(line[0][7:])
    path = os.path.join(ROOT, "profiles", "r2_fp32_tail.json")
    with open(path, "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out, indent=1))
