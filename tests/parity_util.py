"""Shared by tests/test_gpu_pipeline_parity.py, __graft_entry__.smoke() and scripts/fp32_tail_report.py:
one env.step of the fp32 CUDA path against the fp64 oracle from identical states, with every env that leaves
the oracle by more than the north_star tolerance (1e-4 absolute on poses) CLASSIFIED by the discrete decision
that differs.  TEST TOOL: imports oracle/.

A contact simulator is a discontinuous map: SAT axis choice, `dep >= 0` culls, the 4-point manifold reduction,
the clamp / cone-clip pattern of every Gauss-Seidel iteration, the residual early exit and the done thresholds are
decisions, and an input within rounding of such a boundary may fall either way.  north_star's 1e-4 is therefore
enforced as: every env is within 1e-4 of the fp64 oracle UNLESS
 (A) the kernel's own discrete record differs from the oracle's for that env:
     A1 contact counts of some substep (bcp_solve_trace vs the oracle's per-substep info),
     A2 which points the three manifolds hold after the call (get_state vs the oracle's state),
     A3 PGS iteration counts (the residual early exit), AND the oracle forced to run the kernel's counts lands
        within 1e-4 of the kernel - the early exit alone accounts for the difference; or
 (B) the fp64 oracle itself changes a recorded decision (contact counts, SAT codes, iteration counts, the hash of the
     solver's clamp pattern) or moves by more than 1e-4 when its input is perturbed by 1, 4 or 16 fp32 ulps (what
     rounding accumulates to over the ~10^3 dependent operations of a substep).
 (C) (last resort, no decision involved) the state is ill-conditioned for any float evaluation: the oracle's own
     statements compiled in IEEE float - none of the kernel's code - leave the fp64 oracle by more than 1e-4 too and
     land within 1e-5 of the kernel (seen on a pole spinning on one edge through 50 unconverged iterations).
An env above 1e-4 with none of these is a FAILURE (`unexplained`).  Measured at configs[2], 65 536 envs: 0.16 % of the envs
are above 1e-4, 104 of 106 by (A), the other two by (B) (profiles/r2_pipeline_parity.json).
"""
from __future__ import annotations

import numpy as np

from cartpoleplusplus_b200 import config as C

TOL = 1e-4
POSE = np.r_[C.S_CART:C.S_CART + 7, C.S_POLE:C.S_POLE + 7]


def pose_dev(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """max |pose difference| per env over cart and pole (pos3, quat4) of two canonical state arrays"""
    return np.abs(a[:, POSE] - b[:, POSE]).max(axis=1)


def done_margin(cfg, state: np.ndarray) -> np.ndarray:
    """distance of the pole from the nearest done threshold of bullet_cartpole.py:244-257 (position 3.0 m, angle 0.35 rad)"""
    x, y = state[:, C.S_POLE], state[:, C.S_POLE + 1]
    q = state[:, C.S_POLE + 3:C.S_POLE + 7]
    sarg = np.clip(-2.0 * (q[:, 0] * q[:, 2] - q[:, 3] * q[:, 1]), -1, 1)
    roll = np.arctan2(2.0 * (q[:, 1] * q[:, 2] + q[:, 3] * q[:, 0]), q[:, 3] ** 2 - q[:, 0] ** 2 - q[:, 1] ** 2 + q[:, 2] ** 2)
    pitch = np.arcsin(sarg)
    m = np.minimum(np.abs(np.abs(x) - cfg.pos_threshold), np.abs(np.abs(y) - cfg.pos_threshold))
    m = np.minimum(m, np.abs(np.abs(roll) - cfg.angle_threshold))
    return np.minimum(m, np.abs(np.abs(pitch) - cfg.angle_threshold))


def manifold_members(rec: np.ndarray):
    """which contact points the three pair caches hold: per pair the sorted local points on body A, to 1 mm"""
    out = []
    for m in range(3):
        c = int(rec[C.S_MANI_CNT + m])
        base = C.S_MANI + m * C.MANI_STRIDE
        out.append(tuple(sorted(tuple(np.round(rec[base + k * C.MANI_POINT:base + k * C.MANI_POINT + 3] * 1000).astype(int)) for k in range(c))))
    return tuple(out)


def perturbed(state: np.ndarray, rng, ulps: float = 1.0) -> np.ndarray:
    """copies of canonical records with every float field moved by up to `ulps` fp32 ulps (quaternions renormalised,
    counters untouched, empty manifold slots stay zero)"""
    out = state.copy()
    nf = C.S_MANI_CNT
    eps = ulps * 2.0 ** -24
    out[:, :nf] *= 1.0 + rng.uniform(-eps, eps, size=(state.shape[0], nf))
    for b in (C.S_CART, C.S_POLE):
        q = out[:, b + 3:b + 7]
        q /= np.linalg.norm(q, axis=1, keepdims=True)
    return out


def classify_step(env, ora, state: np.ndarray, actions: np.ndarray, rng=None, n_perturb: int = 8, tol: float = TOL) -> dict:
    """One env.step of `env` (fp32 CUDA, through the C-ABI) and of `ora` (fp64 oracle) from `state`.  Returns a report
    dict; report['unexplained'] lists the envs that break the rule in the module docstring."""
    import torch
    rng = rng or np.random.default_rng(0)
    n = env.num_envs
    cfg = ora.cfg
    traced = env.solver_kind != "one-launch"
    env.set_state(state)
    ora.state[:] = state
    if traced:
        env.solve_trace(True)
    a_dev = torch.from_numpy(actions).to(env.device)
    obs, rew, done, info = env.step(a_dev)
    torch.cuda.synchronize()
    trace = env.solve_trace() if traced else None   # (RS, n, 4): contacts gc, cp, gp, iterations
    st_gpu = env.get_state()
    obs, rew, done = obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy()
    reason = info["done_reason"].cpu().numpy()
    obs_o, rew_o, done_o, reason_o, info_o = ora.step(actions, with_info=True)
    st_o = ora.state.copy()
    ran = info_o[:, :, 0] >= 0                     # substeps the env actually ran in this call
    dev = pose_dev(st_gpu, st_o)
    live = reason_o <= 3
    moved = rew_o > 0                              # live envs wrote a new observation
    dev_obs = np.zeros(n)
    if moved.any():
        dev_obs[moved] = np.abs(obs[moved].astype(np.float64) - obs_o[moved].astype(np.float64)).reshape(int(moved.sum()), -1).max(axis=1)
    fresh = reason_o == C.REASON_FRESH
    if fresh.any():
        dev_obs[fresh] = np.abs(obs[fresh].astype(np.float64) - obs_o[fresh].astype(np.float64)).reshape(int(fresh.sum()), -1).max(axis=1)
    dev = np.maximum(dev, dev_obs)

    # (A) the kernel's own discrete record against the oracle's
    cls = {}
    if traced:
        t = np.transpose(trace, (1, 0, 2))          # (n, RS, 4)
        cnt_diff = ((t[:, :, :3] != info_o[:, :, 3:6]) & ran[:, :, None]).any(axis=(1, 2))
        it_diff = ((t[:, :, 3] != info_o[:, :, 0]) & ran).any(axis=1)
    else:
        t = None
        cnt_diff = (st_gpu[:, C.S_MANI_CNT:C.S_MANI_CNT + 3] != st_o[:, C.S_MANI_CNT:C.S_MANI_CNT + 3]).any(axis=1)
        it_diff = np.zeros(n, dtype=bool)
    out = np.flatnonzero(dev > tol)
    member_diff = np.zeros(n, dtype=bool)
    for e in out:
        member_diff[e] = manifold_members(st_gpu[e]) != manifold_members(st_o[e])
    forced_ok = np.zeros(n, dtype=bool)
    if traced and (it_diff[out]).any():
        force = np.zeros((n, t.shape[1]), dtype=np.int32)
        sel = out[it_diff[out]]
        force[sel] = np.where(ran[sel], t[sel, :, 3], 0)
        ora.state[:] = state
        ora.step(actions, force_iterations=force)
        forced_ok[sel] = pose_dev(ora.state[sel], st_gpu[sel]) <= tol
        ora.state[:] = st_o
    # (B) is the fp64 oracle itself discontinuous at (accumulated) fp32 resolution at this state?
    sens = np.zeros(n, dtype=bool)
    spread = np.zeros(n)
    need = out[~(cnt_diff[out] | member_diff[out] | forced_ok[out])]
    if need.size:
        # whole-batch calls keep env_id_offset + index == the env's global id (the reset pushes are keyed by it)
        for ulps in (1, 1, 4, 4, 16, 16, 16, 16)[:n_perturb]:
            ps = state.copy()
            ps[need] = perturbed(state[need], rng, ulps)
            ora.state[:] = ps
            r = ora.step(actions, with_info=True)
            d = pose_dev(ora.state[need], st_o[need])
            spread[need] = np.maximum(spread[need], d)
            sens[need] |= (d > tol) | (r[4][need] != info_o[need]).reshape(need.size, -1).any(axis=1)
        ora.state[:] = st_o
    # (C) what is left: is the state ill-conditioned for ANY float evaluation?  The oracle's own statements compiled
    # in IEEE float (gcc, -ffp-contract=off, none of the kernel's code) must then land where the kernel lands.
    float_agrees = np.zeros(n, dtype=bool)
    left = need[~sens[need]] if need.size else need
    if left.size:
        from oracle import Oracle
        o32 = Oracle(cfg, dtype="fp32")
        o32.state[:] = state
        o32.step(actions)
        float_agrees[left] = (pose_dev(o32.state[left], st_gpu[left]) <= 0.1 * tol) & (pose_dev(o32.state[left], st_o[left]) > tol)
    for e in out:
        why = []
        if cnt_diff[e]:
            why.append("A1-contact-count")
        if member_diff[e]:
            why.append("A2-manifold-members")
        if forced_ok[e]:
            why.append("A3-early-exit-iteration")
        if sens[e]:
            why.append("B-oracle-discontinuous-within-16ulp")
        if float_agrees[e]:
            why.append("C-ill-conditioned-float-oracle-lands-on-the-kernel")
        cls[int(e)] = why
    unexplained = [e for e, why in cls.items() if not why]

    # bookkeeping: reward / done / reason bit-exact wherever the trajectory agrees and the pole is not on a threshold
    ok = dev <= tol
    book_bad = ok & ((rew != rew_o) | (done != done_o) | (reason != reason_o))
    margin = done_margin(cfg, st_o)
    book_unexplained = np.flatnonzero(book_bad & ~(live & (margin <= 10 * tol)))
    counts = {}
    for why in cls.values():
        key = "+".join(why) if why else "none"
        counts[key] = counts.get(key, 0) + 1
    return {
        "n": n, "solver": env.solver_kind,
        "phases": {"live": int(live.sum()), "resetting": int((reason_o == C.REASON_RESETTING).sum()), "fresh": int(fresh.sum())},
        "median": float(np.median(dev)), "q99": float(np.quantile(dev, 0.99)), "max": float(dev.max()),
        "above_tol": int(out.size), "above_tol_frac": float(out.size) / n,
        "classes": counts, "unexplained": [int(e) for e in unexplained],
        "iteration_count_differs": int(it_diff.sum()), "contact_count_differs": int(cnt_diff.sum()),
        "bookkeeping_mismatch": int(book_bad.sum()), "bookkeeping_unexplained": [int(e) for e in book_unexplained],
        "dev": dev, "spread": spread,
    }


def stationary_mix(env, calls: int = 400) -> np.ndarray:
    """canonical records of the batch after `calls` random-agent calls with pipelined resets: live, resetting and fresh
    envs in the proportions the bench times"""
    env.reset(copy=False)
    env.rollout_random(calls)
    return env.get_state()
