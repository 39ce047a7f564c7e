"""CPU tier: the oracle's projected Gauss-Seidel against an INDEPENDENT solution of the same contact problem.

For frictionless contacts (the fork's cart.urdf: lateral_friction 0, so cart-ground and pole-cart rows have no friction)
one substep's normal impulses solve the LCP   w = A lam + q >= 0,  lam >= 0,  lam . w = 0   with the Delassus matrix
A = J M^-1 J^T.  The test rebuilds J, M^-1, the unconstrained velocity and the Baumgarte targets in numpy from the state
alone, solves the LCP by enumerating active sets (<= 8 contacts), and requires of the oracle's converged impulses:
complementarity residual <= 1e-6, and the same constraint-space and linear velocity change as the enumerated solution
(the impulses themselves are not unique: four coplanar contacts are redundant).  scipy's bounded least squares cross-checks the enumeration."""
import itertools

import numpy as np
import pytest
from scipy.optimize import lsq_linear

from cartpoleplusplus_b200 import config as C
from cartpoleplusplus_b200 import default_opts, make_config
from oracle import Oracle

PAIRS = ((None, C.S_CART), (C.S_CART, C.S_POLE), (None, C.S_POLE))  # (body A, body B): ground-cart, cart-pole, ground-pole


def rot(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def contact_problem(cfg, pre, post):
    """J (n x 12 over [cart w, cart v, pole w, pole v]), M^-1 (12 x 12), v_unc (12), targets t (n), oracle impulses (n)"""
    dt = cfg.dt
    bodies = {C.S_CART: (0, cfg.cart), C.S_POLE: (6, cfg.pole)}
    Minv = np.zeros((12, 12))
    v = np.zeros(12)
    for base, (o, b) in bodies.items():
        Rm = rot(pre[base + 3:base + 7])
        Minv[o:o + 3, o:o + 3] = Rm @ np.diag(1.0 / np.array(b.inertia[:])) @ Rm.T
        Minv[o + 3:o + 6, o + 3:o + 6] = np.eye(3) / b.mass
        v[o:o + 3] = pre[base + 10:base + 13]
        v[o + 3:o + 6] = pre[base + 7:base + 10] + dt * np.array(cfg.gravity[:])
        assert not pre[base + 10:base + 13].any()  # no spin: no gyroscopic term to restate
    rows, t, lam = [], [], []
    order = (2, 1, 0)  # row order of the solver (manifold_order 0); irrelevant for the converged solution
    for m in order:
        A, B = PAIRS[m]
        for k in range(int(post[C.S_MANI_CNT + m])):
            p = post[C.S_MANI + m * C.MANI_STRIDE + k * C.MANI_POINT:][:C.MANI_POINT]
            la, lb, n, dist, imp = p[0:3], p[3:6], p[6:9], p[9], p[10]
            row = np.zeros(12)
            for body, local, sign in ((A, la, 1.0), (B, lb, -1.0)):
                if body is None:
                    continue
                o = bodies[body][0]
                r = rot(pre[body + 3:body + 7]) @ local     # contact point relative to the body origin, start-of-substep pose
                row[o:o + 3] = np.cross(r, sign * n)
                row[o + 3:o + 6] = sign * n
            d = dist + cfg.linear_slop
            rows.append(row)
            t.append(-d / dt if d > 0 else -d * cfg.erp2 / dt)
            lam.append(imp)
    return np.array(rows), Minv, v, np.array(t), np.array(lam)


def solve_lcp(A, q):
    n = len(q)
    for k in range(n + 1):
        for act in itertools.combinations(range(n), k):
            lam = np.zeros(n)
            if act:
                idx = list(act)
                lam[idx] = np.linalg.lstsq(A[np.ix_(idx, idx)], -q[idx], rcond=None)[0]
            w = A @ lam + q
            if (lam >= -1e-10).all() and (w >= -1e-9).all() and abs(lam @ w) < 1e-9:
                return lam
    raise AssertionError("no complementary solution")


@pytest.mark.parametrize("seed", [0, 1])
def test_converged_normal_impulses_solve_the_lcp(seed):
    import dataclasses
    from cartpoleplusplus_b200 import urdf
    n = 48
    phys = C.PhysicsParams(linear_damping=0.0, angular_damping=0.0, warmstart_factor=0.0, solver_iterations=20000,
                           residual_threshold=0.0)
    # a 10 cm foot instead of the reference's 1 cm: the same solver code, but the tilt mode converges in thousands of
    # iterations instead of hundreds of thousands (with the 1 cm pole this test passes at 200 000 iterations, 5 minutes)
    ground, cart, pole = urdf.load_scene(None)
    pole = dataclasses.replace(pole, box_size=(0.1, 0.1, 0.5))
    cfg = make_config(default_opts(initial_force=0.0), False, num_envs=n, dtype="fp64", seed=seed, physics=phys,
                      scene=(ground, cart, pole))
    ora = Oracle(cfg, dtype="fp64")
    ora.reset()                           # the settled stack: ground-cart 4 points, cart-pole 4 points
    rng = np.random.default_rng(seed)
    st = ora.state
    for base in (C.S_CART, C.S_POLE):     # knock the bodies about: linear velocities only
        st[:, base + 7:base + 10] = rng.normal(scale=[0.02, 0.02, 0.08], size=(n, 3))
        st[:, base + 10:base + 13] = 0.0
    st[:, C.S_PEND_F:C.S_PEND_F + 6] = 0.0
    pre = st.copy()
    info = ora.substep(with_info=True)
    post = ora.state
    checked = regimes = 0
    for e in range(n):
        J, Minv, v_unc, t, lam = contact_problem(cfg, pre[e], post[e])
        if len(t) == 0:
            continue
        A = J @ Minv @ J.T
        q = J @ v_unc - t
        w = A @ lam + q
        scale = max(1.0, np.abs(lam).max())
        assert (lam >= 0).all()
        assert (w >= -1e-6).all(), (e, w.min())                       # no contact approaches faster than its target allows
        assert np.abs(np.minimum(lam, w)).max() <= 1e-6 * scale, e    # complementarity
        want = solve_lcp(A, q)
        dv, dv_want = Minv @ J.T @ lam, Minv @ J.T @ want
        # in constraint space (what the LCP determines) and on the linear velocities to 1e-6; the pole's tilt rate is seen
        # by the contacts through a short lever (half the foot), so 1e-6 m/s at the contacts is up to 2e-4 rad/s
        assert np.abs(A @ (lam - want)).max() <= 1e-6, (e, np.abs(A @ (lam - want)).max())
        lin = np.r_[3:6, 9:12]
        assert np.abs((dv - dv_want)[lin]).max() <= 1e-6 and np.abs(dv - dv_want).max() <= 2e-4, (e, np.abs(dv - dv_want).max())
        # the oracle's stored post-solve velocities are v_unc + M^-1 J^T lam
        got_v = np.r_[post[e, C.S_CART + 10:C.S_CART + 13], post[e, C.S_CART + 7:C.S_CART + 10],
                      post[e, C.S_POLE + 10:C.S_POLE + 13], post[e, C.S_POLE + 7:C.S_POLE + 10]]
        assert np.abs(got_v - (v_unc + dv)).max() <= 1e-9
        # scipy's bounded least squares on the factored problem: A = B^T B, so min |B lam + c|^2 with B^T c = q
        Bm = np.linalg.cholesky(Minv).T @ J.T
        c = np.linalg.lstsq(Bm.T, q, rcond=None)[0]
        if np.abs(Bm.T @ c - q).max() < 1e-10:                        # q in the range of B^T (it is unless A is singular there)
            res = lsq_linear(Bm, -c, bounds=(0, np.inf), tol=1e-14)
            assert np.abs(Minv @ J.T @ res.x - dv_want).max() <= 1e-5
        checked += 1
        regimes += int((lam > 1e-9).sum() < len(lam))                 # some contact released: the clamp was active
    assert checked >= n // 2 and regimes >= 4, (checked, regimes)
