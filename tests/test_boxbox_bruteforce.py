"""CPU tier: the oracle's narrowphase (dBoxBox2 restated, oracle/bcp_oracle.c box_box) against an INDEPENDENT brute-force
statement in numpy on 10 000 random poses of the scene's three box pairs:
  * separation per candidate axis from projecting all 16 vertices (no closed-form radii),
  * face contact: the incident face clipped against the reference face's side planes with a generic Sutherland-Hodgman
    (no rectangle / quad special case), vertices kept where depth >= 0,
  * edge contact: closest points of the two supporting lines from a 2 x 2 solve.
The selection rule between axes (face axes first, an edge axis must beat them by the 1.05 fudge factor) is the
algorithm's specification and is shared; every number is computed differently.  Contact sets must agree within 1e-9."""
import itertools

import numpy as np
import pytest

from oracle import oracle as O

H = {"ground": np.array([1.5, 1.5, 0.05]), "cart": np.array([0.1, 0.1, 0.025]), "pole": np.array([0.005, 0.005, 0.25])}
SIGNS = np.array(list(itertools.product((-1, 1), repeat=3)), dtype=float)


def rot(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def verts(c, Rm, h):
    return c + (SIGNS * h) @ Rm.T


def clip_polygon(poly, planes):
    """Sutherland-Hodgman: keep the part of `poly` (k x 3) with n.x <= d for every (n, d)"""
    for n, d in planes:
        out = []
        k = len(poly)
        for i in range(k):
            a, b = poly[i], poly[(i + 1) % k]
            da, db = n @ a - d, n @ b - d
            if da <= 0:
                out.append(a)
            if (da < 0 < db) or (db < 0 < da):
                out.append(a + (b - a) * (da / (da - db)))
        poly = out
        if not poly:
            break
    return np.array(poly).reshape(-1, 3)


def brute(cA, RA, hA, cB, RB, hB):
    """-> (kind, axis index, set of (point on B, depth)) or None when separated / ambiguous"""
    VA, VB = verts(cA, RA, hA), verts(cB, RB, hB)
    d = cB - cA
    cands = []  # (separation, unit axis pointing A -> B, code)
    def sep(L):
        rA = np.abs((VA - cA) @ L).max()
        rB = np.abs((VB - cB) @ L).max()
        return abs(d @ L) - (rA + rB)
    for i in range(3):
        cands.append((sep(RA[:, i]), RA[:, i] * np.sign(d @ RA[:, i] or 1.0), 1 + i))
    for j in range(3):
        cands.append((sep(RB[:, j]), RB[:, j] * np.sign(d @ RB[:, j] or 1.0), 4 + j))
    if max(c[0] for c in cands) > 0:
        return None
    # dBoxBox2's selection, which is part of the specification: the six face axes in order, the largest separation wins
    # (strictly); then the nine edge axes in order, each of which must beat the CURRENT best by the 1.05 fudge factor.
    # It also adds 1e-5 to every |R_ij| before the edge tests, inflating the projected radii by up to
    # 1e-5 x (sum of extents) / |ua x ub|: a comparison closer than that is not decidable from here.
    face = sorted(c[0] for c in cands)
    if face[-1] - face[-2] < 1e-6:
        return "ambiguous"
    best = max(cands, key=lambda c: c[0])
    for i in range(3):
        for j in range(3):
            L = np.cross(RA[:, i], RB[:, j])
            ln = np.linalg.norm(L)
            if ln < 1e-3:
                continue  # near-parallel edges: damped by the same 1e-5, not comparable
            L = L / ln
            s = sep(L)
            if s > 1e-9:
                return None
            margin = max(1.6e-5 * (hA.sum() + hB.sum()) / ln, 1e-6)
            if abs(s * 1.05 - best[0]) < margin:
                return "ambiguous"
            if s * 1.05 > best[0]:
                best = (s, L * np.sign(d @ L or 1.0), 7 + 3 * i + j)
    s, n, code = best
    if code >= 7:
        i, j = (code - 7) // 3, (code - 7) % 3
        pa = VA[np.argmax(VA @ n)]      # the vertex of A furthest along the normal lies on the supporting edge
        pb = VB[np.argmin(VB @ n)]
        ua, ub = RA[:, i], RB[:, j]
        # closest points of pa + a ua and pb + b ub
        M = np.array([[ua @ ua, -(ua @ ub)], [ua @ ub, -(ub @ ub)]])
        rhs = np.array([(pb - pa) @ ua, (pb - pa) @ ub])
        a, b = np.linalg.solve(M, rhs)
        return ("edge", code, [(pb + b * ub, -s)])
    ref_is_A = code <= 3
    cR, RR, hR, cI, RI, hI = (cA, RA, hA, cB, RB, hB) if ref_is_A else (cB, RB, hB, cA, RA, hA)
    nR = n if ref_is_A else -n          # outward normal of the reference face
    k = (code - 1) % 3
    # incident face: the face of the other box whose outward normal is most opposed to nR
    dots = RI.T @ nR
    a = int(np.argmax(np.abs(dots)))
    sgn = -np.sign(dots[a])
    centre = cI + sgn * hI[a] * RI[:, a]
    a1, a2 = [x for x in range(3) if x != a]
    poly = np.array([centre + s1 * hI[a1] * RI[:, a1] + s2 * hI[a2] * RI[:, a2] for s1, s2 in ((1, 1), (1, -1), (-1, -1), (-1, 1))])
    k1, k2 = [x for x in range(3) if x != k]
    planes = []
    for kk in (k1, k2):
        for sg in (1, -1):
            nn = sg * RR[:, kk]
            planes.append((nn, nn @ cR + hR[kk]))
    poly = clip_polygon(poly, planes)
    pts = []
    for v in poly:
        dep = hR[k] - nR @ (v - cR)
        if dep >= 0:
            pts.append((v if ref_is_A else v - n * dep, dep))
    return ("face", code, pts)


def quat(rng, tilt):
    ax = rng.normal(size=3)
    ax /= np.linalg.norm(ax)
    ang = rng.normal() * tilt
    return np.r_[ax * np.sin(ang / 2), np.cos(ang / 2)]


def sample(rng, pair):
    a, b = pair
    hA, hB = H[a], H[b]
    qA = np.array([0, 0, 0, 1.0]) if a == "ground" else quat(rng, 0.3)
    qB = quat(rng, rng.choice([0.02, 0.3, 1.5]))
    cA = np.zeros(3) if a == "ground" else rng.uniform(-0.2, 0.2, size=3)
    # B resting on top of A, shifted sideways, pushed in or lifted a little
    RBm = rot(qB)
    reach = np.abs(RBm[2, :]) @ hB
    cB = cA + np.r_[rng.uniform(-1.0, 1.0, size=2) * (hA[:2] + 0.5 * hB[:2]), hA[2] + reach - rng.uniform(-0.002, 0.012)]
    if a != "ground":
        cB = cA + rot(qA) @ (cB - cA)
    return cA, qA, hA, cB, qB, hB


def test_narrowphase_against_brute_force():
    rng = np.random.default_rng(42)
    pairs = [("ground", "cart"), ("cart", "pole"), ("ground", "pole")]
    n_checked = {"face": 0, "edge": 0, "separated": 0}
    n_multi = 0
    worst = 0.0
    for t in range(10000):
        cA, qA, hA, cB, qB, hB = sample(rng, pairs[t % 3])
        n, got = O.box_box(cA, qA, hA, cB, qB, hB)
        want = brute(cA, rot(qA), hA, cB, rot(qB), hB)
        if want == "ambiguous":
            continue
        if want is None:
            if n == 0:
                n_checked["separated"] += 1
            else:   # dBoxBox2 may still report touching boxes within its 1e-5 edge fudge: only shallow ones
                assert np.abs(got[:, 6]).max() < 1e-4
            continue
        kind, code, pts = want
        min_sin = [1.0]
        if kind == "edge":
            i, j = (code - 7) // 3, (code - 7) % 3
            min_sin = [np.linalg.norm(np.cross(rot(qA)[:, i], rot(qB)[:, j]))]
        if not pts:
            assert n == 0
            continue
        assert n >= 1, (t, kind, code)
        P = np.array([p for p, _ in pts])
        D = np.array([dd for _, dd in pts])
        # the oracle's normal is "on B", pointing from B towards A; depth is the signed distance (negative inside)
        for row in got:
            j = int(np.argmin(np.linalg.norm(P - row[3:6], axis=1)))
            # an edge contact's depth carries dBoxBox2's own 1e-5 fudge on |R_ij| (projected radii grow by 1e-5 x extents)
            dtol = 1e-9 if kind == "face" else 2e-5 * (hA.sum() + hB.sum()) / min_sin[0]
            err = np.linalg.norm(P[j] - row[3:6])
            derr = abs(D[j] + row[6])
            worst = max(worst, err if kind == "edge" else max(err, derr))
            assert err <= 1e-9 and derr <= dtol, (t, kind, code, err, derr)
        if len(pts) <= 4:
            assert n == len(pts), (t, kind, code, n, len(pts))
        else:   # more candidates than the manifold takes: four of them, the deepest among them
            assert n == 4 and abs(np.abs(got[:, 6]).max() - D.max()) <= 1e-9
            n_multi += 1
        assert len({tuple(np.round(r[3:6], 9)) for r in got}) == n    # no duplicates
        n_checked[kind] += 1
    assert n_checked["face"] > 4500 and n_checked["edge"] > 150 and n_checked["separated"] > 300, n_checked
    assert n_multi > 10
    print(n_checked, "more than four candidates:", n_multi, "worst error", worst)
