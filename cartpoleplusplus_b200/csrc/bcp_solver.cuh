// bcp_solver.cuh — rigid-body dynamics of one p.stepSimulation().
//
// Device-side statement of btMultiBody (0 links, floating base) forward dynamics,
// btMultiBodyConstraintSolver's contact rows and projected Gauss-Seidel sweep, and
// stepPositionsMultiDof.  [BULLET-RECALL] bullet3 src/BulletDynamics/Featherstone.
#pragma once
#include "bcp_collide.cuh"

namespace bcp {

// solver-side view of a body (index 0 ground, 1 cart, 2 pole)
template <typename T> struct SBody {
  int dyn;
  T mass, inv_mass, I[3];
  T R[9], x[3];
  T vel[6];   // [omega, v], world
  T dvel[6];  // delta velocities accumulated by the sweep
};
template <typename T> struct Row {
  int a, b, fidx;
  T jacA[6], jacB[6], unitA[6], unitB[6];
  T jacDiagABInv, rhs, cfm, applied, lower, upper, friction;
};

// computeAccelerationsArticulatedBodyAlgorithmMultiDof, num_links == 0
template <typename T> BCP_HD void forward_dynamics(const Params<T>& P, SBody<T>& b, const T F_ext[3], const T T_ext[3]) {
  T F[3], wl[3], vl[3], za[3], zl[3], t[3], ra[3], rl[3], sa[3], sl[3], od[3], vd[3], wxv[3];
  for (int i = 0; i < 3; ++i) F[i] = F_ext[i] + P.gravity[i] * b.mass;
  matT_mul_vec(b.R, b.vel, wl);
  matT_mul_vec(b.R, b.vel + 3, vl);
  matT_mul_vec(b.R, T_ext, t);
  for (int i = 0; i < 3; ++i) za[i] = -t[i];
  matT_mul_vec(b.R, F, t);
  for (int i = 0; i < 3; ++i) zl[i] = -t[i];
  T d = dot3(wl, wl);
  T nrm = (d > Lim<T>::eps()) ? bsqrt(d) : T(0.0);
  for (int i = 0; i < 3; ++i) za[i] += ((T(1.0) * b.I[i]) * wl[i]) * (P.ka + P.ka * nrm);
  d = dot3(vl, vl);
  nrm = (d > Lim<T>::eps()) ? bsqrt(d) : T(0.0);
  for (int i = 0; i < 3; ++i) zl[i] += ((T(1.0) * b.mass) * vl[i]) * (P.kl + P.kl * nrm);
  if (P.gyro_term) {
    T Iw[3], g[3];
    for (int i = 0; i < 3; ++i) Iw[i] = b.I[i] * wl[i];
    cross3(wl, Iw, g);
    for (int i = 0; i < 3; ++i) za[i] += g[i];
  }
  cross3(wl, vl, wxv);
  for (int i = 0; i < 3; ++i) zl[i] += b.mass * wxv[i];
  for (int i = 0; i < 3; ++i) ra[i] = za[i] / b.I[i];
  for (int i = 0; i < 3; ++i) rl[i] = zl[i] * b.inv_mass;
  for (int i = 0; i < 3; ++i) { sa[i] = -ra[i]; sl[i] = -rl[i] + wxv[i]; }
  mat_mul_vec(b.R, sa, od);
  mat_mul_vec(b.R, sl, vd);
  for (int i = 0; i < 3; ++i) {
    b.vel[i] = clampv(b.vel[i] + od[i] * P.dt, -P.vmax, P.vmax);
    b.vel[3 + i] = clampv(b.vel[3 + i] + vd[i] * P.dt, -P.vmax, P.vmax);
  }
}

// fillContactJacobianMultiDof + calcAccelerationDeltasMultiDof, one side of a row
template <typename T> BCP_HD void fill_side(const SBody<T>& b, const T pos[3], const T n[3], T jac[6], T unit[6]) {
  T r[3], t[3], u[3];
  for (int i = 0; i < 3; ++i) r[i] = pos[i] - b.x[i];
  cross3(r, n, jac);
  for (int i = 0; i < 3; ++i) jac[3 + i] = n[i];
  if (!b.dyn) {
    for (int i = 0; i < 6; ++i) unit[i] = 0;
    return;
  }
  matT_mul_vec(b.R, jac, t);
  for (int i = 0; i < 3; ++i) u[i] = t[i] / b.I[i];
  mat_mul_vec(b.R, u, unit);
  matT_mul_vec(b.R, jac + 3, t);
  for (int i = 0; i < 3; ++i) u[i] = t[i] * b.inv_mass;
  mat_mul_vec(b.R, u, unit + 3);
}

// setupMultiBodyContactConstraint
template <typename T>
BCP_HD void setup_row(const Params<T>& P, Row<T>& row, const SBody<T>* bodies, int ia, int ib, const T posA[3],
                      const T posB[3], const T axis[3], T distance_in, T mu, T comb_rest, bool is_friction) {
  T denom0 = 0, denom1 = 0, rel_vel = 0, restitution = 0, distance = 0, naxis[3];
  row.a = ia;
  row.b = ib;
  const T cfm = is_friction ? P.cfm_friction : P.cfm_normal;
  const T erp = is_friction ? P.friction_erp : P.erp2;
  fill_side(bodies[ia], posA, axis, row.jacA, row.unitA);
  for (int i = 0; i < 3; ++i) naxis[i] = -axis[i];
  fill_side(bodies[ib], posB, naxis, row.jacB, row.unitB);
  if (bodies[ia].dyn) for (int i = 0; i < 6; ++i) denom0 += row.jacA[i] * row.unitA[i];
  if (bodies[ib].dyn) for (int i = 0; i < 6; ++i) denom1 += row.jacB[i] * row.unitB[i];
  const T d = denom0 + denom1 + cfm;
  row.jacDiagABInv = (d > Lim<T>::eps()) ? T(1.0) / d : T(0.0);
  if (!is_friction) distance = distance_in + P.linear_slop;
  if (bodies[ia].dyn) for (int i = 0; i < 6; ++i) rel_vel += bodies[ia].vel[i] * row.jacA[i];
  if (bodies[ib].dyn) for (int i = 0; i < 6; ++i) rel_vel += bodies[ib].vel[i] * row.jacB[i];
  row.friction = mu;
  if (!is_friction) {
    if (babs(rel_vel) < P.restitution_thr) restitution = 0;
    else restitution = comb_rest * -rel_vel;
    if (restitution <= T(0.0)) restitution = 0;
  }
  T positionalError = 0, velocityError = restitution - rel_vel;
  if (is_friction) {
    positionalError = -distance * erp / P.dt;
  } else if (distance > 0) {
    positionalError = 0;
    velocityError -= distance / P.dt;
  } else {
    positionalError = -distance * erp / P.dt;
  }
  const T penetrationImpulse = positionalError * row.jacDiagABInv;
  const T velocityImpulse = velocityError * row.jacDiagABInv;
  row.rhs = penetrationImpulse + velocityImpulse;
  if (!is_friction) { row.lower = 0; row.upper = T(1e10f); }
  else { row.lower = -row.friction; row.upper = row.friction; }
  row.cfm = cfm * row.jacDiagABInv;
  row.applied = 0;
}

template <typename T> BCP_HD void apply_delta(SBody<T>& b, const T unit[6], T impulse) {
  if (!b.dyn) return;
  for (int i = 0; i < 6; ++i) b.dvel[i] += unit[i] * impulse;
}
template <typename T> BCP_HD T jdotdv(const SBody<T>& b, const T jac[6]) {
  if (!b.dyn) return T(0);
  T s = 0;
  for (int i = 0; i < 6; ++i) s += jac[i] * b.dvel[i];
  return s;
}

// resolveSingleConstraintRowGeneric
template <typename T> BCP_HD T resolve_row(Row<T>& c, SBody<T>* bodies) {
  T deltaImpulse = c.rhs - c.applied * c.cfm;
  const T dA = jdotdv(bodies[c.a], c.jacA);
  const T dB = jdotdv(bodies[c.b], c.jacB);
  deltaImpulse -= dA * c.jacDiagABInv;
  deltaImpulse -= dB * c.jacDiagABInv;
  const T sum = c.applied + deltaImpulse;
  if (sum < c.lower) { deltaImpulse = c.lower - c.applied; c.applied = c.lower; }
  else if (sum > c.upper) { deltaImpulse = c.upper - c.applied; c.applied = c.upper; }
  else c.applied = sum;
  apply_delta(bodies[c.a], c.unitA, deltaImpulse);
  apply_delta(bodies[c.b], c.unitB, deltaImpulse);
  return deltaImpulse / c.jacDiagABInv;
}

// resolveConeFrictionConstraintRows (clip radius evaluated without atan2/sin/cos)
template <typename T> BCP_HD T resolve_cone(Row<T>& cA, Row<T>& cB, SBody<T>* bodies) {
  T dIB = cB.rhs - cB.applied * cB.cfm;
  dIB -= jdotdv(bodies[cB.a], cB.jacA) * cB.jacDiagABInv;
  dIB -= jdotdv(bodies[cB.b], cB.jacB) * cB.jacDiagABInv;
  const T sumB = cB.applied + dIB;
  T dIA = cA.rhs - cA.applied * cA.cfm;
  dIA -= jdotdv(bodies[cA.a], cA.jacA) * cA.jacDiagABInv;
  dIA -= jdotdv(bodies[cA.b], cA.jacB) * cA.jacDiagABInv;
  const T sumA = cA.applied + dIA;
  if (sumA * sumA + sumB * sumB >= cA.lower * cB.lower) {
    const T r = bsqrt(sumA * sumA + sumB * sumB);
    const T sn = (r > T(0.0)) ? sumA / r : T(0.0);
    const T cs = (r > T(0.0)) ? sumB / r : T(1.0);
    const T clipA = babs(cA.lower * sn);
    const T clipB = babs(cB.lower * cs);
    if (sumA < -clipA) { dIA = -clipA - cA.applied; cA.applied = -clipA; }
    else if (sumA > clipA) { dIA = clipA - cA.applied; cA.applied = clipA; }
    else cA.applied = sumA;
    if (sumB < -clipB) { dIB = -clipB - cB.applied; cB.applied = -clipB; }
    else if (sumB > clipB) { dIB = clipB - cB.applied; cB.applied = clipB; }
    else cB.applied = sumB;
  } else {
    cA.applied = sumA;
    cB.applied = sumB;
  }
  apply_delta(bodies[cA.a], cA.unitA, dIA);
  apply_delta(bodies[cA.b], cA.unitB, dIA);
  apply_delta(bodies[cB.a], cB.unitA, dIB);
  apply_delta(bodies[cB.b], cB.unitB, dIB);
  return dIA / cA.jacDiagABInv + dIB / cB.jacDiagABInv;
}

// btPlaneSpace1 + normalize
template <typename T> BCP_HD void plane_space(const T n[3], T p[3], T q[3]) {
  T a, k;
  if (babs(n[2]) > T(0.7071067811865475244008443621048490)) {
    a = n[1] * n[1] + n[2] * n[2];
    k = T(1.0) / bsqrt(a);
    p[0] = 0; p[1] = -n[2] * k; p[2] = n[1] * k;
    q[0] = a * k; q[1] = -n[0] * p[2]; q[2] = n[0] * p[1];
  } else {
    a = n[0] * n[0] + n[1] * n[1];
    k = T(1.0) / bsqrt(a);
    p[0] = -n[1] * k; p[1] = n[0] * k; p[2] = 0;
    q[0] = -n[2] * p[1]; q[1] = n[2] * p[0]; q[2] = a * k;
  }
  T l = T(1.0) / bsqrt(dot3(p, p));
  for (int i = 0; i < 3; ++i) p[i] *= l;
  l = T(1.0) / bsqrt(dot3(q, q));
  for (int i = 0; i < 3; ++i) q[i] *= l;
}

// stepPositionsMultiDof for the base: x += dt v; exponential-map quaternion update
template <typename T> BCP_HD void integrate_body(const Params<T>& P, Body<T>& o) {
  const T dt = P.dt;
  for (int i = 0; i < 3; ++i) o.x[i] += dt * o.v[i];
  T fAngle = bsqrt(dot3(o.w, o.w));
  if (fAngle * dt > T(0.5 * (BCP_PI * 0.5))) fAngle = T(0.5) * T(BCP_PI * 0.5) / dt;
  T f;
  if (fAngle < T(0.001)) f = T(0.5) * dt - (dt * dt * dt) * T(0.020833333333) * fAngle * fAngle;
  else f = sin_core(T(0.5) * fAngle * dt) / fAngle;
  T dq[4], nq[4];
  for (int i = 0; i < 3; ++i) dq[i] = o.w[i] * f;
  dq[3] = cos_core(fAngle * dt * T(0.5));
  nq[0] = ((o.q[3] * dq[0] + o.q[0] * dq[3]) - o.q[1] * dq[2]) + o.q[2] * dq[1];
  nq[1] = ((o.q[3] * dq[1] + o.q[1] * dq[3]) - o.q[2] * dq[0]) + o.q[0] * dq[2];
  nq[2] = ((o.q[3] * dq[2] + o.q[2] * dq[3]) - o.q[0] * dq[1]) + o.q[1] * dq[0];
  nq[3] = ((o.q[3] * dq[3] - o.q[0] * dq[0]) - o.q[1] * dq[1]) - o.q[2] * dq[2];
  const T l = T(1.0) / bsqrt(nq[0] * nq[0] + nq[1] * nq[1] + nq[2] * nq[2] + nq[3] * nq[3]);
  for (int i = 0; i < 4; ++i) o.q[i] = nq[i] * l;
}

}  // namespace bcp
