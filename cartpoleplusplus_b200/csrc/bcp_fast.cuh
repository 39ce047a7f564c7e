// bcp_fast.cuh — fast-mode substep: the same Bullet pipeline as substep<> (bcp_env.cuh),
// re-laid for the SM.
//
// What changes against the parity path (nothing in the algorithm, only its layout):
//   * constraint rows live in a per-env COLUMN of shared memory (`Col`), word-major so a
//     warp's 32 envs touch 32 consecutive banks; nothing of the solver is on the stack;
//   * rows are typed by collision pair - ground-cart and ground-pole rows touch one
//     dynamic body (10 words), cart-pole rows two (13 words) - so the sweep never indexes a
//     body dynamically and never multiplies by a static body's zeros;
//   * the angular half of every Jacobian is kept in the BODY frame, where the inertia is
//     diagonal: M^-1 J^T is an elementwise product with constants instead of R diag R^T;
//   * forward dynamics runs before collision detection (they are independent), so contact
//     rows are built straight from each refreshed manifold, with no world-point arrays.
// Row order (manifold order, normal rows then friction rows), limits, the cone clip and
// the least-squares-residual early exit are those of btMultiBodyConstraintSolver as
// restated in bcp_solver.cuh / oracle/bcp_oracle.c; results agree with the parity path
// to rounding (tests/test_host_kernels.py: fp64 instantiation vs the oracle <= 1e-10).
#pragma once
#include "bcp_env.cuh"

// Warp-uniform control for the solver loops.  Lanes of a warp own different envs whose
// contact counts and iteration counts differ; loops whose trip count is per-lane data make
// the warp split into groups that then run the whole solve one after another.  The fast
// path therefore gives every loop a warp-uniform trip count (fixed 4 contact slots per
// manifold, iteration loop left by a vote) and predicates the bodies.
// All 32 lanes of a warp run substep_fast together (`on` = false for a lane with nothing to
// do): the votes below carry the full mask and double as the warp's reconvergence points.
#if defined(__CUDA_ARCH__)
#define BCP_WARP_ANY(pred) (__any_sync(0xffffffffu, (pred)))
#define BCP_WARP_SYNC() __syncwarp()
#else
#define BCP_WARP_ANY(pred) (pred)
#define BCP_WARP_SYNC() ((void)0)
#endif

namespace bcp {

// word w of this env's scratch column lives at base[w * stride]
template <typename T> struct Col {
  T* base;
  int stride;
  BCP_HD T& operator[](int w) const { return base[(size_t)w * (size_t)stride]; }
};

// row words. one-body: J (local angular jacobian), D (signed linear direction);
// two-body: JA, JB, N (A side linear = +N, B side = -N).  Then the four scalars.
enum { RW1 = 10, RW2 = 13 };
enum { R1_J = 0, R1_D = 3, R1_RHS = 6, R1_INVD = 7, R1_DIAG = 8, R1_APPL = 9 };
enum { R2_JA = 0, R2_JB = 3, R2_N = 6, R2_RHS = 9, R2_INVD = 10, R2_DIAG = 11, R2_APPL = 12 };

// solver-side view of a dynamic body: local-frame angular, world-frame linear
template <typename T> struct FBody {
  T R[9], x[3];
  T wl[3], v[3];    // unconstrained velocities (after forward dynamics)
  T dwl[3], dv[3];  // deltas accumulated by the sweep
  T invI[3], invM;
  T sI[3];  // sqrt(invI): scale of the angular jacobians in the pooled solver's rows (bcp_solve.cuh)
};

template <typename T> BCP_HD void fbody_begin(const Params<T>& P, const Body<T>& ob, const BoxP<T>& bp, const T* Rm,
                                              const T F_ext[3], const T T_ext[3], FBody<T>& fb) {
  SBody<T> b;
  b.dyn = 1;
  b.mass = bp.mass;
  b.inv_mass = bp.inv_mass;
  for (int i = 0; i < 3; ++i) { b.I[i] = bp.I[i]; b.x[i] = ob.x[i]; b.vel[i] = ob.w[i]; b.vel[3 + i] = ob.v[i]; }
  for (int i = 0; i < 9; ++i) b.R[i] = Rm[i];
  forward_dynamics(P, b, F_ext, T_ext);
  for (int i = 0; i < 9; ++i) fb.R[i] = Rm[i];
  matT_mul_vec(Rm, b.vel, fb.wl);
  for (int i = 0; i < 3; ++i) {
    fb.x[i] = ob.x[i];
    fb.v[i] = b.vel[3 + i];
    fb.dwl[i] = 0;
    fb.dv[i] = 0;
    fb.invI[i] = T(1.0) / bp.I[i];
    fb.sI[i] = bsqrt(fb.invI[i]);
  }
  fb.invM = bp.inv_mass;
}

// local angular jacobian of direction `axis` applied at world point `pos`: R^T ((pos - x) x axis)
template <typename T> BCP_HD void local_jac(const FBody<T>& b, const T pos[3], const T axis[3], T sign, T j[3]) {
  T r[3], c[3];
  for (int i = 0; i < 3; ++i) r[i] = pos[i] - b.x[i];
  cross3(r, axis, c);
  matT_mul_vec(b.R, c, j);
  for (int i = 0; i < 3; ++i) j[i] *= sign;
}
template <typename T> BCP_HD T ang_denom(const FBody<T>& b, const T j[3]) {
  return j[0] * j[0] * b.invI[0] + j[1] * j[1] * b.invI[1] + j[2] * j[2] * b.invI[2];
}

// rhs of setupMultiBodyContactConstraint given the row's relative velocity
template <typename T>
BCP_HD T row_rhs(const Params<T>& P, T rel_vel, T distance_in, T comb_rest, T invD, bool is_friction) {
  T restitution = 0, distance = 0;
  if (!is_friction) {
    distance = distance_in + P.linear_slop;
    if (babs(rel_vel) < P.restitution_thr) restitution = 0;
    else restitution = comb_rest * -rel_vel;
    if (restitution <= T(0.0)) restitution = 0;
  }
  T positionalError = 0, velocityError = restitution - rel_vel;
  if (is_friction) positionalError = 0;
  else if (distance > 0) velocityError -= distance / P.dt;
  else positionalError = -distance * P.erp2 / P.dt;
  if (invD < T(0.0)) return positionalError + velocityError;  // velocity units (bcp_solve.cuh rows)
  return positionalError * invD + velocityError * invD;
}

// one-body row (A = static ground, B = body): J = [R^T (rB x -axis), -axis]
template <typename T>
BCP_HD void setup_row1(const Params<T>& P, const Col<T>& col, int off, FBody<T>& B, const T pb[3], const T axis[3],
                       T dist, T rest, bool is_friction, T warm_impulse) {
  T j[3], d[3];
  local_jac(B, pb, axis, T(-1.0), j);
  for (int i = 0; i < 3; ++i) d[i] = -axis[i];
  const T cfm = is_friction ? P.cfm_friction : P.cfm_normal;
  const T den = ang_denom(B, j) + B.invM * dot3(d, d) + cfm;
  const T invD = (den > Lim<T>::eps()) ? T(1.0) / den : T(0.0);
  const T rel_vel = dot3(j, B.wl) + dot3(d, B.v);
  for (int i = 0; i < 3; ++i) { col[off + R1_J + i] = j[i]; col[off + R1_D + i] = d[i]; }
  col[off + R1_RHS] = row_rhs(P, rel_vel, dist, rest, invD, is_friction);
  col[off + R1_INVD] = invD;
  col[off + R1_DIAG] = den;
  col[off + R1_APPL] = warm_impulse;
  if (warm_impulse != T(0.0)) {
    for (int i = 0; i < 3; ++i) { B.dwl[i] += j[i] * B.invI[i] * warm_impulse; B.dv[i] += d[i] * (B.invM * warm_impulse); }
  }
}
// two-body row (A = cart, B = pole): J = [RA^T (rA x axis), +axis | RB^T (rB x -axis), -axis]
template <typename T>
BCP_HD void setup_row2(const Params<T>& P, const Col<T>& col, int off, FBody<T>& A, FBody<T>& B, const T pa[3],
                       const T pb[3], const T axis[3], T dist, T rest, bool is_friction, T warm_impulse) {
  T ja[3], jb[3];
  local_jac(A, pa, axis, T(1.0), ja);
  local_jac(B, pb, axis, T(-1.0), jb);
  const T cfm = is_friction ? P.cfm_friction : P.cfm_normal;
  const T nn = dot3(axis, axis);
  const T den = (ang_denom(A, ja) + A.invM * nn) + (ang_denom(B, jb) + B.invM * nn) + cfm;
  const T invD = (den > Lim<T>::eps()) ? T(1.0) / den : T(0.0);
  const T rel_vel = (dot3(ja, A.wl) + dot3(axis, A.v)) + (dot3(jb, B.wl) - dot3(axis, B.v));
  for (int i = 0; i < 3; ++i) { col[off + R2_JA + i] = ja[i]; col[off + R2_JB + i] = jb[i]; col[off + R2_N + i] = axis[i]; }
  col[off + R2_RHS] = row_rhs(P, rel_vel, dist, rest, invD, is_friction);
  col[off + R2_INVD] = invD;
  col[off + R2_DIAG] = den;
  col[off + R2_APPL] = warm_impulse;
  if (warm_impulse != T(0.0)) {
    for (int i = 0; i < 3; ++i) {
      A.dwl[i] += ja[i] * A.invI[i] * warm_impulse; A.dv[i] += axis[i] * (A.invM * warm_impulse);
      B.dwl[i] += jb[i] * B.invI[i] * warm_impulse; B.dv[i] -= axis[i] * (B.invM * warm_impulse);
    }
  }
}

// ---- row updates (resolveSingleConstraintRowGeneric, per row type) ----
template <typename T> BCP_HD T clamp_impulse(T applied, T& delta, T lo, T hi) {
  const T sum = applied + delta;
  if (sum < lo) { delta = lo - applied; return lo; }
  if (sum > hi) { delta = hi - applied; return hi; }
  return sum;
}
template <typename T> struct Row1 {
  T j[3], d[3], rhs, invD, diag, ap;
};
template <typename T> struct Row2 {
  T ja[3], jb[3], n[3], rhs, invD, diag, ap;
};
template <typename T> BCP_HD void load_row(const Col<T>& col, int off, Row1<T>& r) {
  for (int i = 0; i < 3; ++i) { r.j[i] = col[off + R1_J + i]; r.d[i] = col[off + R1_D + i]; }
  r.rhs = col[off + R1_RHS]; r.invD = col[off + R1_INVD]; r.diag = col[off + R1_DIAG]; r.ap = col[off + R1_APPL];
}
template <typename T> BCP_HD void load_row(const Col<T>& col, int off, Row2<T>& r) {
  for (int i = 0; i < 3; ++i) { r.ja[i] = col[off + R2_JA + i]; r.jb[i] = col[off + R2_JB + i]; r.n[i] = col[off + R2_N + i]; }
  r.rhs = col[off + R2_RHS]; r.invD = col[off + R2_INVD]; r.diag = col[off + R2_DIAG]; r.ap = col[off + R2_APPL];
}
template <typename T> BCP_HD T jdv(const Row1<T>& r, const FBody<T>&, const FBody<T>& B) {
  return dot3(r.j, B.dwl) + dot3(r.d, B.dv);
}
template <typename T> BCP_HD T jdv(const Row2<T>& r, const FBody<T>& A, const FBody<T>& B) {
  T dl[3];
  for (int i = 0; i < 3; ++i) dl[i] = A.dv[i] - B.dv[i];
  return (dot3(r.ja, A.dwl) + dot3(r.jb, B.dwl)) + dot3(r.n, dl);
}
template <typename T> BCP_HD void apply(const Row1<T>& r, FBody<T>&, FBody<T>& B, T delta) {
  const T lm = B.invM * delta;
  for (int i = 0; i < 3; ++i) { B.dwl[i] += r.j[i] * (B.invI[i] * delta); B.dv[i] += r.d[i] * lm; }
}
template <typename T> BCP_HD void apply(const Row2<T>& r, FBody<T>& A, FBody<T>& B, T delta) {
  const T la = A.invM * delta, lb = B.invM * delta;
  for (int i = 0; i < 3; ++i) {
    A.dwl[i] += r.ja[i] * (A.invI[i] * delta);
    B.dwl[i] += r.jb[i] * (B.invI[i] * delta);
    A.dv[i] += r.n[i] * la;
    B.dv[i] -= r.n[i] * lb;
  }
}
template <typename ROW> struct RowTraits;
template <typename T> struct RowTraits<Row1<T>> { enum { W = RW1, APPL = R1_APPL }; };
template <typename T> struct RowTraits<Row2<T>> { enum { W = RW2, APPL = R2_APPL }; };

// one generic row: returns deltaImpulse / jacDiagABInv
template <typename T, typename ROW>
BCP_HD T update_row(const Col<T>& col, int off, FBody<T>& A, FBody<T>& B, T lo, T hi, T cfm) {
  ROW r;
  load_row(col, off, r);
  T delta = r.rhs - (r.ap * cfm + jdv(r, A, B)) * r.invD;
  col[off + RowTraits<ROW>::APPL] = clamp_impulse(r.ap, delta, lo, hi);
  apply(r, A, B, delta);
  return delta * r.diag;
}
// resolveConeFrictionConstraintRows on the two friction rows of one contact; L = mu * normal impulse
template <typename T, typename ROW>
BCP_HD T update_cone(const Col<T>& col, int offA, int offB, FBody<T>& A, FBody<T>& B, T L, T cfm) {
  ROW ra, rb;
  load_row(col, offA, ra);
  load_row(col, offB, rb);
  T dB = rb.rhs - (rb.ap * cfm + jdv(rb, A, B)) * rb.invD;
  T dA = ra.rhs - (ra.ap * cfm + jdv(ra, A, B)) * ra.invD;
  const T sumB = rb.ap + dB, sumA = ra.ap + dA;
  const T ss = sumA * sumA + sumB * sumB;
  T apA = sumA, apB = sumB;
  if (ss >= L * L) {
    const T r = bsqrt(ss);
    const T sn = (r > T(0.0)) ? sumA / r : T(0.0);
    const T cs = (r > T(0.0)) ? sumB / r : T(1.0);
    const T clipA = babs(L * sn), clipB = babs(L * cs);
    apA = clamp_impulse(ra.ap, dA, -clipA, clipA);
    apB = clamp_impulse(rb.ap, dB, -clipB, clipB);
  }
  col[offA + RowTraits<ROW>::APPL] = apA;
  col[offB + RowTraits<ROW>::APPL] = apB;
  apply(ra, A, B, dA);
  apply(rb, A, B, dB);
  return dA * ra.diag + dB * rb.diag;
}

// sweep of one manifold's rows of one kind (normal rows, or the friction rows): four contact
// slots, each skipped by the whole warp when no lane has that many contacts
template <typename T, typename ROW, bool FRICTION_KIND>
BCP_HD void sweep_manifold(const Params<T>& P, const Col<T>& col, int m, int count, FBody<T>& A, FBody<T>& B, T& lsr) {
  // `count` is 0 for a lane that has converged or is off: every lane takes part in every vote
  const int W = RowTraits<ROW>::W, rpc = P.rows_per_contact[m], base = P.row_base[m];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool on = j < count;
    if (!BCP_WARP_ANY(on)) break;
    if (on) {
      const int off = base + j * rpc * W;
      T r;
      if (!FRICTION_KIND) {
        r = update_row<T, ROW>(col, off, A, B, T(0.0), T(1e10f), P.cfm_normal);
      } else {
        const T tot = col[off + RowTraits<ROW>::APPL];
        const T L = P.mu[m] * tot;
        if (rpc == 3 && P.cone_friction) {
          r = update_cone<T, ROW>(col, off + W, off + 2 * W, A, B, L, P.cfm_friction);
        } else {
          r = 0;
          if (tot > T(0.0)) {
            for (int k = 1; k < rpc; ++k) {
              const T rk = update_row<T, ROW>(col, off + k * W, A, B, -L, L, P.cfm_friction);
              const T rk2 = rk * rk;
              lsr = (lsr > rk2) ? lsr : rk2;
            }
          }
        }
      }
      const T r2 = r * r;
      lsr = (lsr > r2) ? lsr : r2;
    }
  }
}

// rows of manifold m from its refreshed points (setupMultiBodyContactConstraint order:
// the normal row, then the friction rows along btPlaneSpace1(normal))
template <typename T, bool FRICTION, bool TWO>
BCP_HD void build_rows(const Params<T>& P, const Col<T>& col, int m, int base, int W, const Manifold<T>& mf,
                       const WorldPt<T>* wp, FBody<T>& A, FBody<T>& B) {
  const int rpc = P.rows_per_contact[m];
  const T rest = P.rest[m];
  for (int j = 0; j < mf.count; ++j) {
    const int off = base + j * rpc * W;
    const T warm = (P.warmstart > T(0.0)) ? mf.p[j].imp * P.warmstart : T(0.0);
    if (TWO) setup_row2(P, col, off, A, B, wp[j].pa, wp[j].pb, mf.p[j].n, mf.p[j].dist, rest, false, warm);
    else setup_row1(P, col, off, B, wp[j].pb, mf.p[j].n, mf.p[j].dist, rest, false, warm);
    if (FRICTION && rpc > 1) {
      T d1[3], d2[3];
      plane_space(mf.p[j].n, d1, d2);
      if (TWO) setup_row2(P, col, off + W, A, B, wp[j].pa, wp[j].pb, d1, T(0), rest, true, T(0));
      else setup_row1(P, col, off + W, B, wp[j].pb, d1, T(0), rest, true, T(0));
      if (rpc > 2) {
        if (TWO) setup_row2(P, col, off + 2 * W, A, B, wp[j].pa, wp[j].pb, d2, T(0), rest, true, T(0));
        else setup_row1(P, col, off + 2 * W, B, wp[j].pb, d2, T(0), rest, true, T(0));
      }
    }
  }
}

template <typename T> BCP_HD void fbody_finish(const Params<T>& P, const FBody<T>& fb, Body<T>& o) {
  T w[3], dw[3];
  mat_mul_vec(fb.R, fb.wl, w);
  mat_mul_vec(fb.R, fb.dwl, dw);
  for (int i = 0; i < 3; ++i) {
    o.w[i] = clampv(w[i] + dw[i], -P.vmax, P.vmax);
    o.v[i] = clampv(fb.v[i] + fb.dv[i], -P.vmax, P.vmax);
  }
  integrate_body(P, o);
}

// Row layout of a column: first word of each pair's rows and the words per row.  The
// shared-memory columns are compact (10 / 13 words per row); the global solve records of
// the split pipeline (bcp_split.cuh) pad rows to 12 / 16 words for 16-byte accesses.
struct RowLayout {
  int base[3], w[3];
};
template <typename T> BCP_HD RowLayout compact_layout(const Params<T>& P) {
  RowLayout L;
  for (int m = 0; m < 3; ++m) { L.base[m] = P.row_base[m]; L.w[m] = (m == BCP_PAIR_CART_POLE) ? (int)RW2 : (int)RW1; }
  return L;
}

// first half of a substep: unconstrained velocities, narrowphase, manifold caches, rows
template <typename T, bool FRICTION>
BCP_HD void fast_setup(const Params<T>& P, Env<T>& st, const Col<T>& col, const RowLayout& L, FBody<T>& cart,
                       FBody<T>& pole, bool on, SubstepInfo* info) {
  Posed<T> gbox, cbox, pbox;
  BCP_WARP_SYNC();
  if (on) {
    for (int i = 0; i < 3; ++i) { gbox.x[i] = P.ground.x0[i]; gbox.h[i] = P.ground.h[i]; }
    for (int i = 0; i < 9; ++i) gbox.R[i] = (i % 4 == 0) ? T(1.0) : T(0.0);
    pose_box(st.cart, P.cart, cbox);
    pose_box(st.pole, P.pole, pbox);
    // gravity + pending external force -> unconstrained velocities
    const T zero3[3] = {0, 0, 0};
    fbody_begin(P, st.cart, P.cart, cbox.R, st.pend_f, st.pend_t, cart);
    fbody_begin(P, st.pole, P.pole, pbox.R, zero3, zero3, pole);
  }
  // narrowphase + manifold cache + rows, pair by pair
  for (int m = 0; m < 3; ++m) {
    BCP_WARP_SYNC();
    if (on) {
      const Posed<T>& A = (m == BCP_PAIR_CART_POLE) ? cbox : gbox;
      const Posed<T>& B = (m == BCP_PAIR_GROUND_CART) ? cbox : pbox;
      T ws[kBoxWs];
      const Contact<T>* cs = box_contacts(ws);
      WorldPt<T> wp[4];
      const int nc = box_box(A.x, A.R, A.h, B.x, B.R, B.h, ws);
      for (int k = 0; k < nc; ++k) manifold_add(st.mani[m], A, B, cs[k], P.breaking[m]);
      manifold_refresh(st.mani[m], A, B, P.breaking[m], wp);
      if (m == BCP_PAIR_GROUND_CART) build_rows<T, FRICTION, false>(P, col, m, L.base[m], L.w[m], st.mani[m], wp, cart, cart);
      else if (m == BCP_PAIR_CART_POLE) build_rows<T, FRICTION, true>(P, col, m, L.base[m], L.w[m], st.mani[m], wp, cart, pole);
      else build_rows<T, FRICTION, false>(P, col, m, L.base[m], L.w[m], st.mani[m], wp, pole, pole);
      if (info) info->contacts[m] = st.mani[m].count;
    }
  }
  BCP_WARP_SYNC();
}

// one projected Gauss-Seidel iteration over a compact column: manifolds in island order,
// normal rows then friction rows.  Contact counts of 0 switch a lane off; every lane takes
// part in every vote.  Returns the iteration's least-squares residual.
template <typename T, bool FRICTION>
BCP_HD T fast_iteration(const Params<T>& P, const Col<T>& col, int n_gc, int n_cp, int n_gp, FBody<T>& cart,
                        FBody<T>& pole) {
  T lsr = 0;
  if (P.manifold_order == 0) {
    sweep_manifold<T, Row1<T>, false>(P, col, BCP_PAIR_GROUND_POLE, n_gp, pole, pole, lsr);
    sweep_manifold<T, Row2<T>, false>(P, col, BCP_PAIR_CART_POLE, n_cp, cart, pole, lsr);
    sweep_manifold<T, Row1<T>, false>(P, col, BCP_PAIR_GROUND_CART, n_gc, cart, cart, lsr);
  } else {
    sweep_manifold<T, Row1<T>, false>(P, col, BCP_PAIR_GROUND_CART, n_gc, cart, cart, lsr);
    sweep_manifold<T, Row2<T>, false>(P, col, BCP_PAIR_CART_POLE, n_cp, cart, pole, lsr);
    sweep_manifold<T, Row1<T>, false>(P, col, BCP_PAIR_GROUND_POLE, n_gp, pole, pole, lsr);
  }
  if (FRICTION) {
    if (P.manifold_order == 0) {
      if (P.rows_per_contact[BCP_PAIR_GROUND_POLE] > 1) sweep_manifold<T, Row1<T>, true>(P, col, BCP_PAIR_GROUND_POLE, n_gp, pole, pole, lsr);
      if (P.rows_per_contact[BCP_PAIR_CART_POLE] > 1) sweep_manifold<T, Row2<T>, true>(P, col, BCP_PAIR_CART_POLE, n_cp, cart, pole, lsr);
      if (P.rows_per_contact[BCP_PAIR_GROUND_CART] > 1) sweep_manifold<T, Row1<T>, true>(P, col, BCP_PAIR_GROUND_CART, n_gc, cart, cart, lsr);
    } else {
      if (P.rows_per_contact[BCP_PAIR_GROUND_CART] > 1) sweep_manifold<T, Row1<T>, true>(P, col, BCP_PAIR_GROUND_CART, n_gc, cart, cart, lsr);
      if (P.rows_per_contact[BCP_PAIR_CART_POLE] > 1) sweep_manifold<T, Row2<T>, true>(P, col, BCP_PAIR_CART_POLE, n_cp, cart, pole, lsr);
      if (P.rows_per_contact[BCP_PAIR_GROUND_POLE] > 1) sweep_manifold<T, Row1<T>, true>(P, col, BCP_PAIR_GROUND_POLE, n_gp, pole, pole, lsr);
    }
  }
  return lsr;
}

// One p.stepSimulation() - bullet_cartpole.py:199,327,330 - fast layout.  Called by all 32
// lanes of a warp; a lane with `on` == false only takes part in the votes.
template <typename T, bool FRICTION>
BCP_HD_NOINLINE void substep_fast(const Params<T>& P, Env<T>& st, const Col<T>& col, bool on, SubstepInfo* info) {
  FBody<T> cart, pole;
  const RowLayout L = compact_layout(P);
  fast_setup<T, FRICTION>(P, st, col, L, cart, pole, on, info);
  const int c_gc = on ? st.mani[BCP_PAIR_GROUND_CART].count : 0, c_cp = on ? st.mani[BCP_PAIR_CART_POLE].count : 0,
            c_gp = on ? st.mani[BCP_PAIR_GROUND_POLE].count : 0;

  // projected Gauss-Seidel with the least-squares-residual early exit.  Every loop has a
  // warp-uniform trip count: the iteration loop is left by a vote, a lane that has converged
  // sweeps with zero contacts (all its row updates predicated off).
  int it = 0;
  bool busy = on;
  for (int k = 0; k < P.solver_iterations; ++k) {
    if (!BCP_WARP_ANY(busy)) break;
    const T lsr = fast_iteration<T, FRICTION>(P, col, busy ? c_gc : 0, busy ? c_cp : 0, busy ? c_gp : 0, cart, pole);
    if (busy) {
      it = k + 1;
      if (lsr <= P.residual_thr || k >= P.solver_iterations - 1) busy = false;
    }
  }
  BCP_WARP_SYNC();
  if (on) {
    if (info) {
      info->iterations = it;
      info->rows = c_gc + c_cp + c_gp;
      int nf = 0;
      for (int m = 0; m < 3; ++m) nf += st.mani[m].count * (P.rows_per_contact[m] - 1);
      info->friction_rows = FRICTION ? nf : 0;
    }
    // applied normal impulses -> manifold cache
    for (int m = 0; m < 3; ++m) {
      const int W = (m == BCP_PAIR_CART_POLE) ? (int)RW2 : (int)RW1;
      const int ap = (m == BCP_PAIR_CART_POLE) ? (int)R2_APPL : (int)R1_APPL;
      for (int j = 0; j < st.mani[m].count; ++j) st.mani[m].p[j].imp = col[P.row_base[m] + j * P.rows_per_contact[m] * W + ap];
    }
    fbody_finish(P, cart, st.cart);
    fbody_finish(P, pole, st.pole);
    for (int i = 0; i < 3; ++i) { st.pend_f[i] = 0; st.pend_t[i] = 0; }
  }
}

// substep functor of the fast path (see StepParity in bcp_episode.cuh)
template <typename T, bool FRICTION> struct StepFast {
  Col<T> col;
  BCP_HD void operator()(const Params<T>& P, Env<T>& st, bool on) const { substep_fast<T, FRICTION>(P, st, col, on, nullptr); }
};

}  // namespace bcp
