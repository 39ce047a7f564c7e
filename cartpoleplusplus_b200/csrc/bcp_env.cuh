// bcp_env.cuh — one environment: substep, force application, reset, step.
//
// The per-env program the kernels run.  Bookkeeping follows bullet_cartpole.py
// (cited per function); the substep is the Bullet pipeline of bcp_collide.cuh /
// bcp_solver.cuh in btDiscreteDynamicsWorld::internalSingleStepSimulation order.
#pragma once
#include "bcp_solver.cuh"

namespace bcp {

struct SubstepInfo {
  int iterations, rows, friction_rows, contacts[3];
};

template <typename T> BCP_HD void pose_box(const Body<T>& b, const BoxP<T>& c, Posed<T>& o) {
  for (int i = 0; i < 3; ++i) { o.x[i] = b.x[i]; o.h[i] = c.h[i]; }
  quat_to_mat(b.q, o.R);
}

// One p.stepSimulation() - bullet_cartpole.py:199,327,330.
// FRICTION=false drops every friction row (all pair frictions are 0, as with the
// fork's cart.urdf): those rows would have limits +-0 and contribute nothing.
template <typename T, bool FRICTION>
BCP_HD_NOINLINE void substep(const Params<T>& P, Env<T>& st, SubstepInfo* info) {
  Posed<T> box[3];  // 0 ground, 1 cart, 2 pole
  WorldPt<T> wp[3][4];
  SBody<T> bodies[3];
  Row<T> nrows[12];
  Row<T> frows[FRICTION ? 24 : 1];
  int n_n = 0, n_f = 0;
  const int pairA[3] = {0, 1, 0}, pairB[3] = {1, 2, 2};

  // ---- 1. collision detection on start-of-step poses ----
  for (int i = 0; i < 3; ++i) { box[0].x[i] = P.ground.x0[i]; box[0].h[i] = P.ground.h[i]; }
  for (int i = 0; i < 9; ++i) box[0].R[i] = (i % 4 == 0) ? T(1.0) : T(0.0);
  pose_box(st.cart, P.cart, box[1]);
  pose_box(st.pole, P.pole, box[2]);
  for (int m = 0; m < 3; ++m) {
    const Posed<T>& A = box[pairA[m]];
    const Posed<T>& B = box[pairB[m]];
    T ws[kBoxWs];
    const Contact<T>* cs = box_contacts(ws);
    census_phase<T>(CENSUS_NARROWPHASE);
    const int nc = box_box(A.x, A.R, A.h, B.x, B.R, B.h, ws);
    census_phase<T>(CENSUS_MANIFOLD);
    for (int k = 0; k < nc; ++k) manifold_add(st.mani[m], A, B, cs[k], P.breaking[m]);
    manifold_refresh(st.mani[m], A, B, P.breaking[m], wp[m]);
    if (info) info->contacts[m] = st.mani[m].count;
  }

  // ---- 2. gravity + external force -> unconstrained velocities ----
  census_phase<T>(CENSUS_DYNAMICS);
  bodies[0].dyn = 0;
  bodies[0].mass = bodies[0].inv_mass = 0;
  for (int i = 0; i < 3; ++i) { bodies[0].x[i] = box[0].x[i]; bodies[0].I[i] = 0; }
  for (int i = 0; i < 6; ++i) { bodies[0].vel[i] = 0; bodies[0].dvel[i] = 0; }
  for (int k = 0; k < 2; ++k) {
    SBody<T>& b = bodies[k + 1];
    const Body<T>& ob = k == 0 ? st.cart : st.pole;
    const BoxP<T>& bp = k == 0 ? P.cart : P.pole;
    b.dyn = 1;
    b.mass = bp.mass;
    b.inv_mass = bp.inv_mass;
    for (int i = 0; i < 3; ++i) { b.I[i] = bp.I[i]; b.x[i] = ob.x[i]; b.vel[i] = ob.w[i]; b.vel[3 + i] = ob.v[i]; }
    for (int i = 0; i < 9; ++i) b.R[i] = box[k + 1].R[i];
    for (int i = 0; i < 6; ++i) b.dvel[i] = 0;
  }
  {
    const T zero3[3] = {0, 0, 0};
    forward_dynamics(P, bodies[1], st.pend_f, st.pend_t);
    forward_dynamics(P, bodies[2], zero3, zero3);
  }

  // ---- 3. contact rows, manifolds in post-island-sort order ----
  census_phase<T>(CENSUS_ROWS);
  for (int k = 0; k < 3; ++k) {
    const int m = (P.manifold_order == 0) ? 2 - k : k;
    const Manifold<T>& mf = st.mani[m];
    const T mu = P.mu[m], rest = P.rest[m];
    for (int j = 0; j < mf.count; ++j) {
      const int fi = n_n;
      Row<T>& row = nrows[n_n];
      setup_row(P, row, bodies, pairA[m], pairB[m], wp[m][j].pa, wp[m][j].pb, mf.p[j].n, mf.p[j].dist, mu, rest, false);
      if (P.warmstart > T(0.0)) {
        row.applied = mf.p[j].imp * P.warmstart;
        if (row.applied != T(0.0)) {
          apply_delta(bodies[row.a], row.unitA, row.applied);
          apply_delta(bodies[row.b], row.unitB, row.applied);
        }
      }
      n_n++;
      if (FRICTION && mu != T(0.0)) {
        T d1[3], d2[3];
        plane_space(mf.p[j].n, d1, d2);
        setup_row(P, frows[n_f], bodies, pairA[m], pairB[m], wp[m][j].pa, wp[m][j].pb, d1, T(0), mu, rest, true);
        frows[n_f].fidx = fi;
        n_f++;
        if (P.friction_dirs == 2) {
          setup_row(P, frows[n_f], bodies, pairA[m], pairB[m], wp[m][j].pa, wp[m][j].pb, d2, T(0), mu, rest, true);
          frows[n_f].fidx = fi;
          n_f++;
        }
      }
    }
  }

  // ---- 4. projected Gauss-Seidel sweep with least-squares-residual early exit ----
  census_phase<T>(CENSUS_PGS);
  int it = 0;
  for (; it < P.solver_iterations; ++it) {
    T lsr = 0;
    for (int j = 0; j < n_n; ++j) {
      const T r = resolve_row(nrows[j], bodies);
      const T r2 = r * r;
      lsr = (lsr > r2) ? lsr : r2;
    }
    if (FRICTION) {
      if (P.friction_dirs == 2 && P.cone_friction) {
        for (int j = 0; j + 1 < n_f; j += 2) {
          const T tot = nrows[frows[j].fidx].applied;
          frows[j].lower = -(frows[j].friction * tot);
          frows[j].upper = frows[j].friction * tot;
          frows[j + 1].lower = -(frows[j + 1].friction * tot);
          frows[j + 1].upper = frows[j + 1].friction * tot;
          const T r = resolve_cone(frows[j], frows[j + 1], bodies);
          const T r2 = r * r;
          lsr = (lsr > r2) ? lsr : r2;
        }
      } else {
        for (int j = 0; j < n_f; ++j) {
          const T tot = nrows[frows[j].fidx].applied;
          if (tot > T(0.0)) {
            frows[j].lower = -(frows[j].friction * tot);
            frows[j].upper = frows[j].friction * tot;
            const T r = resolve_row(frows[j], bodies);
            const T r2 = r * r;
            lsr = (lsr > r2) ? lsr : r2;
          }
        }
      }
    }
    if (lsr <= P.residual_thr || it >= P.solver_iterations - 1) { it++; break; }
  }
  if (info) { info->iterations = it; info->rows = n_n; info->friction_rows = n_f; }

  // applied normal impulses -> manifold cache (solveGroupCacheFriendlyFinish)
  {
    int row = 0;
    for (int k = 0; k < 3; ++k) {
      const int m = (P.manifold_order == 0) ? 2 - k : k;
      for (int j = 0; j < st.mani[m].count; ++j) st.mani[m].p[j].imp = nrows[row++].applied;
    }
  }

  // ---- 5. velocities += deltas; integrate transforms ----
  census_phase<T>(CENSUS_INTEGRATE);
  for (int k = 0; k < 2; ++k) {
    SBody<T>& b = bodies[k + 1];
    Body<T>& o = k == 0 ? st.cart : st.pole;
    for (int i = 0; i < 6; ++i) b.vel[i] = clampv(b.vel[i] + b.dvel[i], -P.vmax, P.vmax);
    for (int i = 0; i < 3; ++i) { o.w[i] = b.vel[i]; o.v[i] = b.vel[3 + i]; }
    integrate_body(P, o);
  }
  // ---- 6. clearForcesAndTorques ----
  for (int i = 0; i < 3; ++i) { st.pend_f[i] = 0; st.pend_t[i] = 0; }
}

// p.applyExternalForce(cart, -1, (fx,fy,0), (0,0,0), frame) - bullet_cartpole.py:202-203,349-351
template <typename T> BCP_HD void apply_cart_force(const Params<T>& P, Env<T>& st, T fx, T fy) {
  T F[3], rel[3], Tq[3];
  const T f_in[3] = {fx, fy, T(0)};
  if (P.force_frame == BCP_FRAME_LINK) {
    T Rm[9];
    const T z[3] = {0, 0, 0};
    quat_to_mat(st.cart.q, Rm);
    mat_mul_vec(Rm, f_in, F);
    mat_mul_vec(Rm, z, rel);
  } else {
    for (int i = 0; i < 3; ++i) { F[i] = f_in[i]; rel[i] = T(0.0) - st.cart.x[i]; }
  }
  cross3(rel, F, Tq);
  for (int i = 0; i < 3; ++i) { st.pend_f[i] += F[i]; st.pend_t[i] += Tq[i]; }
}

// state_fields_of_pose_of(cart), (pole) -> state[r] - bullet_cartpole.py:43-45,307-311
// `o` is 8-byte aligned (14 floats per (repeat, env) block): seven float2 stores.
template <typename T> BCP_HD void write_obs(const Env<T>& st, float* o) {
  float v[14];
  for (int i = 0; i < 3; ++i) v[i] = (float)st.cart.x[i];
  for (int i = 0; i < 4; ++i) v[3 + i] = (float)st.cart.q[i];
  for (int i = 0; i < 3; ++i) v[7 + i] = (float)st.pole.x[i];
  for (int i = 0; i < 4; ++i) v[10 + i] = (float)st.pole.q[i];
#if defined(__CUDA_ARCH__)
  float2* o2 = reinterpret_cast<float2*>(o);
#pragma unroll
  for (int i = 0; i < 7; ++i) o2[i] = make_float2(v[2 * i], v[2 * i + 1]);
#else
  for (int i = 0; i < 14; ++i) o[i] = v[i];
#endif
}

// termination rule of bullet_cartpole.py:243-257, evaluated trig-free
template <typename T> BCP_HD int done_reason(const Params<T>& P, const Env<T>& st) {
  int reason = BCP_REASON_NONE;
  if (P.termination) {
    const T* q = st.pole.q;
    if (babs(st.pole.x[0]) > P.pos_thr || babs(st.pole.x[1]) > P.pos_thr) {
      reason = BCP_REASON_POSITION;
    } else {
      const T sarg = T(-2.0) * (q[0] * q[2] - q[3] * q[1]);
      const T sr = T(2.0) * (q[1] * q[2] + q[3] * q[0]);
      const T cr = q[3] * q[3] - q[0] * q[0] - q[1] * q[1] + q[2] * q[2];
      const bool pitch_out = babs(sarg) > P.sin_thr;
      const bool roll_out = (cr > T(0.0)) ? (babs(sr) > P.tan_thr * cr) : (cr < T(0.0) || sr != T(0.0));
      if (roll_out || pitch_out) reason = BCP_REASON_ORIENTATION;
    }
  }
  if (st.steps >= P.max_len) reason = BCP_REASON_LENGTH;
  return reason;
}

}  // namespace bcp
#include "bcp_episode.cuh"
