// bcp_lambda.cuh — the pooled solver's iteration in IMPULSE space for the contact regimes whose
// rows fit a lane's registers.
//
// solve_iteration (bcp_solve.cuh) runs Bullet's projected Gauss-Seidel the way Bullet does: every
// row reads the bodies' velocity deltas (J.dv), clamps its impulse and writes the deltas back, so a
// row's ~25 instructions form one dependent chain through the bodies, and a solve that needs the
// full 50 iterations costs 50 x rows x that chain whatever else the SM could be doing (measured:
// 75 % of k_solve's lane-iterations are solves that run all 50 - the resting cart+pole stack of a
// resetting env - and the launch lasts as long as one warp needs for them).
// The same iteration on  A = J M^-1 J^T  (the rows' Delassus matrix):
//     res_i = rhsV_i - sum_j A_ij lambda_j            (what the row solve calls  rhsV - J.dv)
//     row i :  u = res_i - cfm lambda_i,  clamp exactly as before,  res_j -= A_ji d(lambda_i)  for all j
// is algebraically the identical sequence of iterates (same row order, same clamps, same cone
// clip, same residual for the early exit), but the only dependent chain from row to row is
//     res_i (one FMA with the previous row's delta) -> u -> delta -> clamp -> d(lambda)
// and the n updates res_j -= A_ji d are independent of each other.  A is built once per solve
// (n (n+1) / 2 dot products of 6..9 terms), the bodies' deltas once at the end.
//
// Two regimes, chosen per env from its contact counts (k_prep files everything else under the
// generic solver):
//   STACK  ground-pole has no points; cart-pole and ground-cart are frictionless (mu = 0):
//          rows = cart-pole normals (two bodies) then ground-cart normals: <= 8, A 8 x 8.
//   SPLIT  cart-pole has no points: the pole's rows (ground-pole normals, then their friction
//          pairs: <= 12, A 12 x 12) and the cart's rows (ground-cart normals: <= 4, A 4 x 4) do
//          not share a body, so the two blocks never couple.
// Slots an env does not use are zero rows (A = 0, invD = 0, rhsV = 0): they solve to a zero impulse
// with a zero residual, as in the generic sweep.  Floating point: the iterates differ from the
// generic sweep's by rounding only (fp64 instantiation vs the oracle <= 1e-10 per substep, same
// iteration counts: tests/test_host_kernels.py).
#pragma once
#include "bcp_solve.cuh"

namespace bcp {

enum { LAM_NONE = 0, LAM_STACK = 1, LAM_SPLIT = 2 };

// can this configuration use the impulse-space solver at all?  (frictionless cart pairs; the pole's
// ground contact frictionless or with the two-direction cone)
template <typename T> BCP_HD bool lam_config_ok(const Params<T>& P) {
  return P.rows_per_contact[BCP_PAIR_GROUND_CART] == 1 && P.rows_per_contact[BCP_PAIR_CART_POLE] == 1 &&
         (P.rows_per_contact[BCP_PAIR_GROUND_POLE] == 1 || P.rows_per_contact[BCP_PAIR_GROUND_POLE] == 3) && P.manifold_order == 0;
}
BCP_HD int lam_mode_of(int n_gc, int n_cp, int n_gp) {
  (void)n_gc;
  if (n_cp == 0) return LAM_SPLIT;
  if (n_gp == 0) return LAM_STACK;
  return LAM_NONE;
}

BCP_HD constexpr int tri(int i, int j) { return i >= j ? i * (i + 1) / 2 + j : j * (j + 1) / 2 + i; }

// per-lane system.  STACK: lam 0..3 cart-pole, 4..7 ground-cart; A[0..36) packed lower triangle.
// SPLIT: lam 0..3 ground-pole normals, 4 + 2c / 5 + 2c the friction pair of contact c, 12..15
// ground-cart; A[0..78) pole block, A[78..88) cart block.
template <typename T> struct LamSys {
  T A[88];
  T lam[16], res[16], invd[16];
};

// continuation buffer: word w of slot s at cont[w * n + s] (lanes of a warp hold consecutive slots)
enum { LAM_WORDS = 88 + 48 };
template <typename T> BCP_HD void lam_store(T* cont, int n, int slot, const LamSys<T>& S) {
#pragma unroll
  for (int w = 0; w < 88; ++w) cont[(size_t)w * n + slot] = S.A[w];
#pragma unroll
  for (int w = 0; w < 16; ++w) {
    cont[(size_t)(88 + w) * n + slot] = S.lam[w];
    cont[(size_t)(104 + w) * n + slot] = S.res[w];
    cont[(size_t)(120 + w) * n + slot] = S.invd[w];
  }
}
template <typename T> BCP_HD void lam_fetch(const T* cont, int n, int slot, LamSys<T>& S) {
#pragma unroll
  for (int w = 0; w < 88; ++w) S.A[w] = cont[(size_t)w * n + slot];
#pragma unroll
  for (int w = 0; w < 16; ++w) {
    S.lam[w] = cont[(size_t)(88 + w) * n + slot];
    S.res[w] = cont[(size_t)(104 + w) * n + slot];
    S.invd[w] = cont[(size_t)(120 + w) * n + slot];
  }
}

template <bool SMEM, typename T> BCP_HD void lam_row1(const T* p, T j[3], T d[3], T& rhs, T& invd) {
  SRow1<T> r;
  load_row<SMEM>(p, r);
  j[0] = r.a[0]; j[1] = r.a[1]; j[2] = r.a[2];
  d[0] = r.a[3]; d[1] = r.b[0]; d[2] = r.b[1];
  rhs = r.b[2]; invd = r.b[3];
}

// ---- STACK ----
template <bool SMEM, typename T>
BCP_HD void lam_build_stack(const Params<T>& P, const T* colp, const SolveLayout& L, const T* apn, LamSys<T>& S) {
  const T* const cp = colp + L.base[BCP_PAIR_CART_POLE];
  const T* const gc = colp + L.base[BCP_PAIR_GROUND_CART];
  const T imc = P.cart.inv_mass, imp = P.pole.inv_mass;
  T ja[4][3], jb[4][3], n[4][3], j[4][3], d[4][3], rhs[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    SRow2<T> r;
    load_row<SMEM>(cp + i * SW2, r);
    ja[i][0] = r.a[0]; ja[i][1] = r.a[1]; ja[i][2] = r.a[2];
    jb[i][0] = r.a[3]; jb[i][1] = r.b[0]; jb[i][2] = r.b[1];
    n[i][0] = r.b[2]; n[i][1] = r.b[3]; n[i][2] = r.c[0];
    rhs[i] = r.c[1]; S.invd[i] = r.c[2];
    lam_row1<SMEM>(gc + i * SW1, j[i], d[i], rhs[4 + i], S.invd[4 + i]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int k = 0; k <= i; ++k) {
      S.A[tri(i, k)] = (dot3(ja[i], ja[k]) + dot3(jb[i], jb[k])) + (imc + imp) * dot3(n[i], n[k]);       // cart-pole x cart-pole
      S.A[tri(4 + i, 4 + k)] = dot3(j[i], j[k]) + imc * dot3(d[i], d[k]);                                 // ground-cart x ground-cart
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) S.A[tri(4 + i, k)] = dot3(j[i], ja[k]) + imc * dot3(d[i], n[k]);           // through the cart
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) S.lam[i] = apn[(i < 4 ? (int)SLOT_CP : (int)SLOT_GC) + (i & 3)];  // warm start
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    T acc = rhs[i];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc -= S.A[tri(i, k)] * S.lam[k];
    S.res[i] = acc;
  }
}
template <typename T> BCP_HD T lam_normal(LamSys<T>& S, int i, T cfm, T& delta) {
  const T D = fast_rcp(S.invd[i]);
  const T u = S.res[i] - S.lam[i] * cfm;
  delta = u * S.invd[i];
  return clamp_normal(S.lam[i], delta, u, D);
}
// one iteration; n_cp / n_gc: warp-uniform trip counts as in solve_iteration (rows in pairs)
template <typename T> BCP_HD T lam_round_stack(const Params<T>& P, int mx_cp, int mx_gc, LamSys<T>& S) {
  T lsr = 0;
  const T cfm = P.cfm_normal;
  if (mx_cp > 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (i == 2 && mx_cp <= 2) break;
      T dl;
      const T r = lam_normal(S, i, cfm, dl);
      const T r2 = r * r;
      lsr = lsr > r2 ? lsr : r2;
#pragma unroll
      for (int k = 0; k < 8; ++k) S.res[k] -= S.A[tri(k, i)] * dl;
    }
  }
  if (mx_gc > 0) {
#pragma unroll
    for (int i = 4; i < 8; ++i) {
      if (i == 6 && mx_gc <= 2) break;
      T dl;
      const T r = lam_normal(S, i, cfm, dl);
      const T r2 = r * r;
      lsr = lsr > r2 ? lsr : r2;
#pragma unroll
      for (int k = 0; k < 8; ++k) S.res[k] -= S.A[tri(k, i)] * dl;
    }
  }
  return lsr;
}
template <bool SMEM, typename T>
BCP_HD void lam_finish_stack(const Params<T>& P, const T* colp, const SolveLayout& L, const LamSys<T>& S, SolveRegs<T>& R) {
  const T* const cp = colp + L.base[BCP_PAIR_CART_POLE];
  const T* const gc = colp + L.base[BCP_PAIR_GROUND_CART];
  for (int i = 0; i < 3; ++i) { R.cart.wt[i] = 0; R.cart.dv[i] = 0; R.pole.wt[i] = 0; R.pole.dv[i] = 0; }
  for (int s = 0; s < N_SLOTS; ++s) R.apn[s] = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    SRow2<T> r;
    load_row<SMEM>(cp + i * SW2, r);
    const T l = S.lam[i], la = P.cart.inv_mass * l, lb = P.pole.inv_mass * l;
    R.cart.wt[0] += r.a[0] * l; R.cart.wt[1] += r.a[1] * l; R.cart.wt[2] += r.a[2] * l;
    R.pole.wt[0] += r.a[3] * l; R.pole.wt[1] += r.b[0] * l; R.pole.wt[2] += r.b[1] * l;
    R.cart.dv[0] += r.b[2] * la; R.cart.dv[1] += r.b[3] * la; R.cart.dv[2] += r.c[0] * la;
    R.pole.dv[0] -= r.b[2] * lb; R.pole.dv[1] -= r.b[3] * lb; R.pole.dv[2] -= r.c[0] * lb;
    R.apn[SLOT_CP + i] = l;
    T j[3], d[3], rhs, invd;
    lam_row1<SMEM>(gc + i * SW1, j, d, rhs, invd);
    const T g = S.lam[4 + i], lm = P.cart.inv_mass * g;
    R.cart.wt[0] += j[0] * g; R.cart.wt[1] += j[1] * g; R.cart.wt[2] += j[2] * g;
    R.cart.dv[0] += d[0] * lm; R.cart.dv[1] += d[1] * lm; R.cart.dv[2] += d[2] * lm;
    R.apn[SLOT_GC + i] = g;
  }
}

// ---- SPLIT ----
template <bool SMEM, bool FRICTION, typename T>
BCP_HD void lam_build_split(const Params<T>& P, const T* colp, const SolveLayout& L, const T* apn, LamSys<T>& S) {
  const T* const gp = colp + L.base[BCP_PAIR_GROUND_POLE];
  const T* const gc = colp + L.base[BCP_PAIR_GROUND_CART];
  const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
  const T* const fgp = colp + (fr ? L.fbase[BCP_PAIR_GROUND_POLE] : 0);
  const T imc = P.cart.inv_mass, imp = P.pole.inv_mass;
  T j[12][3], d[12][3], rhs[16];
#pragma unroll
  for (int i = 0; i < 4; ++i) lam_row1<SMEM>(gp + i * SW1, j[i], d[i], rhs[i], S.invd[i]);
#pragma unroll
  for (int i = 4; i < 12; ++i) {
    if (fr) lam_row1<SMEM>(fgp + (i - 4) * SW1, j[i], d[i], rhs[i], S.invd[i]);
    else { j[i][0] = j[i][1] = j[i][2] = 0; d[i][0] = d[i][1] = d[i][2] = 0; rhs[i] = 0; S.invd[i] = 0; }
  }
#pragma unroll
  for (int i = 0; i < 12; ++i)
#pragma unroll
    for (int k = 0; k <= i; ++k) S.A[tri(i, k)] = dot3(j[i], j[k]) + imp * dot3(d[i], d[k]);
  {
    T jc[4][3], dc[4][3];
#pragma unroll
    for (int i = 0; i < 4; ++i) lam_row1<SMEM>(gc + i * SW1, jc[i], dc[i], rhs[12 + i], S.invd[12 + i]);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int k = 0; k <= i; ++k) S.A[78 + tri(i, k)] = dot3(jc[i], jc[k]) + imc * dot3(dc[i], dc[k]);
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) S.lam[i] = i < 4 ? apn[SLOT_GP + i] : (i >= 12 ? apn[SLOT_GC + (i - 12)] : T(0.0));  // friction starts at 0
#pragma unroll
  for (int i = 0; i < 12; ++i) {
    T acc = rhs[i];
#pragma unroll
    for (int k = 0; k < 4; ++k) acc -= S.A[tri(i, k)] * S.lam[k];
    S.res[i] = acc;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    T acc = rhs[12 + i];
#pragma unroll
    for (int k = 0; k < 4; ++k) acc -= S.A[78 + tri(i, k)] * S.lam[12 + k];
    S.res[12 + i] = acc;
  }
}
template <bool FRICTION, typename T> BCP_HD T lam_round_split(const Params<T>& P, bool fr, int mx_gp, int mx_gc, LamSys<T>& S) {
  T lsr = 0;
  const T cfm = P.cfm_normal, cfm_f = P.cfm_friction;
#define BCP_LACC(r_) { const T r__ = (r_); const T r2__ = r__ * r__; lsr = lsr > r2__ ? lsr : r2__; }
  // normal rows: ground-pole, then ground-cart (solve_iteration's order with cart-pole empty)
  if (mx_gp > 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (i == 2 && mx_gp <= 2) break;
      T dl;
      BCP_LACC(lam_normal(S, i, cfm, dl));
#pragma unroll
      for (int k = 0; k < 12; ++k) S.res[k] -= S.A[tri(k, i)] * dl;
    }
  }
  if (mx_gc > 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (i == 2 && mx_gc <= 2) break;
      T dl;
      BCP_LACC(lam_normal(S, 12 + i, cfm, dl));
#pragma unroll
      for (int k = 0; k < 4; ++k) S.res[12 + k] -= S.A[78 + tri(k, i)] * dl;
    }
  }
  // friction pairs of the pole's ground contacts (both rows of a pair from the same state, cone clip)
  if (FRICTION && fr && mx_gp > 0) {
    const T mu = P.mu[BCP_PAIR_GROUND_POLE];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (c == 2 && mx_gp <= 2) break;
      const int a = 4 + 2 * c, b = 5 + 2 * c;
      const T DA = fast_rcp(S.invd[a]), DB = fast_rcp(S.invd[b]);
      const T uB = S.res[b] - S.lam[b] * cfm_f, uA = S.res[a] - S.lam[a] * cfm_f;
      T dB = uB * S.invd[b], dA = uA * S.invd[a];
      BCP_LACC(cone_clip(S.lam[a], S.lam[b], dA, dB, uA, uB, DA, DB, mu * S.lam[c]));
#pragma unroll
      for (int k = 0; k < 12; ++k) { S.res[k] -= S.A[tri(k, a)] * dA; S.res[k] -= S.A[tri(k, b)] * dB; }
    }
  }
#undef BCP_LACC
  return lsr;
}
template <bool SMEM, bool FRICTION, typename T>
BCP_HD void lam_finish_split(const Params<T>& P, const T* colp, const SolveLayout& L, const LamSys<T>& S, SolveRegs<T>& R) {
  const T* const gp = colp + L.base[BCP_PAIR_GROUND_POLE];
  const T* const gc = colp + L.base[BCP_PAIR_GROUND_CART];
  const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
  const T* const fgp = colp + (fr ? L.fbase[BCP_PAIR_GROUND_POLE] : 0);
  for (int i = 0; i < 3; ++i) { R.cart.wt[i] = 0; R.cart.dv[i] = 0; R.pole.wt[i] = 0; R.pole.dv[i] = 0; }
  for (int s = 0; s < N_SLOTS; ++s) R.apn[s] = 0;
#pragma unroll
  for (int i = 0; i < 12; ++i) {
    if (i >= 4 && !fr) break;
    T j[3], d[3], rhs, invd;
    lam_row1<SMEM>(i < 4 ? gp + i * SW1 : fgp + (i - 4) * SW1, j, d, rhs, invd);
    const T l = S.lam[i], lm = P.pole.inv_mass * l;
    R.pole.wt[0] += j[0] * l; R.pole.wt[1] += j[1] * l; R.pole.wt[2] += j[2] * l;
    R.pole.dv[0] += d[0] * lm; R.pole.dv[1] += d[1] * lm; R.pole.dv[2] += d[2] * lm;
    if (i < 4) R.apn[SLOT_GP + i] = l;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    T j[3], d[3], rhs, invd;
    lam_row1<SMEM>(gc + i * SW1, j, d, rhs, invd);
    const T l = S.lam[12 + i], lm = P.cart.inv_mass * l;
    R.cart.wt[0] += j[0] * l; R.cart.wt[1] += j[1] * l; R.cart.wt[2] += j[2] * l;
    R.cart.dv[0] += d[0] * lm; R.cart.dv[1] += d[1] * lm; R.cart.dv[2] += d[2] * lm;
    R.apn[SLOT_GC + i] = l;
  }
}

// ---- GEN: the regime neither STACK nor SPLIT covers - the pole touches the cart AND the ground ----
// Up to 20 rows that all couple (cart: cart-pole + ground-cart rows; pole: cart-pole + ground-pole rows and their friction
// pairs): the 162 entries of the Delassus matrix that are not structurally zero (ground-cart and ground-pole rows share no
// body) do not leave room for anything else in a lane's registers, which is fine for what this regime is - a few dozen envs
// per launch (a pole leaning between cart and ground) that never reach the residual threshold and whose 50 row-space
// iterations on the side stream were the longest chain of the whole solve phase.  In impulse space the chain from row to
// row is the same 6 instructions as in STACK / SPLIT.
//   pole block  (rows 0..11: ground-pole normals, then friction pair 4 + 2c / 5 + 2c)   Ap  12 x 12 packed
//   cart-pole   (rows 12..15)  Ac 4 x 4 packed        ground-cart (rows 16..19)  Ag 4 x 4 packed
//   Acg[i * 4 + k]  cart-pole i x ground-cart k (through the cart)      Acp[i * 12 + k]  cart-pole i x pole-block k (the pole)
template <typename T> struct LamGen {
  T Ap[78], Ac[10], Ag[10], Acg[16], Acp[48];
  T lam[20], res[20], invd[20];
};
template <bool SMEM, bool FRICTION, typename T>
BCP_HD void lam_build_gen(const Params<T>& P, const T* colp, const SolveLayout& L, const T* apn, LamGen<T>& S) {
  const T* const gp = colp + L.base[BCP_PAIR_GROUND_POLE];
  const T* const cp = colp + L.base[BCP_PAIR_CART_POLE];
  const T* const gc = colp + L.base[BCP_PAIR_GROUND_CART];
  const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
  const T* const fgp = colp + (fr ? L.fbase[BCP_PAIR_GROUND_POLE] : 0);
  const T imc = P.cart.inv_mass, imp = P.pole.inv_mass;
  T rhs[20];
  {  // pole block, and its coupling to the cart-pole rows through the pole
    T j[12][3], d[12][3];
#pragma unroll
    for (int i = 0; i < 4; ++i) lam_row1<SMEM>(gp + i * SW1, j[i], d[i], rhs[i], S.invd[i]);
#pragma unroll
    for (int i = 4; i < 12; ++i) {
      if (fr) lam_row1<SMEM>(fgp + (i - 4) * SW1, j[i], d[i], rhs[i], S.invd[i]);
      else { j[i][0] = j[i][1] = j[i][2] = 0; d[i][0] = d[i][1] = d[i][2] = 0; rhs[i] = 0; S.invd[i] = 0; }
    }
#pragma unroll
    for (int i = 0; i < 12; ++i)
#pragma unroll
      for (int k = 0; k <= i; ++k) S.Ap[tri(i, k)] = dot3(j[i], j[k]) + imp * dot3(d[i], d[k]);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      SRow2<T> r;
      load_row<SMEM>(cp + i * SW2, r);
      const T jb[3] = {r.a[3], r.b[0], r.b[1]}, n[3] = {r.b[2], r.b[3], r.c[0]};
#pragma unroll
      for (int k = 0; k < 12; ++k) S.Acp[i * 12 + k] = dot3(jb, j[k]) - imp * dot3(n, d[k]);  // the normal pushes the pole with -n
    }
  }
  {  // the cart's side
    T ja[4][3], jb[4][3], n[4][3], jg[4][3], dg[4][3];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      SRow2<T> r;
      load_row<SMEM>(cp + i * SW2, r);
      ja[i][0] = r.a[0]; ja[i][1] = r.a[1]; ja[i][2] = r.a[2];
      jb[i][0] = r.a[3]; jb[i][1] = r.b[0]; jb[i][2] = r.b[1];
      n[i][0] = r.b[2]; n[i][1] = r.b[3]; n[i][2] = r.c[0];
      rhs[12 + i] = r.c[1]; S.invd[12 + i] = r.c[2];
      lam_row1<SMEM>(gc + i * SW1, jg[i], dg[i], rhs[16 + i], S.invd[16 + i]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
      for (int k = 0; k <= i; ++k) {
        S.Ac[tri(i, k)] = (dot3(ja[i], ja[k]) + dot3(jb[i], jb[k])) + (imc + imp) * dot3(n[i], n[k]);
        S.Ag[tri(i, k)] = dot3(jg[i], jg[k]) + imc * dot3(dg[i], dg[k]);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) S.Acg[i * 4 + k] = dot3(ja[i], jg[k]) + imc * dot3(n[i], dg[k]);
    }
  }
#pragma unroll
  for (int i = 0; i < 20; ++i)  // warm start; friction starts at 0
    S.lam[i] = i < 4 ? apn[SLOT_GP + i] : (i < 12 ? T(0.0) : (i < 16 ? apn[SLOT_CP + (i - 12)] : apn[SLOT_GC + (i - 16)]));
#pragma unroll
  for (int i = 0; i < 12; ++i) {
    T acc = rhs[i];
#pragma unroll
    for (int k = 0; k < 4; ++k) acc -= S.Ap[tri(i, k)] * S.lam[k];
#pragma unroll
    for (int m = 0; m < 4; ++m) acc -= S.Acp[m * 12 + i] * S.lam[12 + m];
    S.res[i] = acc;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    T acc = rhs[12 + i], ag = rhs[16 + i];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      acc -= S.Ac[tri(i, k)] * S.lam[12 + k];
      acc -= S.Acg[i * 4 + k] * S.lam[16 + k];
      acc -= S.Acp[i * 12 + k] * S.lam[k];
      ag -= S.Ag[tri(i, k)] * S.lam[16 + k];
      ag -= S.Acg[k * 4 + i] * S.lam[12 + k];
    }
    S.res[12 + i] = acc;
    S.res[16 + i] = ag;
  }
}
template <typename T> BCP_HD T gen_normal(LamGen<T>& S, int i, T cfm, T& delta) {
  const T D = fast_rcp(S.invd[i]);
  const T u = S.res[i] - S.lam[i] * cfm;
  delta = u * S.invd[i];
  return clamp_normal(S.lam[i], delta, u, D);
}
// one iteration, solve_iteration's order: ground-pole, cart-pole, ground-cart normals, then the ground-pole friction pairs;
// all four slots of every block (slots an env does not use are zero rows)
template <bool FRICTION, typename T> BCP_HD T lam_round_gen(const Params<T>& P, bool fr, LamGen<T>& S) {
  T lsr = 0;
  const T cfm = P.cfm_normal, cfm_f = P.cfm_friction;
#define BCP_GACC(r_) { const T r__ = (r_); const T r2__ = r__ * r__; lsr = lsr > r2__ ? lsr : r2__; }
#pragma unroll
  for (int i = 0; i < 4; ++i) {  // ground-pole normals
    T dl;
    BCP_GACC(gen_normal(S, i, cfm, dl));
#pragma unroll
    for (int k = 0; k < 12; ++k) S.res[k] -= S.Ap[tri(k, i)] * dl;
#pragma unroll
    for (int m = 0; m < 4; ++m) S.res[12 + m] -= S.Acp[m * 12 + i] * dl;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {  // cart-pole
    T dl;
    BCP_GACC(gen_normal(S, 12 + i, cfm, dl));
#pragma unroll
    for (int k = 0; k < 4; ++k) { S.res[12 + k] -= S.Ac[tri(k, i)] * dl; S.res[16 + k] -= S.Acg[i * 4 + k] * dl; }
#pragma unroll
    for (int k = 0; k < 12; ++k) S.res[k] -= S.Acp[i * 12 + k] * dl;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {  // ground-cart
    T dl;
    BCP_GACC(gen_normal(S, 16 + i, cfm, dl));
#pragma unroll
    for (int k = 0; k < 4; ++k) { S.res[16 + k] -= S.Ag[tri(k, i)] * dl; S.res[12 + k] -= S.Acg[k * 4 + i] * dl; }
  }
  if (FRICTION && fr) {
    const T mu = P.mu[BCP_PAIR_GROUND_POLE];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int a = 4 + 2 * c, b = 5 + 2 * c;
      const T DA = fast_rcp(S.invd[a]), DB = fast_rcp(S.invd[b]);
      const T uB = S.res[b] - S.lam[b] * cfm_f, uA = S.res[a] - S.lam[a] * cfm_f;
      T dB = uB * S.invd[b], dA = uA * S.invd[a];
      BCP_GACC(cone_clip(S.lam[a], S.lam[b], dA, dB, uA, uB, DA, DB, mu * S.lam[c]));
#pragma unroll
      for (int k = 0; k < 12; ++k) { S.res[k] -= S.Ap[tri(k, a)] * dA; S.res[k] -= S.Ap[tri(k, b)] * dB; }
#pragma unroll
      for (int m = 0; m < 4; ++m) { S.res[12 + m] -= S.Acp[m * 12 + a] * dA; S.res[12 + m] -= S.Acp[m * 12 + b] * dB; }
    }
  }
#undef BCP_GACC
  return lsr;
}
template <bool SMEM, bool FRICTION, typename T>
BCP_HD void lam_finish_gen(const Params<T>& P, const T* colp, const SolveLayout& L, const LamGen<T>& S, SolveRegs<T>& R) {
  const T* const gp = colp + L.base[BCP_PAIR_GROUND_POLE];
  const T* const cp = colp + L.base[BCP_PAIR_CART_POLE];
  const T* const gc = colp + L.base[BCP_PAIR_GROUND_CART];
  const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
  const T* const fgp = colp + (fr ? L.fbase[BCP_PAIR_GROUND_POLE] : 0);
  for (int i = 0; i < 3; ++i) { R.cart.wt[i] = 0; R.cart.dv[i] = 0; R.pole.wt[i] = 0; R.pole.dv[i] = 0; }
  for (int s = 0; s < N_SLOTS; ++s) R.apn[s] = 0;
#pragma unroll
  for (int i = 0; i < 12; ++i) {
    if (i >= 4 && !fr) break;
    T j[3], d[3], rhs, invd;
    lam_row1<SMEM>(i < 4 ? gp + i * SW1 : fgp + (i - 4) * SW1, j, d, rhs, invd);
    const T l = S.lam[i], lm = P.pole.inv_mass * l;
    R.pole.wt[0] += j[0] * l; R.pole.wt[1] += j[1] * l; R.pole.wt[2] += j[2] * l;
    R.pole.dv[0] += d[0] * lm; R.pole.dv[1] += d[1] * lm; R.pole.dv[2] += d[2] * lm;
    if (i < 4) R.apn[SLOT_GP + i] = l;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    SRow2<T> r;
    load_row<SMEM>(cp + i * SW2, r);
    const T l = S.lam[12 + i], la = P.cart.inv_mass * l, lb = P.pole.inv_mass * l;
    R.cart.wt[0] += r.a[0] * l; R.cart.wt[1] += r.a[1] * l; R.cart.wt[2] += r.a[2] * l;
    R.pole.wt[0] += r.a[3] * l; R.pole.wt[1] += r.b[0] * l; R.pole.wt[2] += r.b[1] * l;
    R.cart.dv[0] += r.b[2] * la; R.cart.dv[1] += r.b[3] * la; R.cart.dv[2] += r.c[0] * la;
    R.pole.dv[0] -= r.b[2] * lb; R.pole.dv[1] -= r.b[3] * lb; R.pole.dv[2] -= r.c[0] * lb;
    R.apn[SLOT_CP + i] = l;
    T j[3], d[3], rhs, invd;
    lam_row1<SMEM>(gc + i * SW1, j, d, rhs, invd);
    const T g = S.lam[16 + i], lm = P.cart.inv_mass * g;
    R.cart.wt[0] += j[0] * g; R.cart.wt[1] += j[1] * g; R.cart.wt[2] += j[2] * g;
    R.cart.dv[0] += d[0] * lm; R.cart.dv[1] += d[1] * lm; R.cart.dv[2] += d[2] * lm;
    R.apn[SLOT_GC + i] = g;
  }
}

}  // namespace bcp
