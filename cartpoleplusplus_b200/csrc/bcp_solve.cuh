// bcp_solve.cuh — the pooled solver's row format and its branch-free iteration.
//
// Rows of the split pipeline (bcp_split.cuh) are stored for 16-byte accesses, in "velocity
// units" so that nothing but the applied impulse changes during a solve:
//   one-body row  (8 words) : j[3] d[3] rhsV invD            rhsV = rhs / invD
//   two-body row (12 words) : ja[3] jb[3] n[3] rhsV invD pad
// and   deltaImpulse = (rhsV - applied*cfm - J.dv) * invD,
//       residual     =  rhsV - applied*cfm - J.dv            (= deltaImpulse / jacDiagABInv)
// unless the impulse is clamped (then deltaImpulse_clamped / invD).  Applied impulses live
// in registers (every slot index below is a compile-time constant).  Contact slots a lane
// does not use are ZERO rows: they solve to a zero impulse with a zero residual, so the
// iteration has no per-lane branch at all - only warp-uniform trip counts (the largest
// contact count per manifold among the warp's envs).  Same row order, limits, cone clip and
// early exit as bcp_fast.cuh / the oracle.
#pragma once
#include "bcp_fast.cuh"

namespace bcp {

enum { SW1 = 8, SW2 = 12 };  // words per one-body / two-body row
enum { S1_J = 0, S1_D = 3, S1_RHSV = 6, S1_INVD = 7 };
enum { S2_JA = 0, S2_JB = 3, S2_N = 6, S2_RHSV = 9, S2_INVD = 10 };
enum { SLOT_GC = 0, SLOT_CP = 4, SLOT_GP = 8, N_SLOTS = 12 };  // applied-impulse slots: pair * 4 + contact

// word offsets of the three pairs' row blocks inside a column / a record's row area
struct SolveLayout {
  int base[3];   // first word of the pair's 4 normal rows (stride SW1 / SW2)
  int fbase[3];  // first word of the pair's 4 friction-row pairs (stride 2 rows), -1 if none
  int rpc[3], words;
};
template <typename T> BCP_HD SolveLayout solve_layout(const Params<T>& P) {
  SolveLayout L;
  int at = 0;
  for (int m = 0; m < 3; ++m) {  // normal rows of every pair first: fixed strides, and the
    L.base[m] = at;              // live part of each block is a prefix (contacts 0..count-1)
    L.rpc[m] = P.rows_per_contact[m];
    at += 4 * (m == BCP_PAIR_CART_POLE ? (int)SW2 : (int)SW1);
  }
  for (int m = 0; m < 3; ++m) {
    L.fbase[m] = L.rpc[m] > 1 ? at : -1;
    at += 4 * (L.rpc[m] - 1) * (m == BCP_PAIR_CART_POLE ? (int)SW2 : (int)SW1);
  }
  L.words = at;
  return L;
}

template <typename T> BCP_HD void ld4(const T* p, T o[4]) {
  o[0] = p[0]; o[1] = p[1]; o[2] = p[2]; o[3] = p[3];
}
#if defined(__CUDA_ARCH__)
template <> __device__ __forceinline__ void ld4<float>(const float* p, float o[4]) {
  const float4 v = *reinterpret_cast<const float4*>(p);
  o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
}
#endif
// row load of the solver sweep.  SMEM (the pooled kernel: `p` is a shared-memory column): a
// volatile ld.shared, which the compiler keeps where it is written - one row AHEAD of its use -
// instead of sinking it below the warp-uniform loop exits next to the first consumer.
template <bool SMEM, typename T> BCP_HD void ld4_row(const T* p, T o[4]) { ld4(p, o); }
#if defined(__CUDA_ARCH__)
template <> __device__ __forceinline__ void ld4_row<true, float>(const float* p, float o[4]) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(p);
  asm volatile("ld.volatile.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]) : "r"(a));
}
#endif
template <typename T> BCP_HD void st4(T* p, const T v[4]) {
  p[0] = v[0]; p[1] = v[1]; p[2] = v[2]; p[3] = v[3];
}
#if defined(__CUDA_ARCH__)
template <> __device__ __forceinline__ void st4<float>(float* p, const float v[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
#endif
// reciprocal / reciprocal square root for the residual and the cone scale: one MUFU each on
// the device (2 ulp; neither feeds a clamp decision), exact on the host and in double
BCP_HD float fast_rcp(float x) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return 1.0f / x;
#endif
}
BCP_HD double fast_rcp(double x) { return 1.0 / x; }
BCP_HD float fast_rsqrt(float x) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return 1.0f / sqrtf(x);
#endif
}
BCP_HD double fast_rsqrt(double x) { return 1.0 / sqrt(x); }
BCP_HD float bmin(float a, float b) { return fminf(a, b); }
BCP_HD float bmax(float a, float b) { return fmaxf(a, b); }
BCP_HD double bmin(double a, double b) { return fmin(a, b); }
BCP_HD double bmax(double a, double b) { return fmax(a, b); }

// solver-side body.  Angular deltas are kept SCALED, wt = dw_local / sqrt(invI) per axis, and the
// rows hold angular jacobians scaled the other way (j * sqrt(invI)): J.dw and the update
// dw += invI j delta become plain dot products / FMAs without the per-row invI multiplies.
template <typename T> struct SBodyLite {
  T wt[3], dv[3], invM;
};
// everything a lane carries through a solve
template <typename T> struct SolveRegs {
  SBodyLite<T> cart, pole;
  T apn[N_SLOTS];     // applied normal impulses
  T apf[N_SLOTS][2];  // applied friction impulses
};
template <typename T> struct SRow1 { T a[4], b[4]; };        // a = j0 j1 j2 d0 | b = d1 d2 rhsV invD
template <typename T> struct SRow2 { T a[4], b[4], c[4]; };  // a = ja0 ja1 ja2 jb0 | b = jb1 jb2 n0 n1 | c = n2 rhsV invD -
template <bool SMEM, typename T> BCP_HD void load_row(const T* p, SRow1<T>& r) { ld4_row<SMEM>(p, r.a); ld4_row<SMEM>(p + 4, r.b); }
template <bool SMEM, typename T> BCP_HD void load_row(const T* p, SRow2<T>& r) {
  ld4_row<SMEM>(p, r.a); ld4_row<SMEM>(p + 4, r.b); ld4_row<SMEM>(p + 8, r.c);
}

// normal row: accumulated impulse clamped at 0 from below (Bullet's upper limit of 1e10 is
// never reached by a finite state); returns the residual of the row
template <typename T> BCP_HD T clamp_normal(T& ap, T& delta, T u, T D) {  // D = 1 / invD
  const T sum = ap + delta;
  const T nv = bmax(sum, T(0.0));
  delta = nv - ap;
  ap = nv;
  return (nv != sum) ? delta * D : u;
}
template <typename T> BCP_HD T solve_row1(const SRow1<T>& r, SBodyLite<T>& B, T& ap, T cfm) {
  const T* a = r.a; const T* b = r.b;
  const T D = fast_rcp(b[3]);  // off the dependent chain: issued with the row's first products
  const T jdv = (a[0] * B.wt[0] + a[1] * B.wt[1] + a[2] * B.wt[2]) + (a[3] * B.dv[0] + b[0] * B.dv[1] + b[1] * B.dv[2]);
  const T u = b[2] - ap * cfm - jdv;
  T delta = u * b[3];
  const T res = clamp_normal(ap, delta, u, D);
  const T lm = B.invM * delta;
  B.wt[0] += a[0] * delta; B.wt[1] += a[1] * delta; B.wt[2] += a[2] * delta;
  B.dv[0] += a[3] * lm; B.dv[1] += b[0] * lm; B.dv[2] += b[1] * lm;
  return res;
}
template <typename T> BCP_HD T solve_row2(const SRow2<T>& r, SBodyLite<T>& A, SBodyLite<T>& B, T& ap, T cfm) {
  const T* a = r.a; const T* b = r.b; const T* c = r.c;
  const T D = fast_rcp(c[2]);
  const T jdv = (a[0] * A.wt[0] + a[1] * A.wt[1] + a[2] * A.wt[2]) + (a[3] * B.wt[0] + b[0] * B.wt[1] + b[1] * B.wt[2]) +
                (b[2] * (A.dv[0] - B.dv[0]) + b[3] * (A.dv[1] - B.dv[1]) + c[0] * (A.dv[2] - B.dv[2]));
  const T u = c[1] - ap * cfm - jdv;
  T delta = u * c[2];
  const T res = clamp_normal(ap, delta, u, D);
  const T la = A.invM * delta, lb = B.invM * delta;
  A.wt[0] += a[0] * delta; A.wt[1] += a[1] * delta; A.wt[2] += a[2] * delta;
  B.wt[0] += a[3] * delta; B.wt[1] += b[0] * delta; B.wt[2] += b[1] * delta;
  A.dv[0] += b[2] * la; A.dv[1] += b[3] * la; A.dv[2] += c[0] * la;
  B.dv[0] -= b[2] * lb; B.dv[1] -= b[3] * lb; B.dv[2] -= c[0] * lb;
  return res;
}

// cone clip of the two friction impulses of one contact (resolveConeFrictionConstraintRows)
template <typename T>
BCP_HD T cone_clip(T& apA, T& apB, T& dA, T& dB, T uA, T uB, T DA, T DB, T L) {  // DA, DB = 1 / invD
  // Bullet clips (sumA, sumB) to the circle of radius |L| through the angle (sin, cos) =
  // (sumA, sumB) / r; outside the circle both clamps bind, i.e. the pair is scaled by |L| / r
  const T sumA = apA + dA, sumB = apB + dB;
  const T ss = sumA * sumA + sumB * sumB;
  const bool clip = ss >= L * L && ss > T(0.0);
  const T rs = babs(L) * fast_rsqrt(ss);  // issued without waiting for the clip decision
  const T sc = clip ? rs : T(1.0);
  const T nA = sumA * sc, nB = sumB * sc;
  dA = nA - apA;
  dB = nB - apB;
  apA = nA;
  apB = nB;
  const T rA = clip ? dA * DA : uA;
  const T rB = clip ? dB * DB : uB;
  return rA + rB;
}
template <typename T> BCP_HD T solve_cone1(const SRow1<T>& ra, const SRow1<T>& rb, SBodyLite<T>& B, T& apA, T& apB, T L, T cfm) {
  const T* a = ra.a; const T* b = ra.b; const T* c = rb.a; const T* d = rb.b;
  const T DA = fast_rcp(b[3]), DB = fast_rcp(d[3]);
  const T jdvB = (c[0] * B.wt[0] + c[1] * B.wt[1] + c[2] * B.wt[2]) + (c[3] * B.dv[0] + d[0] * B.dv[1] + d[1] * B.dv[2]);
  const T jdvA = (a[0] * B.wt[0] + a[1] * B.wt[1] + a[2] * B.wt[2]) + (a[3] * B.dv[0] + b[0] * B.dv[1] + b[1] * B.dv[2]);
  const T uB = d[2] - apB * cfm - jdvB, uA = b[2] - apA * cfm - jdvA;
  T dB = uB * d[3], dA = uA * b[3];
  const T res = cone_clip(apA, apB, dA, dB, uA, uB, DA, DB, L);
  const T lA = B.invM * dA, lB = B.invM * dB;
  B.wt[0] += a[0] * dA; B.wt[1] += a[1] * dA; B.wt[2] += a[2] * dA;
  B.dv[0] += a[3] * lA; B.dv[1] += b[0] * lA; B.dv[2] += b[1] * lA;
  B.wt[0] += c[0] * dB; B.wt[1] += c[1] * dB; B.wt[2] += c[2] * dB;
  B.dv[0] += c[3] * lB; B.dv[1] += d[0] * lB; B.dv[2] += d[1] * lB;
  return res;
}
template <typename T>
BCP_HD T solve_cone2(const SRow2<T>& ra, const SRow2<T>& rb, SBodyLite<T>& A, SBodyLite<T>& B, T& apA, T& apB, T L, T cfm) {
  const T* a = ra.a; const T* b = ra.b; const T* c = ra.c; const T* e = rb.a; const T* f = rb.b; const T* g = rb.c;
  const T DA = fast_rcp(c[2]), DB = fast_rcp(g[2]);
  const T dl[3] = {A.dv[0] - B.dv[0], A.dv[1] - B.dv[1], A.dv[2] - B.dv[2]};
  const T jdvB = (e[0] * A.wt[0] + e[1] * A.wt[1] + e[2] * A.wt[2]) + (e[3] * B.wt[0] + f[0] * B.wt[1] + f[1] * B.wt[2]) +
                 (f[2] * dl[0] + f[3] * dl[1] + g[0] * dl[2]);
  const T jdvA = (a[0] * A.wt[0] + a[1] * A.wt[1] + a[2] * A.wt[2]) + (a[3] * B.wt[0] + b[0] * B.wt[1] + b[1] * B.wt[2]) +
                 (b[2] * dl[0] + b[3] * dl[1] + c[0] * dl[2]);
  const T uB = g[1] - apB * cfm - jdvB, uA = c[1] - apA * cfm - jdvA;
  T dB = uB * g[2], dA = uA * c[2];
  const T res = cone_clip(apA, apB, dA, dB, uA, uB, DA, DB, L);
  {
    const T la = A.invM * dA, lb = B.invM * dA;
    A.wt[0] += a[0] * dA; A.wt[1] += a[1] * dA; A.wt[2] += a[2] * dA;
    B.wt[0] += a[3] * dA; B.wt[1] += b[0] * dA; B.wt[2] += b[1] * dA;
    A.dv[0] += b[2] * la; A.dv[1] += b[3] * la; A.dv[2] += c[0] * la;
    B.dv[0] -= b[2] * lb; B.dv[1] -= b[3] * lb; B.dv[2] -= c[0] * lb;
  }
  {
    const T la = A.invM * dB, lb = B.invM * dB;
    A.wt[0] += e[0] * dB; A.wt[1] += e[1] * dB; A.wt[2] += e[2] * dB;
    B.wt[0] += e[3] * dB; B.wt[1] += f[0] * dB; B.wt[2] += f[1] * dB;
    A.dv[0] += f[2] * la; A.dv[1] += f[3] * la; A.dv[2] += g[0] * la;
    B.dv[0] -= f[2] * lb; B.dv[1] -= f[3] * lb; B.dv[2] -= g[0] * lb;
  }
  return res;
}

// One projected Gauss-Seidel iteration over a lane's column: ground-pole, cart-pole,
// ground-cart normal rows, then their friction rows (manifold_order 0).  mx[m] = slots to
// sweep per pair (warp-uniform; a lane's unused slots are zero rows).  Every row is loaded one
// row ahead of its use (the first rows of the three blocks at the top), so the shared-memory
// latency hides behind the previous row's dependent arithmetic.  Returns the largest squared
// residual.
template <typename T, bool FRICTION, bool SMEM = false>
BCP_HD T solve_iteration(const Params<T>& P, const T* colp, const SolveLayout& L, const int mx[3], SolveRegs<T>& R) {
  T lsr = 0;
  const T cfm_n = P.cfm_normal, cfm_f = P.cfm_friction;
#define BCP_ACC(r_) { const T r__ = (r_); const T r2__ = r__ * r__; lsr = (lsr > r2__) ? lsr : r2__; }
// rows are swept two at a time between the warp-uniform exits (slots past a lane's count are zero
// rows), through two register buffers used alternately: row j + 1 is in flight while row j runs
#define BCP_SWEEP(PAIR, N_, LOAD_NEXT, SOLVE)             \
  if (mx[PAIR] > 0) {                                      \
    _Pragma("unroll") for (int j = 0; j < 4; ++j) {        \
      if (j == 2 && mx[PAIR] <= 2) break;                  \
      if (j < 3) { LOAD_NEXT; }                            \
      BCP_ACC(SOLVE);                                      \
    }                                                      \
  }
  const T* const gp = colp + L.base[BCP_PAIR_GROUND_POLE];
  const T* const cp = colp + L.base[BCP_PAIR_CART_POLE];
  const T* const gc = colp + L.base[BCP_PAIR_GROUND_CART];
  SRow1<T> r_gp[2], r_gc[2];
  SRow2<T> r_cp[2];
  load_row<SMEM>(gp, r_gp[0]);
  load_row<SMEM>(cp, r_cp[0]);
  load_row<SMEM>(gc, r_gc[0]);
  // first friction pair of every pair that has friction rows, also at the top
  SRow1<T> f_gp[2][2], f_gc[2][2];
  SRow2<T> f_cp[2][2];
  const T* const fgp = colp + L.fbase[BCP_PAIR_GROUND_POLE];
  const T* const fcp = colp + L.fbase[BCP_PAIR_CART_POLE];
  const T* const fgc = colp + L.fbase[BCP_PAIR_GROUND_CART];
  if (FRICTION) {
    if (L.fbase[BCP_PAIR_GROUND_POLE] >= 0) { load_row<SMEM>(fgp, f_gp[0][0]); load_row<SMEM>(fgp + SW1, f_gp[0][1]); }
    if (L.fbase[BCP_PAIR_CART_POLE] >= 0) { load_row<SMEM>(fcp, f_cp[0][0]); load_row<SMEM>(fcp + SW2, f_cp[0][1]); }
    if (L.fbase[BCP_PAIR_GROUND_CART] >= 0) { load_row<SMEM>(fgc, f_gc[0][0]); load_row<SMEM>(fgc + SW1, f_gc[0][1]); }
  }
  BCP_SWEEP(BCP_PAIR_GROUND_POLE, 4, load_row<SMEM>(gp + (j + 1) * SW1, r_gp[(j + 1) & 1]),
            solve_row1(r_gp[j & 1], R.pole, R.apn[SLOT_GP + j], cfm_n))
  BCP_SWEEP(BCP_PAIR_CART_POLE, 4, load_row<SMEM>(cp + (j + 1) * SW2, r_cp[(j + 1) & 1]),
            solve_row2(r_cp[j & 1], R.cart, R.pole, R.apn[SLOT_CP + j], cfm_n))
  BCP_SWEEP(BCP_PAIR_GROUND_CART, 4, load_row<SMEM>(gc + (j + 1) * SW1, r_gc[(j + 1) & 1]),
            solve_row1(r_gc[j & 1], R.cart, R.apn[SLOT_GC + j], cfm_n))
  if (FRICTION) {
    if (L.fbase[BCP_PAIR_GROUND_POLE] >= 0) {
      const T mu = P.mu[BCP_PAIR_GROUND_POLE];
      BCP_SWEEP(BCP_PAIR_GROUND_POLE, 4,
                load_row<SMEM>(fgp + (j + 1) * 2 * SW1, f_gp[(j + 1) & 1][0]); load_row<SMEM>(fgp + (j + 1) * 2 * SW1 + SW1, f_gp[(j + 1) & 1][1]),
                solve_cone1(f_gp[j & 1][0], f_gp[j & 1][1], R.pole, R.apf[SLOT_GP + j][0], R.apf[SLOT_GP + j][1], mu * R.apn[SLOT_GP + j], cfm_f))
    }
    if (L.fbase[BCP_PAIR_CART_POLE] >= 0) {
      const T mu = P.mu[BCP_PAIR_CART_POLE];
      BCP_SWEEP(BCP_PAIR_CART_POLE, 4,
                load_row<SMEM>(fcp + (j + 1) * 2 * SW2, f_cp[(j + 1) & 1][0]); load_row<SMEM>(fcp + (j + 1) * 2 * SW2 + SW2, f_cp[(j + 1) & 1][1]),
                solve_cone2(f_cp[j & 1][0], f_cp[j & 1][1], R.cart, R.pole, R.apf[SLOT_CP + j][0], R.apf[SLOT_CP + j][1], mu * R.apn[SLOT_CP + j], cfm_f))
    }
    if (L.fbase[BCP_PAIR_GROUND_CART] >= 0) {
      const T mu = P.mu[BCP_PAIR_GROUND_CART];
      BCP_SWEEP(BCP_PAIR_GROUND_CART, 4,
                load_row<SMEM>(fgc + (j + 1) * 2 * SW1, f_gc[(j + 1) & 1][0]); load_row<SMEM>(fgc + (j + 1) * 2 * SW1 + SW1, f_gc[(j + 1) & 1][1]),
                solve_cone1(f_gc[j & 1][0], f_gc[j & 1][1], R.cart, R.apf[SLOT_GC + j][0], R.apf[SLOT_GC + j][1], mu * R.apn[SLOT_GC + j], cfm_f))
    }
  }
#undef BCP_SWEEP
#undef BCP_ACC
  return lsr;
}

// ---- writing rows in this format (k_prep) ----
// one-body row (A = static ground, B = body): J = [R^T (rB x -axis), -axis]
template <typename T>
BCP_HD T put_row1(const Params<T>& P, T* row, FBody<T>& B, const T pb[3], const T axis[3], T dist, T rest,
                  bool is_friction, T warm_impulse) {
  T j[3], d[3];
  local_jac(B, pb, axis, T(-1.0), j);
  for (int i = 0; i < 3; ++i) d[i] = -axis[i];
  const T cfm = is_friction ? P.cfm_friction : P.cfm_normal;
  const T den = ang_denom(B, j) + B.invM * dot3(d, d) + cfm;
  const T invD = (den > Lim<T>::eps()) ? T(1.0) / den : T(0.0);
  const T rel_vel = dot3(j, B.wl) + dot3(d, B.v);
  T w[SW1];  // assembled in registers, stored as whole 16-byte pieces
  for (int i = 0; i < 3; ++i) { w[S1_J + i] = j[i] * B.sI[i]; w[S1_D + i] = d[i]; }
  w[S1_RHSV] = (invD > T(0.0)) ? row_rhs(P, rel_vel, dist, rest, T(-1.0), is_friction) : T(0.0);
  w[S1_INVD] = invD;
  st4(row, w);
  st4(row + 4, w + 4);
  if (warm_impulse != T(0.0)) {
    for (int i = 0; i < 3; ++i) { B.dwl[i] += j[i] * B.invI[i] * warm_impulse; B.dv[i] += d[i] * (B.invM * warm_impulse); }
  }
  return warm_impulse;
}
template <typename T>
BCP_HD T put_row2(const Params<T>& P, T* row, FBody<T>& A, FBody<T>& B, const T pa[3], const T pb[3], const T axis[3],
                  T dist, T rest, bool is_friction, T warm_impulse) {
  T ja[3], jb[3];
  local_jac(A, pa, axis, T(1.0), ja);
  local_jac(B, pb, axis, T(-1.0), jb);
  const T cfm = is_friction ? P.cfm_friction : P.cfm_normal;
  const T nn = dot3(axis, axis);
  const T den = (ang_denom(A, ja) + A.invM * nn) + (ang_denom(B, jb) + B.invM * nn) + cfm;
  const T invD = (den > Lim<T>::eps()) ? T(1.0) / den : T(0.0);
  const T rel_vel = (dot3(ja, A.wl) + dot3(axis, A.v)) + (dot3(jb, B.wl) - dot3(axis, B.v));
  T w[SW2];
  for (int i = 0; i < 3; ++i) { w[S2_JA + i] = ja[i] * A.sI[i]; w[S2_JB + i] = jb[i] * B.sI[i]; w[S2_N + i] = axis[i]; }
  w[S2_RHSV] = (invD > T(0.0)) ? row_rhs(P, rel_vel, dist, rest, T(-1.0), is_friction) : T(0.0);
  w[S2_INVD] = invD;
  w[11] = 0;
  st4(row, w);
  st4(row + 4, w + 4);
  st4(row + 8, w + 8);
  if (warm_impulse != T(0.0)) {
    for (int i = 0; i < 3; ++i) {
      A.dwl[i] += ja[i] * A.invI[i] * warm_impulse; A.dv[i] += axis[i] * (A.invM * warm_impulse);
      B.dwl[i] += jb[i] * B.invI[i] * warm_impulse; B.dv[i] -= axis[i] * (B.invM * warm_impulse);
    }
  }
  return warm_impulse;
}
// rows of manifold m - normal rows into the pair's normal block, the friction rows along
// btPlaneSpace1(normal) pairwise into its friction block - and the warm-start impulses of the
// normal rows into apn[4]
template <typename T, bool FRICTION, bool TWO>
BCP_HD void put_rows(const Params<T>& P, T* nrows, T* frows, T* apn, int m, int rpc, const Manifold<T>& mf, const WorldPt<T>* wp,
                     FBody<T>& A, FBody<T>& B) {
  const int W = TWO ? (int)SW2 : (int)SW1;
  const T rest = P.rest[m];
#pragma unroll
  for (int j = 0; j < 4; ++j) {  // compile-time indices into mf.p / wp (they stay in registers)
    if (j >= mf.count) break;
    T* row = nrows + j * W;
    const T warm = (P.warmstart > T(0.0)) ? mf.p[j].imp * P.warmstart : T(0.0);
    if (TWO) apn[j] = put_row2(P, row, A, B, wp[j].pa, wp[j].pb, mf.p[j].n, mf.p[j].dist, rest, false, warm);
    else apn[j] = put_row1(P, row, B, wp[j].pb, mf.p[j].n, mf.p[j].dist, rest, false, warm);
    if (FRICTION && rpc > 1) {
      T d1[3], d2[3];
      plane_space(mf.p[j].n, d1, d2);
      T* fr = frows + j * (rpc - 1) * W;
      if (TWO) put_row2(P, fr, A, B, wp[j].pa, wp[j].pb, d1, T(0), rest, true, T(0));
      else put_row1(P, fr, B, wp[j].pb, d1, T(0), rest, true, T(0));
      if (rpc > 2) {
        if (TWO) put_row2(P, fr + W, A, B, wp[j].pa, wp[j].pb, d2, T(0), rest, true, T(0));
        else put_row1(P, fr + W, B, wp[j].pb, d2, T(0), rest, true, T(0));
      }
    }
  }
}

}  // namespace bcp
