// bcp_boxbox.cuh — dBoxBox2 (included by bcp_collide.cuh).
// Box 1 = body A of the pair, box 2 = body B.  Emits <= 4 contacts in Bullet's
// emission order; the normal is normalOnB (pointing from B towards A).
#pragma once

/* ATTRIBUTION.  The routine below follows, step for step, `dBoxBox2` and its helpers (`dLineClosestApproach`,
 * `intersectRectQuad2`, `cullPoints2`) as published in Bullet Physics' btBoxBoxDetector.cpp - itself taken from the
 * Open Dynamics Engine.  Reproducing that routine's decisions (axis order, fudge factors, emission order) is what the
 * path requires ("reproduces Bullet's ..."), so its structure and names are kept recognisable.
 *   Open Dynamics Engine, Copyright (C) 2001-2003 Russell L. Smith. All rights reserved.
 *     Used under the BSD-style license of ODE (LICENSE-BSD.TXT): redistribution in source and binary forms, with or
 *     without modification, permitted provided the copyright notice, the conditions and the disclaimer are retained;
 *     THIS SOFTWARE IS PROVIDED "AS IS", WITHOUT WARRANTY OF ANY KIND.
 *   Bullet Continuous Collision Detection and Physics Library, Copyright (c) 2003-2006 Erwin Coumans.
 *     zlib license: provided 'as-is' without warranty; altered source versions must be plainly marked as such - THIS IS
 *     AN ALTERED VERSION (restated from the published algorithm, re-laid for a per-thread workspace), not the original
 *     software; this notice may not be removed. */

namespace bcp {

// Workspace of one call, supplied by the caller (k_prep: the thread's shared-memory frame, so
// that the clipper's register-indexed arrays never touch the thread's stack; elsewhere a local
// array).  The <= 4 contacts come back in its first words, over the clipper's arrays, which are
// dead by then.
constexpr int kBoxWs = 72;
template <typename T> BCP_HD const Contact<T>* box_contacts(const T* ws) { return reinterpret_cast<const Contact<T>*>(ws); }

template <typename T>
BCP_HD_NOINLINE int box_box(const T p1[3], const T* R1, const T h1[3], const T p2[3], const T* R2, const T h2[3], T* ws) {
  Contact<T>* const out = reinterpret_cast<Contact<T>*>(ws);                                   // [0, 28)
  T *const quad = ws, *const ret = ws + 8, *const buffer = ws + 24, *const point = ws + 40, *const dep = ws + 64;
  const T fudge_factor = T(1.05);
  const T fudge2 = (T)1.0e-5f;
  T p[3], pp[3], normalC[3] = {0, 0, 0}, normal[3];
  const T* normalR = nullptr;
  int normalR_col = 0;
  T A[3], B[3], Rm[3][3], Q[3][3], s, s2, l;
  int invert_normal = 0, code = 0;

  p[0] = p2[0] - p1[0]; p[1] = p2[1] - p1[1]; p[2] = p2[2] - p1[2];
  pp[0] = dot_cv(R1, 0, p); pp[1] = dot_cv(R1, 1, p); pp[2] = dot_cv(R1, 2, p);
  for (int i = 0; i < 3; ++i) { A[i] = (T(2.0) * h1[i]) * T(0.5); B[i] = (T(2.0) * h2[i]) * T(0.5); }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      Rm[i][j] = dot_cc(R1, i, R2, j);
      Q[i][j] = babs(Rm[i][j]);
    }
  s = -3.402823466e+38f;  // -dInfinity

#define BCP_TST1(expr1, expr2, nR, col, cc) \
  {                                         \
    const T e1 = (expr1);                   \
    s2 = babs(e1) - (expr2);                \
    if (s2 > 0) return 0;                   \
    if (s2 > s) {                           \
      s = s2;                               \
      normalR = nR;                         \
      normalR_col = col;                    \
      invert_normal = (e1 < 0);             \
      code = (cc);                          \
    }                                       \
  }
  BCP_TST1(pp[0], (A[0] + B[0] * Q[0][0] + B[1] * Q[0][1] + B[2] * Q[0][2]), R1, 0, 1);
  BCP_TST1(pp[1], (A[1] + B[0] * Q[1][0] + B[1] * Q[1][1] + B[2] * Q[1][2]), R1, 1, 2);
  BCP_TST1(pp[2], (A[2] + B[0] * Q[2][0] + B[1] * Q[2][1] + B[2] * Q[2][2]), R1, 2, 3);
  BCP_TST1(dot_cv(R2, 0, p), (A[0] * Q[0][0] + A[1] * Q[1][0] + A[2] * Q[2][0] + B[0]), R2, 0, 4);
  BCP_TST1(dot_cv(R2, 1, p), (A[0] * Q[0][1] + A[1] * Q[1][1] + A[2] * Q[2][1] + B[1]), R2, 1, 5);
  BCP_TST1(dot_cv(R2, 2, p), (A[0] * Q[0][2] + A[1] * Q[1][2] + A[2] * Q[2][2] + B[2]), R2, 2, 6);
#undef BCP_TST1

#define BCP_TST2(expr1, expr2, n1, n2, n3, cc)             \
  {                                                        \
    const T e1 = (expr1);                                  \
    s2 = babs(e1) - (expr2);                               \
    if (s2 > Lim<T>::eps()) return 0;                      \
    l = bsqrt((n1) * (n1) + (n2) * (n2) + (n3) * (n3));    \
    if (l > Lim<T>::eps()) {                               \
      s2 /= l;                                             \
      if (s2 * fudge_factor > s) {                         \
        s = s2;                                            \
        normalR = nullptr;                                 \
        normalC[0] = (n1) / l;                             \
        normalC[1] = (n2) / l;                             \
        normalC[2] = (n3) / l;                             \
        invert_normal = (e1 < 0);                          \
        code = (cc);                                       \
      }                                                    \
    }                                                      \
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Q[i][j] += fudge2;
  BCP_TST2(pp[2] * Rm[1][0] - pp[1] * Rm[2][0], (A[1] * Q[2][0] + A[2] * Q[1][0] + B[1] * Q[0][2] + B[2] * Q[0][1]), T(0.0), -Rm[2][0], Rm[1][0], 7);
  BCP_TST2(pp[2] * Rm[1][1] - pp[1] * Rm[2][1], (A[1] * Q[2][1] + A[2] * Q[1][1] + B[0] * Q[0][2] + B[2] * Q[0][0]), T(0.0), -Rm[2][1], Rm[1][1], 8);
  BCP_TST2(pp[2] * Rm[1][2] - pp[1] * Rm[2][2], (A[1] * Q[2][2] + A[2] * Q[1][2] + B[0] * Q[0][1] + B[1] * Q[0][0]), T(0.0), -Rm[2][2], Rm[1][2], 9);
  BCP_TST2(pp[0] * Rm[2][0] - pp[2] * Rm[0][0], (A[0] * Q[2][0] + A[2] * Q[0][0] + B[1] * Q[1][2] + B[2] * Q[1][1]), Rm[2][0], T(0.0), -Rm[0][0], 10);
  BCP_TST2(pp[0] * Rm[2][1] - pp[2] * Rm[0][1], (A[0] * Q[2][1] + A[2] * Q[0][1] + B[0] * Q[1][2] + B[2] * Q[1][0]), Rm[2][1], T(0.0), -Rm[0][1], 11);
  BCP_TST2(pp[0] * Rm[2][2] - pp[2] * Rm[0][2], (A[0] * Q[2][2] + A[2] * Q[0][2] + B[0] * Q[1][1] + B[1] * Q[1][0]), Rm[2][2], T(0.0), -Rm[0][2], 12);
  BCP_TST2(pp[1] * Rm[0][0] - pp[0] * Rm[1][0], (A[0] * Q[1][0] + A[1] * Q[0][0] + B[1] * Q[2][2] + B[2] * Q[2][1]), -Rm[1][0], Rm[0][0], T(0.0), 13);
  BCP_TST2(pp[1] * Rm[0][1] - pp[0] * Rm[1][1], (A[0] * Q[1][1] + A[1] * Q[0][1] + B[0] * Q[2][2] + B[2] * Q[2][0]), -Rm[1][1], Rm[0][1], T(0.0), 14);
  BCP_TST2(pp[1] * Rm[0][2] - pp[0] * Rm[1][2], (A[0] * Q[1][2] + A[1] * Q[0][2] + B[0] * Q[2][1] + B[1] * Q[2][0]), -Rm[1][2], Rm[0][2], T(0.0), 15);
#undef BCP_TST2

  if (!code) return 0;

  if (normalR) {
    normal[0] = BCP_COL(normalR, normalR_col, 0);
    normal[1] = BCP_COL(normalR, normalR_col, 1);
    normal[2] = BCP_COL(normalR, normalR_col, 2);
  } else {
    mat_mul_vec(R1, normalC, normal);
  }
  if (invert_normal) { normal[0] = -normal[0]; normal[1] = -normal[1]; normal[2] = -normal[2]; }
  const T depth = -s;

  if (code > 6) {
    // edge-edge: a single contact, at the closest point on box 2's edge
    T pa[3], pb[3], ua[3], ub[3], alpha, beta, sign;
    for (int i = 0; i < 3; ++i) pa[i] = p1[i];
    for (int j = 0; j < 3; ++j) {
      sign = (dot_vc(normal, R1, j) > 0) ? T(1.0) : T(-1.0);
      for (int i = 0; i < 3; ++i) pa[i] += sign * A[j] * BCP_COL(R1, j, i);
    }
    for (int i = 0; i < 3; ++i) pb[i] = p2[i];
    for (int j = 0; j < 3; ++j) {
      sign = (dot_vc(normal, R2, j) > 0) ? T(-1.0) : T(1.0);
      for (int i = 0; i < 3; ++i) pb[i] += sign * B[j] * BCP_COL(R2, j, i);
    }
    const int ca = (code - 7) / 3, cb = (code - 7) % 3;
    for (int i = 0; i < 3; ++i) ua[i] = BCP_COL(R1, ca, i);
    for (int i = 0; i < 3; ++i) ub[i] = BCP_COL(R2, cb, i);
    line_closest_approach(pa, ua, pb, ub, alpha, beta);
    for (int i = 0; i < 3; ++i) pa[i] += ua[i] * alpha;
    for (int i = 0; i < 3; ++i) pb[i] += ub[i] * beta;
    for (int i = 0; i < 3; ++i) { out[0].n[i] = -normal[i]; out[0].pb[i] = pb[i]; }
    out[0].depth = -depth;
    return 1;
  }

  // face-something: reference face on box a, incident face on box b
  const T *Ra, *Rb, *pa, *pb, *Sa, *Sb;
  T normal2[3], nr[3], anr[3], center[3], rect[2];
  int lanr, a1, a2, codeN, code1, code2, maxc = 4;
  if (code <= 3) { Ra = R1; Rb = R2; pa = p1; pb = p2; Sa = A; Sb = B; }
  else { Ra = R2; Rb = R1; pa = p2; pb = p1; Sa = B; Sb = A; }
  if (code <= 3) { normal2[0] = normal[0]; normal2[1] = normal[1]; normal2[2] = normal[2]; }
  else { normal2[0] = -normal[0]; normal2[1] = -normal[1]; normal2[2] = -normal[2]; }
  nr[0] = dot_cv(Rb, 0, normal2); nr[1] = dot_cv(Rb, 1, normal2); nr[2] = dot_cv(Rb, 2, normal2);
  anr[0] = babs(nr[0]); anr[1] = babs(nr[1]); anr[2] = babs(nr[2]);
  if (anr[1] > anr[0]) {
    if (anr[1] > anr[2]) { a1 = 0; lanr = 1; a2 = 2; }
    else { a1 = 0; a2 = 1; lanr = 2; }
  } else {
    if (anr[0] > anr[2]) { lanr = 0; a1 = 1; a2 = 2; }
    else { a1 = 0; a2 = 1; lanr = 2; }
  }
  if (nr[lanr] < 0) {
    for (int i = 0; i < 3; ++i) center[i] = pb[i] - pa[i] + Sb[lanr] * BCP_COL(Rb, lanr, i);
  } else {
    for (int i = 0; i < 3; ++i) center[i] = pb[i] - pa[i] - Sb[lanr] * BCP_COL(Rb, lanr, i);
  }
  codeN = (code <= 3) ? code - 1 : code - 4;
  if (codeN == 0) { code1 = 1; code2 = 2; }
  else if (codeN == 1) { code1 = 0; code2 = 2; }
  else { code1 = 0; code2 = 1; }
  const T c1 = dot_vc(center, Ra, code1);
  const T c2 = dot_vc(center, Ra, code2);
  T m11 = dot_cc(Ra, code1, Rb, a1);
  T m12 = dot_cc(Ra, code1, Rb, a2);
  T m21 = dot_cc(Ra, code2, Rb, a1);
  T m22 = dot_cc(Ra, code2, Rb, a2);
  {
    const T k1 = m11 * Sb[a1], k2 = m21 * Sb[a1], k3 = m12 * Sb[a2], k4 = m22 * Sb[a2];
    quad[0] = c1 - k1 - k3; quad[1] = c2 - k2 - k4;
    quad[2] = c1 - k1 + k3; quad[3] = c2 - k2 + k4;
    quad[4] = c1 + k1 + k3; quad[5] = c2 + k2 + k4;
    quad[6] = c1 + k1 - k3; quad[7] = c2 + k2 - k4;
  }
  rect[0] = Sa[code1];
  rect[1] = Sa[code2];
  // The incident quad strictly inside the reference rectangle - a foot on the cart, a body well inside the ground
  // plate - is the common case, and there intersectRectQuad2 is the identity: each of its four passes finds every point
  // on the inner side (sign * p[dir] < h[dir] for both signs <=> |p[dir]| < h[dir]), copies it, and sees no crossing.
  // Same four points in the same order, without the clipper.
  bool inside = true;
  for (int i = 0; i < 4; ++i) inside = inside && (babs(quad[2 * i]) < rect[0]) && (babs(quad[2 * i + 1]) < rect[1]);
  int n = 4;
  if (inside) {
    for (int i = 0; i < 8; ++i) ret[i] = quad[i];
  } else {
    n = intersect_rect_quad(rect, quad, ret, buffer);
  }
  if (n < 1) return 0;
  const T det1 = T(1.0) / (m11 * m22 - m12 * m21);
  m11 *= det1; m12 *= det1; m21 *= det1; m22 *= det1;
  int cnum = 0;
  for (int j = 0; j < n; ++j) {
    const T k1 = m22 * (ret[j * 2] - c1) - m12 * (ret[j * 2 + 1] - c2);
    const T k2 = -m21 * (ret[j * 2] - c1) + m11 * (ret[j * 2 + 1] - c2);
    for (int i = 0; i < 3; ++i) point[cnum * 3 + i] = center[i] + k1 * BCP_COL(Rb, a1, i) + k2 * BCP_COL(Rb, a2, i);
    dep[cnum] = Sa[codeN] - dot3(normal2, point + cnum * 3);
    if (dep[cnum] >= 0) {
      ret[cnum * 2] = ret[j * 2];
      ret[cnum * 2 + 1] = ret[j * 2 + 1];
      cnum++;
    }
  }
  if (cnum < 1) return 0;
  if (maxc > cnum) maxc = cnum;
  if (maxc < 1) maxc = 1;
  if (cnum <= maxc) {
    for (int j = 0; j < cnum; ++j) {
      for (int i = 0; i < 3; ++i) {
        out[j].n[i] = -normal[i];
        if (code < 4) out[j].pb[i] = point[j * 3 + i] + pa[i];
        else out[j].pb[i] = point[j * 3 + i] + pa[i] - normal[i] * dep[j];
      }
      out[j].depth = -dep[j];
    }
    return cnum;
  }
  int i1 = 0;
  T maxdepth = dep[0];
  for (int i = 1; i < cnum; ++i)
    if (dep[i] > maxdepth) { maxdepth = dep[i]; i1 = i; }
  const unsigned picks = cull_points(cnum, ret, maxc, i1, buffer);
  for (int j = 0; j < maxc; ++j) {
    const int ij = (int)((picks >> (4 * j)) & 15u);
    T pw[3];
    for (int i = 0; i < 3; ++i) pw[i] = point[ij * 3 + i] + pa[i];
    for (int i = 0; i < 3; ++i) {
      out[j].n[i] = -normal[i];
      if (code < 4) out[j].pb[i] = pw[i];
      else out[j].pb[i] = pw[i] - normal[i] * dep[ij];
    }
    out[j].depth = -dep[ij];
  }
  return maxc;
}

}  // namespace bcp
