/*
 * bcp_oracle.c — CPU ORACLE.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the BulletCartpole reset/step path of
 * rebcabin/cartpoleplusplus (`bullet_cartpole.py`) INCLUDING the part the
 * reference delegates to the third-party `pybullet` module (un-vendored,
 * un-pinned; not importable in this image).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library.
 *
 * PARITY STATUS: **parity unpinned** against pybullet.  The reference ships no
 * golden vectors for this path and pybullet cannot be imported here, so every
 * engine-side fact below tagged [BULLET-RECALL] restates the published Bullet
 * Physics SDK algorithm (bullet3, `src/BulletDynamics/Featherstone`,
 * `src/BulletCollision`, `examples/SharedMemory/PhysicsServerCommandProcessor.cpp`)
 * from its public sources and is held as a PARAMETER in bcp_config, never a
 * literal.  What IS pinned: the bookkeeping/call order of bullet_cartpole.py
 * (cited per function), the model constants of models/{ground,cart,pole}.urdf
 * (tests/golden/model_constants.json), the README's exact known-answer stability
 * protocol (README.md:77-80) and analytic known answers (tests/test_oracle_kat.py).
 *
 * Compiled twice (double, and float with -DBCPO_FP32); every exported symbol
 * carries the precision suffix.  Compile with -ffp-contract=off: the arithmetic
 * order written here is the specification the FP64 CUDA kernel reproduces
 * bit-for-bit.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/bcp.h"

#ifdef BCPO_FP32
typedef float real;
#define SFX(name) name##_f32
#define R_EPS 1.1920928955078125e-7f /* SIMD_EPSILON = FLT_EPSILON */
#define R_LARGE 1e18f                /* BT_LARGE_FLOAT */
#define R_SQRT(x) sqrtf(x)
#define R_FABS(x) fabsf(x)
#else
typedef double real;
#define SFX(name) name##_f64
#define R_EPS 2.220446049250313e-16 /* SIMD_EPSILON = DBL_EPSILON */
#define R_LARGE 1e18
#define R_SQRT(x) sqrt(x)
#define R_FABS(x) fabs(x)
#endif
#define R(x) ((real)(x))

/* ======================================================================== */
/* deterministic elementary functions                                        */
/* (shared *definition* with the CUDA kernel, so FP64 parity can be bit-exact: */
/*  libm and CUDA libdevice agree only to an ulp)                             */
/* ======================================================================== */

/* sin/cos on |x| <= pi/4: Taylor in Horner form, one rounding per operation */
static real sin_core(real x) {
  const real x2 = x * x;
  real p = R(-2.8114572543455206e-15); /* -1/17! */
  p = p * x2 + R(7.647163731819816e-13);  /* +1/15! */
  p = p * x2 + R(-1.6059043836821613e-10); /* -1/13! */
  p = p * x2 + R(2.505210838544172e-08);  /* +1/11! */
  p = p * x2 + R(-2.7557319223985893e-06); /* -1/9! */
  p = p * x2 + R(0.0001984126984126984);  /* +1/7! */
  p = p * x2 + R(-0.008333333333333333);  /* -1/5! */
  p = p * x2 + R(0.16666666666666666);    /* +1/3! */
  return x - (x * x2) * p; /* x - x^3 (1/3! - x^2/5! + ...) */
}
static real cos_core(real x) {
  const real x2 = x * x;
  real p = R(4.779477332387385e-14);  /* +1/16! */
  p = p * x2 + R(-1.1470745597729725e-11); /* -1/14! */
  p = p * x2 + R(2.08767569878681e-09);   /* +1/12! */
  p = p * x2 + R(-2.755731922398589e-07); /* -1/10! */
  p = p * x2 + R(2.48015873015873e-05);   /* +1/8! */
  p = p * x2 + R(-0.001388888888888889);  /* -1/6! */
  p = p * x2 + R(0.041666666666666664);   /* +1/4! */
  p = p * x2 + R(-0.5);                   /* -1/2! */
  return R(1.0) + x2 * p;
}

#define PI_D 3.141592653589793

/* (sin, cos) of theta = u * 2*pi for u in [0,1) given exactly (u = k/2^24).
 * Octant reduction is exact in binary; the core runs on |r| <= pi/4. */
static void sincos_2pi(real u, real* s_out, real* c_out) {
  const real t = u * R(8.0);          /* exact */
  const int oct = (int)t;             /* 0..7 */
  const int q = (oct + 1) >> 1;       /* nearest multiple of pi/2: 0..4 */
  const real f = t - R(2.0) * (real)q; /* exact, in [-1,1) */
  const real r = f * R(PI_D / 4.0);
  const real s = sin_core(r), c = cos_core(r);
  switch (q & 3) {
    case 0: *s_out = s; *c_out = c; break;
    case 1: *s_out = c; *c_out = -s; break;
    case 2: *s_out = -s; *c_out = -c; break;
    default: *s_out = -c; *c_out = s; break;
  }
}

/* atan2 used only to order clipped polygon vertices (cullPoints2).  Three
 * half-angle reductions bring the argument under 0.1, then Taylor. */
static real atan_unit(real t) { /* t in [0,1] */
  int k;
  real p, t2;
  for (k = 0; k < 3; ++k) t = t / (R(1.0) + R_SQRT(R(1.0) + t * t));
  t2 = t * t;
  p = R(1.0 / 17.0);
  p = R(-1.0 / 15.0) + t2 * p;
  p = R(1.0 / 13.0) + t2 * p;
  p = R(-1.0 / 11.0) + t2 * p;
  p = R(1.0 / 9.0) + t2 * p;
  p = R(-1.0 / 7.0) + t2 * p;
  p = R(1.0 / 5.0) + t2 * p;
  p = R(-1.0 / 3.0) + t2 * p;
  p = R(1.0) + t2 * p;
  return R(8.0) * (t * p);
}
static real det_atan2(real y, real x) {
  const real ay = R_FABS(y), ax = R_FABS(x);
  real a;
  if (ax == R(0.0) && ay == R(0.0)) return R(0.0);
  if (ay <= ax)
    a = atan_unit(ay / ax);
  else
    a = R(PI_D / 2.0) - atan_unit(ax / ay);
  if (x < R(0.0)) a = R(PI_D) - a;
  return (y < R(0.0)) ? -a : a;
}

/* ======================================================================== */
/* Philox4x32-10 counter RNG (Salmon et al., SC'11), the in-kernel generator  */
/* ======================================================================== */
static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  int r;
  for (r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
    const uint32_t n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    const uint32_t n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
}
#define RNG_PUSH 1u
#define RNG_ACTION 2u
static void env_rng(const bcp_config* cfg, int64_t global_env, uint32_t episode,
                    uint32_t purpose, uint32_t index, uint32_t out[4]) {
  out[0] = (uint32_t)((uint64_t)global_env & 0xffffffffu);
  out[1] = (uint32_t)((uint64_t)global_env >> 32);
  out[2] = episode;
  out[3] = (purpose << 24) | (index & 0xffffffu);
  philox4x32_10(out, (uint32_t)(cfg->seed & 0xffffffffu), (uint32_t)(cfg->seed >> 32));
}
static real u24(uint32_t x) { return (real)(x >> 8) * R(5.9604644775390625e-08); } /* [0,1) */

/* ======================================================================== */
/* small vector helpers (operation order is part of the spec)                 */
/* ======================================================================== */
static real dot3(const real* a, const real* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static void cross3(const real* a, const real* b, real* o) {
  const real x = a[1] * b[2] - a[2] * b[1];
  const real y = a[2] * b[0] - a[0] * b[2];
  const real z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
/* btMatrix3x3::setRotation(q), q = (x,y,z,w): body->world basis, row major */
static void quat_to_mat(const real* q, real* m) {
  const real d = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  const real s = R(2.0) / d;
  const real xs = q[0] * s, ys = q[1] * s, zs = q[2] * s;
  const real wx = q[3] * xs, wy = q[3] * ys, wz = q[3] * zs;
  const real xx = q[0] * xs, xy = q[0] * ys, xz = q[0] * zs;
  const real yy = q[1] * ys, yz = q[1] * zs, zz = q[2] * zs;
  m[0] = R(1.0) - (yy + zz); m[1] = xy - wz; m[2] = xz + wy;
  m[3] = xy + wz; m[4] = R(1.0) - (xx + zz); m[5] = yz - wx;
  m[6] = xz - wy; m[7] = yz + wx; m[8] = R(1.0) - (xx + yy);
}
static void mat_mul_vec(const real* m, const real* v, real* o) { /* o = M v */
  const real x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  const real y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  const real z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
static void matT_mul_vec(const real* m, const real* v, real* o) { /* o = M^T v */
  const real x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2];
  const real y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2];
  const real z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}

/* ======================================================================== */
/* working state                                                              */
/* ======================================================================== */
typedef struct {
  real x[3], q[4], v[3], w[3];
} obody;
typedef struct {
  real la[3], lb[3], n[3], dist; /* localPointA/B, normalWorldOnB, distance1 */
  real imp;                      /* m_appliedImpulse (warm-start cache) */
} opoint;
typedef struct {
  int count;
  opoint p[4];
} omanifold;
typedef struct {
  obody cart, pole;
  real pend_f[3], pend_t[3];
  omanifold mani[3];
  int steps, done, episode;
  int reset_left; /* pipelined auto-reset only (bcp.h BCP_AUTO_RESET_PIPELINED): reset substeps still to run */
} ostate;

static void load_body(const double* s, obody* b) {
  int i;
  for (i = 0; i < 3; ++i) b->x[i] = (real)s[i];
  for (i = 0; i < 4; ++i) b->q[i] = (real)s[3 + i];
  for (i = 0; i < 3; ++i) b->v[i] = (real)s[7 + i];
  for (i = 0; i < 3; ++i) b->w[i] = (real)s[10 + i];
}
static void store_body(const obody* b, double* s) {
  int i;
  for (i = 0; i < 3; ++i) s[i] = b->x[i];
  for (i = 0; i < 4; ++i) s[3 + i] = b->q[i];
  for (i = 0; i < 3; ++i) s[7 + i] = b->v[i];
  for (i = 0; i < 3; ++i) s[10 + i] = b->w[i];
}
static void load_state(const double* s, ostate* st) {
  int m, k, i;
  load_body(s + BCP_S_CART, &st->cart);
  load_body(s + BCP_S_POLE, &st->pole);
  for (i = 0; i < 3; ++i) st->pend_f[i] = (real)s[BCP_S_PEND_F + i];
  for (i = 0; i < 3; ++i) st->pend_t[i] = (real)s[BCP_S_PEND_T + i];
  for (m = 0; m < 3; ++m) {
    st->mani[m].count = (int)s[BCP_S_MANI_CNT + m];
    for (k = 0; k < 4; ++k) {
      const double* p = s + BCP_S_MANI + m * BCP_MANI_STRIDE + k * BCP_MANI_POINT;
      opoint* o = &st->mani[m].p[k];
      for (i = 0; i < 3; ++i) { o->la[i] = (real)p[i]; o->lb[i] = (real)p[3 + i]; o->n[i] = (real)p[6 + i]; }
      o->dist = (real)p[9];
      o->imp = (real)p[10];
    }
  }
  st->steps = (int)s[BCP_S_STEPS];
  st->done = (int)s[BCP_S_DONE];
  st->episode = (int)s[BCP_S_EPISODE];
  st->reset_left = (int)s[BCP_S_RESET_LEFT];
}
static void store_state(const ostate* st, double* s) {
  int m, k, i;
  store_body(&st->cart, s + BCP_S_CART);
  store_body(&st->pole, s + BCP_S_POLE);
  for (i = 0; i < 3; ++i) s[BCP_S_PEND_F + i] = st->pend_f[i];
  for (i = 0; i < 3; ++i) s[BCP_S_PEND_T + i] = st->pend_t[i];
  for (m = 0; m < 3; ++m) {
    s[BCP_S_MANI_CNT + m] = st->mani[m].count;
    for (k = 0; k < 4; ++k) {
      double* p = s + BCP_S_MANI + m * BCP_MANI_STRIDE + k * BCP_MANI_POINT;
      const opoint* o = &st->mani[m].p[k];
      const int live = k < st->mani[m].count;
      for (i = 0; i < 3; ++i) {
        p[i] = live ? o->la[i] : 0.0;
        p[3 + i] = live ? o->lb[i] : 0.0;
        p[6 + i] = live ? o->n[i] : 0.0;
      }
      p[9] = live ? o->dist : 0.0;
      p[10] = live ? o->imp : 0.0;
    }
  }
  s[BCP_S_STEPS] = st->steps;
  s[BCP_S_DONE] = st->done;
  s[BCP_S_EPISODE] = st->episode;
  for (i = BCP_S_EPISODE + 1; i < BCP_STATE_DOUBLES; ++i) s[i] = 0.0;
  s[BCP_S_RESET_LEFT] = st->reset_left;
}

/* ======================================================================== */
/* Collision: btBoxBoxDetector (ODE dBoxBox) + btPersistentManifold           */
/* [BULLET-RECALL] src/BulletCollision/CollisionDispatch/btBoxBoxDetector.cpp  */
/* ======================================================================== */
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
typedef struct {
  real n[3];     /* normalOnBInWorld */
  real pb[3];    /* pointInWorld (on B) */
  real depth;    /* signed distance (negative = penetration) */
} ocontact;

/* intersectRectQuad2: clip quad p against rectangle +-h; returns point count */
static int intersect_rect_quad(const real h[2], real p[8], real ret[16]) {
  int nq = 4, nr = 0, dir, sign, i;
  real buffer[16];
  real* q = p;
  real* r = ret;
  for (dir = 0; dir <= 1; ++dir) {
    for (sign = -1; sign <= 1; sign += 2) {
      real* pq = q;
      real* pr = r;
      nr = 0;
      for (i = nq; i > 0; --i) {
        real* nextq;
        if ((real)sign * pq[dir] < h[dir]) {
          pr[0] = pq[0];
          pr[1] = pq[1];
          pr += 2;
          nr++;
          if (nr & 8) { q = r; goto done; }
        }
        nextq = (i > 1) ? pq + 2 : q;
        if (((real)sign * pq[dir] < h[dir]) ^ ((real)sign * nextq[dir] < h[dir])) {
          pr[1 - dir] = pq[1 - dir] +
                        (nextq[1 - dir] - pq[1 - dir]) / (nextq[dir] - pq[dir]) * ((real)sign * h[dir] - pq[dir]);
          pr[dir] = (real)sign * h[dir];
          pr += 2;
          nr++;
          if (nr & 8) { q = r; goto done; }
        }
        pq += 2;
      }
      q = r;
      r = (q == ret) ? buffer : ret;
      nq = nr;
    }
  }
done:
  if (q != ret) memcpy(ret, q, (size_t)nr * 2 * sizeof(real));
  return nr;
}

/* cullPoints2: pick m of n polygon points spread in angle around the centroid */
static void cull_points(int n, const real p[], int m, int i0, int iret[]) {
  int i, j;
  real a, cx, cy, q;
  real A[8];
  int avail[8];
  const float M__PI = 3.14159265f; /* Bullet's float literal */
  if (n == 1) {
    cx = p[0];
    cy = p[1];
  } else if (n == 2) {
    cx = R(0.5) * (p[0] + p[2]);
    cy = R(0.5) * (p[1] + p[3]);
  } else {
    a = 0; cx = 0; cy = 0;
    for (i = 0; i < (n - 1); ++i) {
      q = p[i * 2] * p[i * 2 + 3] - p[i * 2 + 2] * p[i * 2 + 1];
      a += q;
      cx += q * (p[i * 2] + p[i * 2 + 2]);
      cy += q * (p[i * 2 + 1] + p[i * 2 + 3]);
    }
    q = p[n * 2 - 2] * p[1] - p[0] * p[n * 2 - 1];
    if (R_FABS(a + q) > R_EPS)
      a = R(1.0) / (R(3.0) * (a + q));
    else
      a = R_LARGE;
    cx = a * (cx + q * (p[n * 2 - 2] + p[0]));
    cy = a * (cy + q * (p[n * 2 - 1] + p[1]));
  }
  for (i = 0; i < n; ++i) A[i] = det_atan2(p[i * 2 + 1] - cy, p[i * 2] - cx);
  for (i = 0; i < n; ++i) avail[i] = 1;
  avail[i0] = 0;
  iret[0] = i0;
  iret++;
  for (j = 1; j < m; ++j) {
    real maxdiff = R(1e9), diff;
    a = (real)j * (real)(2 * M__PI / (float)m) + A[i0];
    if (a > (real)M__PI) a -= (real)(2 * M__PI);
    *iret = i0;
    for (i = 0; i < n; ++i) {
      if (avail[i]) {
        diff = R_FABS(A[i] - a);
        if (diff > (real)M__PI) diff = (real)(2 * M__PI) - diff;
        if (diff < maxdiff) {
          maxdiff = diff;
          *iret = i;
        }
      }
    }
    avail[*iret] = 0;
    iret++;
  }
}

/* dLineClosestApproach */
static void line_closest_approach(const real pa[3], const real ua[3], const real pb[3], const real ub[3],
                                  real* alpha, real* beta) {
  real p[3], uaub, q1, q2, d;
  p[0] = pb[0] - pa[0];
  p[1] = pb[1] - pa[1];
  p[2] = pb[2] - pa[2];
  uaub = dot3(ua, ub);
  q1 = dot3(ua, p);
  q2 = -dot3(ub, p);
  d = R(1.0) - uaub * uaub;
  if (d <= R(0.0001f)) {
    *alpha = 0;
    *beta = 0;
  } else {
    d = R(1.0) / d;
    *alpha = (q1 + uaub * q2) * d;
    *beta = (uaub * q1 + q2) * d;
  }
}

/* column j of a row-major 3x3 */
#define COL(M, j, i) ((M)[(i) * 3 + (j)])
static real dot_col_col(const real* A, int ja, const real* B, int jb) { /* dDOT44 */
  return COL(A, ja, 0) * COL(B, jb, 0) + COL(A, ja, 1) * COL(B, jb, 1) + COL(A, ja, 2) * COL(B, jb, 2);
}
static real dot_col_vec(const real* A, int ja, const real* v) { /* dDOT41 */
  return COL(A, ja, 0) * v[0] + COL(A, ja, 1) * v[1] + COL(A, ja, 2) * v[2];
}
static real dot_vec_col(const real* v, const real* A, int ja) { /* dDOT14 */
  return v[0] * COL(A, ja, 0) + v[1] * COL(A, ja, 1) + v[2] * COL(A, ja, 2);
}

/* dBoxBox2.  p1,R1,h1: box 1 (= body A); p2,R2,h2: box 2 (= body B); h = half
 * extents (Bullet passes 2*halfExtents and halves it again: exact).  Returns the
 * number of contacts written (0..4), in Bullet's emission order. */
static __thread int g_sat_code; /* the separating-axis code of the last box_box call (0: separated), for test instrumentation */
static int box_box_impl(const real p1[3], const real* R1, const real h1[3], const real p2[3], const real* R2,
                        const real h2[3], ocontact out[4], int* code_out);
static int box_box(const real p1[3], const real* R1, const real h1[3], const real p2[3], const real* R2,
                   const real h2[3], ocontact out[4]) {
  int code = 0;
  const int n = box_box_impl(p1, R1, h1, p2, R2, h2, out, &code);
  g_sat_code = code;
  return n;
}
static int box_box_impl(const real p1[3], const real* R1, const real h1[3], const real p2[3], const real* R2,
                        const real h2[3], ocontact out[4], int* code_out) {
  const real fudge_factor = R(1.05);
  real p[3], pp[3], normalC[3] = {0, 0, 0}, normal[3];
  const real* normalR = 0;
  int normalR_col = 0;
  real A[3], B[3], Rm[3][3], Q[3][3], s, s2, l, depth;
  int i, j, invert_normal, code;
  const real fudge2 = (real)1.0e-5f;

  p[0] = p2[0] - p1[0]; p[1] = p2[1] - p1[1]; p[2] = p2[2] - p1[2];
  pp[0] = dot_col_vec(R1, 0, p); pp[1] = dot_col_vec(R1, 1, p); pp[2] = dot_col_vec(R1, 2, p);
  for (i = 0; i < 3; ++i) { A[i] = (R(2.0) * h1[i]) * R(0.5); B[i] = (R(2.0) * h2[i]) * R(0.5); }
  for (i = 0; i < 3; ++i)
    for (j = 0; j < 3; ++j) {
      Rm[i][j] = dot_col_col(R1, i, R2, j);
      Q[i][j] = R_FABS(Rm[i][j]);
    }

  s = -3.402823466e+38f; /* -dInfinity (FLT_MAX) */
  invert_normal = 0;
  code = 0;

#define TST1(expr1, expr2, nR, col, cc)      \
  {                                          \
    const real e1 = (expr1);                 \
    s2 = R_FABS(e1) - (expr2);               \
    if (s2 > 0) return 0;                    \
    if (s2 > s) {                            \
      s = s2;                                \
      normalR = nR;                          \
      normalR_col = col;                     \
      invert_normal = (e1 < 0);              \
      code = (cc);                           \
    }                                        \
  }
  TST1(pp[0], (A[0] + B[0] * Q[0][0] + B[1] * Q[0][1] + B[2] * Q[0][2]), R1, 0, 1);
  TST1(pp[1], (A[1] + B[0] * Q[1][0] + B[1] * Q[1][1] + B[2] * Q[1][2]), R1, 1, 2);
  TST1(pp[2], (A[2] + B[0] * Q[2][0] + B[1] * Q[2][1] + B[2] * Q[2][2]), R1, 2, 3);
  TST1(dot_col_vec(R2, 0, p), (A[0] * Q[0][0] + A[1] * Q[1][0] + A[2] * Q[2][0] + B[0]), R2, 0, 4);
  TST1(dot_col_vec(R2, 1, p), (A[0] * Q[0][1] + A[1] * Q[1][1] + A[2] * Q[2][1] + B[1]), R2, 1, 5);
  TST1(dot_col_vec(R2, 2, p), (A[0] * Q[0][2] + A[1] * Q[1][2] + A[2] * Q[2][2] + B[2]), R2, 2, 6);
#undef TST1

#define TST2(expr1, expr2, n1, n2, n3, cc)                      \
  {                                                             \
    const real e1 = (expr1);                                    \
    s2 = R_FABS(e1) - (expr2);                                  \
    if (s2 > R_EPS) return 0;                                   \
    l = R_SQRT((n1) * (n1) + (n2) * (n2) + (n3) * (n3));        \
    if (l > R_EPS) {                                            \
      s2 /= l;                                                  \
      if (s2 * fudge_factor > s) {                              \
        s = s2;                                                 \
        normalR = 0;                                            \
        normalC[0] = (n1) / l;                                  \
        normalC[1] = (n2) / l;                                  \
        normalC[2] = (n3) / l;                                  \
        invert_normal = (e1 < 0);                               \
        code = (cc);                                            \
      }                                                         \
    }                                                           \
  }
  for (i = 0; i < 3; ++i)
    for (j = 0; j < 3; ++j) Q[i][j] += fudge2;
  /* u1 x (v1,v2,v3) */
  TST2(pp[2] * Rm[1][0] - pp[1] * Rm[2][0], (A[1] * Q[2][0] + A[2] * Q[1][0] + B[1] * Q[0][2] + B[2] * Q[0][1]), R(0.0), -Rm[2][0], Rm[1][0], 7);
  TST2(pp[2] * Rm[1][1] - pp[1] * Rm[2][1], (A[1] * Q[2][1] + A[2] * Q[1][1] + B[0] * Q[0][2] + B[2] * Q[0][0]), R(0.0), -Rm[2][1], Rm[1][1], 8);
  TST2(pp[2] * Rm[1][2] - pp[1] * Rm[2][2], (A[1] * Q[2][2] + A[2] * Q[1][2] + B[0] * Q[0][1] + B[1] * Q[0][0]), R(0.0), -Rm[2][2], Rm[1][2], 9);
  /* u2 x (v1,v2,v3) */
  TST2(pp[0] * Rm[2][0] - pp[2] * Rm[0][0], (A[0] * Q[2][0] + A[2] * Q[0][0] + B[1] * Q[1][2] + B[2] * Q[1][1]), Rm[2][0], R(0.0), -Rm[0][0], 10);
  TST2(pp[0] * Rm[2][1] - pp[2] * Rm[0][1], (A[0] * Q[2][1] + A[2] * Q[0][1] + B[0] * Q[1][2] + B[2] * Q[1][0]), Rm[2][1], R(0.0), -Rm[0][1], 11);
  TST2(pp[0] * Rm[2][2] - pp[2] * Rm[0][2], (A[0] * Q[2][2] + A[2] * Q[0][2] + B[0] * Q[1][1] + B[1] * Q[1][0]), Rm[2][2], R(0.0), -Rm[0][2], 12);
  /* u3 x (v1,v2,v3) */
  TST2(pp[1] * Rm[0][0] - pp[0] * Rm[1][0], (A[0] * Q[1][0] + A[1] * Q[0][0] + B[1] * Q[2][2] + B[2] * Q[2][1]), -Rm[1][0], Rm[0][0], R(0.0), 13);
  TST2(pp[1] * Rm[0][1] - pp[0] * Rm[1][1], (A[0] * Q[1][1] + A[1] * Q[0][1] + B[0] * Q[2][2] + B[2] * Q[2][0]), -Rm[1][1], Rm[0][1], R(0.0), 14);
  TST2(pp[1] * Rm[0][2] - pp[0] * Rm[1][2], (A[0] * Q[1][2] + A[1] * Q[0][2] + B[0] * Q[2][1] + B[1] * Q[2][0]), -Rm[1][2], Rm[0][2], R(0.0), 15);
#undef TST2

  if (!code) return 0;
  *code_out = code;

  if (normalR) {
    normal[0] = COL(normalR, normalR_col, 0);
    normal[1] = COL(normalR, normalR_col, 1);
    normal[2] = COL(normalR, normalR_col, 2);
  } else {
    mat_mul_vec(R1, normalC, normal);
  }
  if (invert_normal) { normal[0] = -normal[0]; normal[1] = -normal[1]; normal[2] = -normal[2]; }
  depth = -s;

  if (code > 6) {
    /* edge-edge: one contact at the closest point on box 2's edge */
    real pa[3], pb[3], ua[3], ub[3], alpha, beta, sign;
    for (i = 0; i < 3; ++i) pa[i] = p1[i];
    for (j = 0; j < 3; ++j) {
      sign = (dot_vec_col(normal, R1, j) > 0) ? R(1.0) : R(-1.0);
      for (i = 0; i < 3; ++i) pa[i] += sign * A[j] * COL(R1, j, i);
    }
    for (i = 0; i < 3; ++i) pb[i] = p2[i];
    for (j = 0; j < 3; ++j) {
      sign = (dot_vec_col(normal, R2, j) > 0) ? R(-1.0) : R(1.0);
      for (i = 0; i < 3; ++i) pb[i] += sign * B[j] * COL(R2, j, i);
    }
    for (i = 0; i < 3; ++i) ua[i] = COL(R1, (code - 7) / 3, i);
    for (i = 0; i < 3; ++i) ub[i] = COL(R2, (code - 7) % 3, i);
    line_closest_approach(pa, ua, pb, ub, &alpha, &beta);
    for (i = 0; i < 3; ++i) pa[i] += ua[i] * alpha;
    for (i = 0; i < 3; ++i) pb[i] += ub[i] * beta;
    for (i = 0; i < 3; ++i) { out[0].n[i] = -normal[i]; out[0].pb[i] = pb[i]; }
    out[0].depth = -depth;
    return 1;
  }

  {
    /* face-something: reference face on box a, incident face on box b */
    const real *Ra, *Rb, *pa, *pb, *Sa, *Sb;
    real normal2[3], nr[3], anr[3], center[3], quad[8], rect[2], ret[16], point[3 * 8], dep[8];
    real c1, c2, m11, m12, m21, m22, det1;
    int lanr, a1, a2, codeN, code1, code2, n, cnum, maxc = 4;
    if (code <= 3) { Ra = R1; Rb = R2; pa = p1; pb = p2; Sa = A; Sb = B; }
    else { Ra = R2; Rb = R1; pa = p2; pb = p1; Sa = B; Sb = A; }
    if (code <= 3) { normal2[0] = normal[0]; normal2[1] = normal[1]; normal2[2] = normal[2]; }
    else { normal2[0] = -normal[0]; normal2[1] = -normal[1]; normal2[2] = -normal[2]; }
    nr[0] = dot_col_vec(Rb, 0, normal2); nr[1] = dot_col_vec(Rb, 1, normal2); nr[2] = dot_col_vec(Rb, 2, normal2);
    anr[0] = R_FABS(nr[0]); anr[1] = R_FABS(nr[1]); anr[2] = R_FABS(nr[2]);
    if (anr[1] > anr[0]) {
      if (anr[1] > anr[2]) { a1 = 0; lanr = 1; a2 = 2; }
      else { a1 = 0; a2 = 1; lanr = 2; }
    } else {
      if (anr[0] > anr[2]) { lanr = 0; a1 = 1; a2 = 2; }
      else { a1 = 0; a2 = 1; lanr = 2; }
    }
    if (nr[lanr] < 0) {
      for (i = 0; i < 3; ++i) center[i] = pb[i] - pa[i] + Sb[lanr] * COL(Rb, lanr, i);
    } else {
      for (i = 0; i < 3; ++i) center[i] = pb[i] - pa[i] - Sb[lanr] * COL(Rb, lanr, i);
    }
    codeN = (code <= 3) ? code - 1 : code - 4;
    if (codeN == 0) { code1 = 1; code2 = 2; }
    else if (codeN == 1) { code1 = 0; code2 = 2; }
    else { code1 = 0; code2 = 1; }
    c1 = dot_vec_col(center, Ra, code1);
    c2 = dot_vec_col(center, Ra, code2);
    m11 = dot_col_col(Ra, code1, Rb, a1);
    m12 = dot_col_col(Ra, code1, Rb, a2);
    m21 = dot_col_col(Ra, code2, Rb, a1);
    m22 = dot_col_col(Ra, code2, Rb, a2);
    {
      const real k1 = m11 * Sb[a1], k2 = m21 * Sb[a1], k3 = m12 * Sb[a2], k4 = m22 * Sb[a2];
      quad[0] = c1 - k1 - k3; quad[1] = c2 - k2 - k4;
      quad[2] = c1 - k1 + k3; quad[3] = c2 - k2 + k4;
      quad[4] = c1 + k1 + k3; quad[5] = c2 + k2 + k4;
      quad[6] = c1 + k1 - k3; quad[7] = c2 + k2 - k4;
    }
    rect[0] = Sa[code1];
    rect[1] = Sa[code2];
    n = intersect_rect_quad(rect, quad, ret);
    if (n < 1) return 0;
    det1 = R(1.0) / (m11 * m22 - m12 * m21);
    m11 *= det1; m12 *= det1; m21 *= det1; m22 *= det1;
    cnum = 0;
    for (j = 0; j < n; ++j) {
      const real k1 = m22 * (ret[j * 2] - c1) - m12 * (ret[j * 2 + 1] - c2);
      const real k2 = -m21 * (ret[j * 2] - c1) + m11 * (ret[j * 2 + 1] - c2);
      for (i = 0; i < 3; ++i) point[cnum * 3 + i] = center[i] + k1 * COL(Rb, a1, i) + k2 * COL(Rb, a2, i);
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
      for (j = 0; j < cnum; ++j) {
        for (i = 0; i < 3; ++i) {
          out[j].n[i] = -normal[i];
          if (code < 4) out[j].pb[i] = point[j * 3 + i] + pa[i];
          else out[j].pb[i] = point[j * 3 + i] + pa[i] - normal[i] * dep[j];
        }
        out[j].depth = -dep[j];
      }
      return cnum;
    } else {
      int i1 = 0, iret[8];
      real maxdepth = dep[0];
      for (i = 1; i < cnum; ++i)
        if (dep[i] > maxdepth) { maxdepth = dep[i]; i1 = i; }
      cull_points(cnum, ret, maxc, i1, iret);
      for (j = 0; j < maxc; ++j) {
        real pw[3];
        for (i = 0; i < 3; ++i) pw[i] = point[iret[j] * 3 + i] + pa[i];
        for (i = 0; i < 3; ++i) {
          out[j].n[i] = -normal[i];
          if (code < 4) out[j].pb[i] = pw[i];
          else out[j].pb[i] = pw[i] - normal[i] * dep[iret[j]];
        }
        out[j].depth = -dep[iret[j]];
      }
      return maxc;
    }
  }
}

/* a posed box as the narrowphase sees it */
typedef struct {
  real x[3];
  real R[9];
  real h[3];
} oposed;

/* btPersistentManifold::sortCachedPoints (gContactCalcArea3Points = true) */
static int sort_cached_points(const omanifold* m, const opoint* pt) {
  int maxPenetrationIndex = -1, i, best;
  real maxPenetration = pt->dist, res[4] = {0, 0, 0, 0}, a[3], b[3], c[3], maxVal;
  for (i = 0; i < 4; ++i)
    if (m->p[i].dist < maxPenetration) { maxPenetrationIndex = i; maxPenetration = m->p[i].dist; }
#define SUB3(o, u, v) { o[0] = (u)[0] - (v)[0]; o[1] = (u)[1] - (v)[1]; o[2] = (u)[2] - (v)[2]; }
  if (maxPenetrationIndex != 0) { SUB3(a, pt->la, m->p[1].la); SUB3(b, m->p[3].la, m->p[2].la); cross3(a, b, c); res[0] = dot3(c, c); }
  if (maxPenetrationIndex != 1) { SUB3(a, pt->la, m->p[0].la); SUB3(b, m->p[3].la, m->p[2].la); cross3(a, b, c); res[1] = dot3(c, c); }
  if (maxPenetrationIndex != 2) { SUB3(a, pt->la, m->p[0].la); SUB3(b, m->p[3].la, m->p[1].la); cross3(a, b, c); res[2] = dot3(c, c); }
  if (maxPenetrationIndex != 3) { SUB3(a, pt->la, m->p[0].la); SUB3(b, m->p[2].la, m->p[1].la); cross3(a, b, c); res[3] = dot3(c, c); }
#undef SUB3
  /* btVector4::closestAxis4 = maxAxis4 of |.| */
  best = -1;
  maxVal = -R_LARGE;
  for (i = 0; i < 4; ++i) {
    const real v = R_FABS(res[i]);
    if (i < 3) { if (v > maxVal) { best = i; maxVal = v; } }
    else { if (v > maxVal) { best = i; } }
  }
  return best;
}

/* btManifoldResult::addContactPoint + btPersistentManifold cache logic */
static void manifold_add(omanifold* m, const oposed* A, const oposed* B, const ocontact* c, real breaking) {
  opoint np;
  real pointA[3], v[3], shortest;
  int i, nearest = -1, insert;
  if (c->depth > breaking) return;
  for (i = 0; i < 3; ++i) pointA[i] = c->pb[i] + c->n[i] * c->depth;
  for (i = 0; i < 3; ++i) v[i] = pointA[i] - A->x[i];
  matT_mul_vec(A->R, v, np.la); /* invXform */
  for (i = 0; i < 3; ++i) v[i] = c->pb[i] - B->x[i];
  matT_mul_vec(B->R, v, np.lb);
  for (i = 0; i < 3; ++i) np.n[i] = c->n[i];
  np.dist = c->depth;
  np.imp = 0;
  /* getCacheEntry */
  shortest = breaking * breaking;
  for (i = 0; i < m->count; ++i) {
    real d[3], dd;
    d[0] = m->p[i].la[0] - np.la[0]; d[1] = m->p[i].la[1] - np.la[1]; d[2] = m->p[i].la[2] - np.la[2];
    dd = dot3(d, d);
    if (dd < shortest) { shortest = dd; nearest = i; }
  }
  if (nearest >= 0) {
    np.imp = m->p[nearest].imp; /* replaceContactPoint keeps m_appliedImpulse */
    m->p[nearest] = np;
  } else {
    insert = m->count;
    if (insert == 4) insert = sort_cached_points(m, &np);
    else m->count++;
    if (insert < 0) insert = 0;
    m->p[insert] = np;
  }
}

/* btPersistentManifold::refreshContactPoints; fills world points for the solver */
typedef struct {
  real pa[3], pb[3]; /* positionWorldOnA / OnB */
} oworld;
static void manifold_refresh(omanifold* m, const oposed* A, const oposed* B, real breaking, oworld wp[4]) {
  int i, k;
  for (i = m->count - 1; i >= 0; --i) {
    opoint* p = &m->p[i];
    real d[3];
    mat_mul_vec(A->R, p->la, wp[i].pa);
    mat_mul_vec(B->R, p->lb, wp[i].pb);
    for (k = 0; k < 3; ++k) { wp[i].pa[k] += A->x[k]; wp[i].pb[k] += B->x[k]; }
    for (k = 0; k < 3; ++k) d[k] = wp[i].pa[k] - wp[i].pb[k];
    p->dist = dot3(d, p->n);
  }
  for (i = m->count - 1; i >= 0; --i) {
    opoint* p = &m->p[i];
    int remove = 0;
    if (!(p->dist <= breaking)) {
      remove = 1;
    } else {
      real proj[3], diff[3], d2;
      for (k = 0; k < 3; ++k) proj[k] = wp[i].pa[k] - p->n[k] * p->dist;
      for (k = 0; k < 3; ++k) diff[k] = wp[i].pb[k] - proj[k];
      d2 = dot3(diff, diff);
      if (d2 > breaking * breaking) remove = 1;
    }
    if (remove) {
      const int last = m->count - 1;
      if (i != last) { m->p[i] = m->p[last]; wp[i] = wp[last]; }
      m->count--;
    }
  }
}

/* ======================================================================== */
/* Dynamics: btMultiBody (0 links, floating base) + btMultiBodyConstraintSolver */
/* [BULLET-RECALL] src/BulletDynamics/Featherstone/btMultiBody.cpp,            */
/*                 btMultiBodyConstraintSolver.cpp, btMultiBodyDynamicsWorld.cpp */
/* ======================================================================== */
typedef struct {
  int dyn;        /* 0 => static */
  real mass, I[3];
  real R[9];      /* body->world at the start of the step */
  real x[3];
  real vel[6];    /* [omega, v] world (m_realBuf) */
  real dvel[6];   /* solver delta velocities */
} sbody;

typedef struct {
  int a, b;        /* indices into sbody array */
  real jacA[6], jacB[6], unitA[6], unitB[6];
  real jacDiagABInv, rhs, cfm, applied, lower, upper, friction;
  int friction_index; /* normal row this friction row belongs to */
} orow;

static real clampr(real v, real lo, real hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* computeAccelerationsArticulatedBodyAlgorithmMultiDof for num_links = 0 */
static void forward_dynamics(const bcp_config* cfg, sbody* b, const real F_ext[3], const real T_ext[3]) {
  const real dt = (real)cfg->dt;
  const real kl = (real)cfg->linear_damping, ka = (real)cfg->angular_damping;
  const real vmax = (real)cfg->max_coordinate_velocity;
  real F[3], wl[3], vl[3], za[3], zl[3], t[3], nrm, d, ra[3], rl[3], sa[3], sl[3], od[3], vd[3], wxv[3], Iw[3];
  int i;
  for (i = 0; i < 3; ++i) F[i] = F_ext[i] + (real)cfg->gravity[i] * b->mass;
  matT_mul_vec(b->R, b->vel, wl);     /* base-frame omega */
  matT_mul_vec(b->R, b->vel + 3, vl); /* base-frame v */
  matT_mul_vec(b->R, T_ext, t);
  for (i = 0; i < 3; ++i) za[i] = -t[i];
  matT_mul_vec(b->R, F, t);
  for (i = 0; i < 3; ++i) zl[i] = -t[i];
  /* damping */
  d = dot3(wl, wl);
  nrm = (d > R_EPS) ? R_SQRT(d) : R(0.0);
  for (i = 0; i < 3; ++i) za[i] += ((R(1.0) * b->I[i]) * wl[i]) * (ka + ka * nrm);
  d = dot3(vl, vl);
  nrm = (d > R_EPS) ? R_SQRT(d) : R(0.0);
  for (i = 0; i < 3; ++i) zl[i] += ((R(1.0) * b->mass) * vl[i]) * (kl + kl * nrm);
  /* gyroscopic */
  if (cfg->gyro_term) {
    real g[3];
    for (i = 0; i < 3; ++i) Iw[i] = b->I[i] * wl[i];
    cross3(wl, Iw, g);
    for (i = 0; i < 3; ++i) za[i] += g[i];
  }
  cross3(wl, vl, wxv);
  for (i = 0; i < 3; ++i) zl[i] += b->mass * wxv[i];
  /* solveImatrix */
  for (i = 0; i < 3; ++i) ra[i] = za[i] / b->I[i];
  {
    const real inv_m = R(1.0) / b->mass;
    for (i = 0; i < 3; ++i) rl[i] = zl[i] * inv_m;
  }
  for (i = 0; i < 3; ++i) { sa[i] = -ra[i]; sl[i] = -rl[i] + wxv[i]; }
  mat_mul_vec(b->R, sa, od);
  mat_mul_vec(b->R, sl, vd);
  for (i = 0; i < 3; ++i) {
    b->vel[i] = clampr(b->vel[i] + od[i] * dt, -vmax, vmax);
    b->vel[3 + i] = clampr(b->vel[3 + i] + vd[i] * dt, -vmax, vmax);
  }
}

/* fillContactJacobianMultiDof + calcAccelerationDeltasMultiDof for one side */
static void fill_side(const sbody* b, const real pos[3], const real n[3], real jac[6], real unit[6]) {
  real r[3], t[3], u[3];
  int i;
  for (i = 0; i < 3; ++i) r[i] = pos[i] - b->x[i];
  cross3(r, n, jac);
  for (i = 0; i < 3; ++i) jac[3 + i] = n[i];
  if (!b->dyn) {
    for (i = 0; i < 6; ++i) unit[i] = 0;
    return;
  }
  matT_mul_vec(b->R, jac, t);
  for (i = 0; i < 3; ++i) u[i] = t[i] / b->I[i];
  mat_mul_vec(b->R, u, unit);
  matT_mul_vec(b->R, jac + 3, t);
  {
    const real inv_m = R(1.0) / b->mass;
    for (i = 0; i < 3; ++i) u[i] = t[i] * inv_m;
  }
  mat_mul_vec(b->R, u, unit + 3);
}

/* setupMultiBodyContactConstraint */
static void setup_row(const bcp_config* cfg, orow* row, sbody* bodies, int ia, int ib, const real posA[3],
                      const real posB[3], const real axis[3], real distance_in, real comb_friction,
                      real comb_restitution, int is_friction) {
  const real dt = (real)cfg->dt;
  const real inv_dt = R(1.0) / dt;
  real cfm, erp, denom0 = 0, denom1 = 0, d, rel_vel = 0, restitution = 0, distance = 0, naxis[3];
  real positionalError, velocityError, penetrationImpulse, velocityImpulse;
  int i;
  row->a = ia;
  row->b = ib;
  if (is_friction) { cfm = (real)cfg->friction_cfm; erp = (real)cfg->friction_erp; }
  else { cfm = (real)cfg->global_cfm; erp = (real)cfg->erp2; }
  cfm *= inv_dt;
  fill_side(&bodies[ia], posA, axis, row->jacA, row->unitA);
  for (i = 0; i < 3; ++i) naxis[i] = -axis[i];
  fill_side(&bodies[ib], posB, naxis, row->jacB, row->unitB);
  if (bodies[ia].dyn) for (i = 0; i < 6; ++i) denom0 += row->jacA[i] * row->unitA[i];
  if (bodies[ib].dyn) for (i = 0; i < 6; ++i) denom1 += row->jacB[i] * row->unitB[i];
  d = denom0 + denom1 + cfm;
  row->jacDiagABInv = (d > R_EPS) ? R(1.0) / d : R(0.0);
  if (!is_friction) distance = distance_in + (real)cfg->linear_slop;
  if (bodies[ia].dyn) for (i = 0; i < 6; ++i) rel_vel += bodies[ia].vel[i] * row->jacA[i];
  if (bodies[ib].dyn) for (i = 0; i < 6; ++i) rel_vel += bodies[ib].vel[i] * row->jacB[i];
  row->friction = comb_friction;
  if (!is_friction) {
    /* restitutionCurve */
    if (R_FABS(rel_vel) < (real)cfg->restitution_velocity_threshold) restitution = 0;
    else restitution = comb_restitution * -rel_vel;
    if (restitution <= R(0.0)) restitution = 0;
  }
  positionalError = 0;
  velocityError = restitution - rel_vel;
  if (is_friction) {
    positionalError = -distance * erp / dt;
  } else if (distance > 0) {
    positionalError = 0;
    velocityError -= distance / dt;
  } else {
    positionalError = -distance * erp / dt;
  }
  penetrationImpulse = positionalError * row->jacDiagABInv;
  velocityImpulse = velocityError * row->jacDiagABInv;
  row->rhs = penetrationImpulse + velocityImpulse;
  if (!is_friction) { row->lower = 0; row->upper = R(1e10f); }
  else { row->lower = -row->friction; row->upper = row->friction; }
  row->cfm = cfm * row->jacDiagABInv;
  row->applied = 0; /* contact warm starting is disabled for btMultiBody; see warm_start_row */
  (void)erp;
}

static void apply_delta(sbody* b, const real unit[6], real impulse) {
  int i;
  if (!b->dyn) return;
  for (i = 0; i < 6; ++i) b->dvel[i] += unit[i] * impulse;
}
static real jdotdv(const sbody* b, const real jac[6]) {
  real s = 0;
  int i;
  if (!b->dyn) return 0;
  for (i = 0; i < 6; ++i) s += jac[i] * b->dvel[i];
  return s;
}

/* optional contact warm starting (SOLVER_USE_ARTICULATED_WARMSTARTING path of
 * setupMultiBodyContactConstraint): cached impulse x factor applied up front */
static void warm_start_row(orow* c, sbody* bodies, real cached, real factor) {
  c->applied = cached * factor;
  if (c->applied != R(0.0)) {
    apply_delta(&bodies[c->a], c->unitA, c->applied);
    apply_delta(&bodies[c->b], c->unitB, c->applied);
  }
}

static __thread unsigned g_clamp_hash; /* instrumentation only: never read by the arithmetic */
#define CLAMP_NOTE(bit) (g_clamp_hash = (g_clamp_hash ^ (unsigned)(bit)) * 16777619u)
/* resolveSingleConstraintRowGeneric */
static real resolve_row(orow* c, sbody* bodies) {
  real deltaImpulse = c->rhs - c->applied * c->cfm;
  const real dA = jdotdv(&bodies[c->a], c->jacA);
  const real dB = jdotdv(&bodies[c->b], c->jacB);
  real sum;
  deltaImpulse -= dA * c->jacDiagABInv;
  deltaImpulse -= dB * c->jacDiagABInv;
  sum = c->applied + deltaImpulse;
  if (sum < c->lower) { deltaImpulse = c->lower - c->applied; c->applied = c->lower; CLAMP_NOTE(1); }
  else if (sum > c->upper) { deltaImpulse = c->upper - c->applied; c->applied = c->upper; CLAMP_NOTE(2); }
  else { c->applied = sum; CLAMP_NOTE(0); }
  apply_delta(&bodies[c->a], c->unitA, deltaImpulse);
  apply_delta(&bodies[c->b], c->unitB, deltaImpulse);
  return deltaImpulse / c->jacDiagABInv;
}

/* resolveConeFrictionConstraintRows; the clip radius is evaluated without
 * atan2/sin/cos: sin(atan2(a,b)) = a/hypot, cos = b/hypot (identical in exact
 * arithmetic; keeps the kernel and this oracle on the same IEEE operations). */
static real resolve_cone(orow* cA, orow* cB, sbody* bodies) {
  real dIB = cB->rhs - cB->applied * cB->cfm, dIA, sumA, sumB;
  dIB -= jdotdv(&bodies[cB->a], cB->jacA) * cB->jacDiagABInv;
  dIB -= jdotdv(&bodies[cB->b], cB->jacB) * cB->jacDiagABInv;
  sumB = cB->applied + dIB;
  dIA = cA->rhs - cA->applied * cA->cfm;
  dIA -= jdotdv(&bodies[cA->a], cA->jacA) * cA->jacDiagABInv;
  dIA -= jdotdv(&bodies[cA->b], cA->jacB) * cA->jacDiagABInv;
  sumA = cA->applied + dIA;
  CLAMP_NOTE(sumA * sumA + sumB * sumB >= cA->lower * cB->lower ? 5 : 4);
  if (sumA * sumA + sumB * sumB >= cA->lower * cB->lower) {
    const real r = R_SQRT(sumA * sumA + sumB * sumB);
    const real sn = (r > R(0.0)) ? sumA / r : R(0.0);
    const real cs = (r > R(0.0)) ? sumB / r : R(1.0);
    const real clipA = R_FABS(cA->lower * sn);
    const real clipB = R_FABS(cB->lower * cs);
    if (sumA < -clipA) { dIA = -clipA - cA->applied; cA->applied = -clipA; }
    else if (sumA > clipA) { dIA = clipA - cA->applied; cA->applied = clipA; }
    else cA->applied = sumA;
    if (sumB < -clipB) { dIB = -clipB - cB->applied; cB->applied = -clipB; }
    else if (sumB > clipB) { dIB = clipB - cB->applied; cB->applied = clipB; }
    else cB->applied = sumB;
  } else {
    cA->applied = sumA;
    cB->applied = sumB;
  }
  apply_delta(&bodies[cA->a], cA->unitA, dIA);
  apply_delta(&bodies[cA->b], cA->unitB, dIA);
  apply_delta(&bodies[cB->a], cB->unitA, dIB);
  apply_delta(&bodies[cB->b], cB->unitB, dIB);
  return dIA / cA->jacDiagABInv + dIB / cB->jacDiagABInv;
}

/* btPlaneSpace1 + normalize */
static void plane_space(const real n[3], real p[3], real q[3]) {
  real a, k, l;
  int i;
  if (R_FABS(n[2]) > R(0.7071067811865475244008443621048490)) {
    a = n[1] * n[1] + n[2] * n[2];
    k = R(1.0) / R_SQRT(a);
    p[0] = 0; p[1] = -n[2] * k; p[2] = n[1] * k;
    q[0] = a * k; q[1] = -n[0] * p[2]; q[2] = n[0] * p[1];
  } else {
    a = n[0] * n[0] + n[1] * n[1];
    k = R(1.0) / R_SQRT(a);
    p[0] = -n[1] * k; p[1] = n[0] * k; p[2] = 0;
    q[0] = -n[2] * p[1]; q[1] = n[2] * p[0]; q[2] = a * k;
  }
  l = R(1.0) / R_SQRT(dot3(p, p));
  for (i = 0; i < 3; ++i) p[i] *= l;
  l = R(1.0) / R_SQRT(dot3(q, q));
  for (i = 0; i < 3; ++i) q[i] *= l;
}

static real pair_breaking(const bcp_config* cfg, const bcp_body* a, const bcp_body* b) {
  /* min over the two shapes of gContactBreakingThreshold * angular motion disc */
  const double ra = sqrt(4 * a->half_extents[0] * a->half_extents[0] + 4 * a->half_extents[1] * a->half_extents[1] +
                         4 * a->half_extents[2] * a->half_extents[2]) * 0.5;
  const double rb = sqrt(4 * b->half_extents[0] * b->half_extents[0] + 4 * b->half_extents[1] * b->half_extents[1] +
                         4 * b->half_extents[2] * b->half_extents[2]) * 0.5;
  const double ta = cfg->contact_breaking_factor * ra, tb = cfg->contact_breaking_factor * rb;
  return (real)(ta < tb ? ta : tb);
}
static real combine_friction(const bcp_config* cfg, const bcp_body* a, const bcp_body* b) {
  double f = a->friction * b->friction;
  if (f < -cfg->max_friction) f = -cfg->max_friction;
  if (f > cfg->max_friction) f = cfg->max_friction;
  return (real)f;
}

static void pose_box(const obody* b, const bcp_body* c, oposed* o) {
  int i;
  for (i = 0; i < 3; ++i) { o->x[i] = b->x[i]; o->h[i] = (real)c->half_extents[i]; }
  quat_to_mat(b->q, o->R);
}

#include <stdio.h>
static int g_debug = 0;
/* optional instrumentation for tests */
typedef struct {
  int iterations, rows, friction_rows;
  int contacts[3];
  int code[3]; /* dBoxBox2's separating-axis code per pair: 0 separated, 1-3 face of A, 4-6 face of B, 7-15 edge-edge */
  int active_hash; /* hash over all iterations of the solver's clamp decisions (normal row at its lower limit, friction pair
                      clipped to the cone): the active set of the projected Gauss-Seidel */
  int force_it;    /* INPUT, > 0: run exactly this many iterations, ignoring the residual early exit (test instrumentation:
                      "the oracle with the kernel's iteration count") */
} osubstep_info;

/* One p.stepSimulation() — bullet_cartpole.py:199,327,330.
 * [BULLET-RECALL] btDiscreteDynamicsWorld::internalSingleStepSimulation order:
 * collide (start-of-step poses) -> forces+unconstrained velocity -> contact
 * solve -> integrate transforms -> clear forces. */
static void substep(const bcp_config* cfg, ostate* st, osubstep_info* info) {
  oposed G, C, P;
  oworld wp[3][4];
  sbody bodies[3]; /* 0 ground, 1 cart, 2 pole */
  orow nrows[12], frows[24];
  int n_n = 0, n_f = 0, m, k, i, order[3], it;
  const int pairA[3] = {0, 1, 0}, pairB[3] = {1, 2, 2};
  const bcp_body* cb[3] = {&cfg->ground, &cfg->cart, &cfg->pole};
  const real dt = (real)cfg->dt;
  const real vmax = (real)cfg->max_coordinate_velocity;
  oposed* posed[3];
  real zero3[3] = {0, 0, 0};

  /* ---- 1. collision detection on start-of-step poses ---- */
  for (i = 0; i < 3; ++i) { G.x[i] = (real)cfg->ground.init_pos[i]; G.h[i] = (real)cfg->ground.half_extents[i]; }
  for (i = 0; i < 9; ++i) G.R[i] = (i % 4 == 0) ? R(1.0) : R(0.0);
  pose_box(&st->cart, &cfg->cart, &C);
  pose_box(&st->pole, &cfg->pole, &P);
  posed[0] = &G; posed[1] = &C; posed[2] = &P;
  for (m = 0; m < 3; ++m) {
    const oposed *A = posed[pairA[m]], *B = posed[pairB[m]];
    const real breaking = pair_breaking(cfg, cb[pairA[m]], cb[pairB[m]]);
    ocontact cs[4];
    const int nc = box_box(A->x, A->R, A->h, B->x, B->R, B->h, cs);
    if (info) info->code[m] = g_sat_code;
    for (k = 0; k < nc; ++k) manifold_add(&st->mani[m], A, B, &cs[k], breaking);
    manifold_refresh(&st->mani[m], A, B, breaking, wp[m]);
    if (info) info->contacts[m] = st->mani[m].count;
  }

  /* ---- 2. forces + unconstrained velocities ---- */
  memset(bodies, 0, sizeof(bodies));
  bodies[0].dyn = 0;
  for (i = 0; i < 3; ++i) bodies[0].x[i] = G.x[i];
  {
    const obody* ob[2] = {&st->cart, &st->pole};
    const oposed* op[2] = {&C, &P};
    for (k = 0; k < 2; ++k) {
      sbody* b = &bodies[k + 1];
      b->dyn = 1;
      b->mass = (real)cb[k + 1]->mass;
      for (i = 0; i < 3; ++i) { b->I[i] = (real)cb[k + 1]->inertia[i]; b->x[i] = ob[k]->x[i]; }
      for (i = 0; i < 9; ++i) b->R[i] = op[k]->R[i];
      for (i = 0; i < 3; ++i) { b->vel[i] = ob[k]->w[i]; b->vel[3 + i] = ob[k]->v[i]; }
    }
  }
  forward_dynamics(cfg, &bodies[1], st->pend_f, st->pend_t);
  forward_dynamics(cfg, &bodies[2], zero3, zero3);

  /* ---- 3. constraint rows (manifold order after the island sort) ---- */
  if (cfg->manifold_order == 0) { order[0] = 2; order[1] = 1; order[2] = 0; }
  else { order[0] = 0; order[1] = 1; order[2] = 2; }
  for (k = 0; k < 3; ++k) {
    const omanifold* mf;
    real mu, rest;
    int j;
    m = order[k];
    mf = &st->mani[m];
    mu = combine_friction(cfg, cb[pairA[m]], cb[pairB[m]]);
    if (m == BCP_PAIR_GROUND_CART && cfg->cart_ground_friction >= 0.0) { /* compound cart: the lower half touches the ground */
      bcp_body lower = cfg->cart;
      lower.friction = cfg->cart_ground_friction;
      mu = combine_friction(cfg, &cfg->ground, &lower);
    }
    rest = (real)(cb[pairA[m]]->restitution * cb[pairB[m]]->restitution);
    for (j = 0; j < mf->count; ++j) {
      const int fi = n_n;
      setup_row(cfg, &nrows[n_n], bodies, pairA[m], pairB[m], wp[m][j].pa, wp[m][j].pb, mf->p[j].n, mf->p[j].dist,
                mu, rest, 0);
      if (cfg->warmstart_factor > 0.0) warm_start_row(&nrows[n_n], bodies, mf->p[j].imp, (real)cfg->warmstart_factor);
      n_n++;
      if (mu != R(0.0)) {
        real d1[3], d2[3];
        plane_space(mf->p[j].n, d1, d2);
        setup_row(cfg, &frows[n_f], bodies, pairA[m], pairB[m], wp[m][j].pa, wp[m][j].pb, d1, 0, mu, rest, 1);
        frows[n_f].friction_index = fi;
        n_f++;
        if (cfg->friction_dirs == 2) {
          setup_row(cfg, &frows[n_f], bodies, pairA[m], pairB[m], wp[m][j].pa, wp[m][j].pb, d2, 0, mu, rest, 1);
          frows[n_f].friction_index = fi;
          n_f++;
        }
      }
    }
  }
  /* rows with combined friction 0 are skipped above: their limits are +-0, so
   * they would apply a zero impulse and report a zero residual. */

  /* ---- 4. projected Gauss-Seidel ---- */
  g_clamp_hash = 2166136261u;
  for (it = 0; it < cfg->solver_iterations; ++it) {
    real lsr = 0;
    int j;
    for (j = 0; j < n_n; ++j) {
      const real r = resolve_row(&nrows[j], bodies);
      const real r2 = r * r;
      lsr = (lsr > r2) ? lsr : r2;
    }
    if (cfg->friction_dirs == 2 && cfg->cone_friction) {
      for (j = 0; j + 1 < n_f; j += 2) {
        const real tot = nrows[frows[j].friction_index].applied;
        real r, r2;
        frows[j].lower = -(frows[j].friction * tot);
        frows[j].upper = frows[j].friction * tot;
        frows[j + 1].lower = -(frows[j + 1].friction * tot);
        frows[j + 1].upper = frows[j + 1].friction * tot;
        r = resolve_cone(&frows[j], &frows[j + 1], bodies);
        r2 = r * r;
        lsr = (lsr > r2) ? lsr : r2;
      }
    } else {
      for (j = 0; j < n_f; ++j) {
        const real tot = nrows[frows[j].friction_index].applied;
        if (tot > R(0.0)) {
          real r, r2;
          frows[j].lower = -(frows[j].friction * tot);
          frows[j].upper = frows[j].friction * tot;
          r = resolve_row(&frows[j], bodies);
          r2 = r * r;
          lsr = (lsr > r2) ? lsr : r2;
        }
      }
    }
    if (g_debug) {
      fprintf(stderr, "it %d lsr %.3e imp:", it, (double)lsr);
      for (j = 0; j < n_n; ++j) fprintf(stderr, " %.5f", (double)nrows[j].applied);
      fprintf(stderr, "\n");
    }
    if (info && info->force_it > 0) {
      if (it + 1 >= info->force_it) { it++; break; }
      continue;
    }
    if (lsr <= (real)cfg->residual_threshold || it >= cfg->solver_iterations - 1) { it++; break; }
  }
  if (info) { info->iterations = it; info->rows = n_n; info->friction_rows = n_f; info->active_hash = (int)(g_clamp_hash & 0x7fffffffu); }

  /* solveGroupCacheFriendlyFinish: applied normal impulses -> manifold cache */
  {
    int row = 0, j;
    for (k = 0; k < 3; ++k) {
      m = order[k];
      for (j = 0; j < st->mani[m].count; ++j) st->mani[m].p[j].imp = nrows[row++].applied;
    }
  }

  /* ---- 5. write back velocities, integrate transforms ---- */
  {
    obody* ob[2] = {&st->cart, &st->pole};
    for (k = 0; k < 2; ++k) {
      sbody* b = &bodies[k + 1];
      obody* o = ob[k];
      real fAngle, axis[3], dq[4], nq[4], l;
      for (i = 0; i < 6; ++i) b->vel[i] = clampr(b->vel[i] + b->dvel[i], -vmax, vmax);
      for (i = 0; i < 3; ++i) { o->w[i] = b->vel[i]; o->v[i] = b->vel[3 + i]; }
      /* stepPositionsMultiDof */
      for (i = 0; i < 3; ++i) o->x[i] += dt * o->v[i];
      fAngle = R_SQRT(dot3(o->w, o->w));
      if (fAngle * dt > R(0.5 * (PI_D * 0.5))) fAngle = R(0.5) * R(PI_D * 0.5) / dt;
      if (fAngle < R(0.001)) {
        const real f = R(0.5) * dt - (dt * dt * dt) * R(0.020833333333) * fAngle * fAngle;
        for (i = 0; i < 3; ++i) axis[i] = o->w[i] * f;
      } else {
        const real f = sin_core(R(0.5) * fAngle * dt) / fAngle;
        for (i = 0; i < 3; ++i) axis[i] = o->w[i] * f;
      }
      dq[0] = axis[0]; dq[1] = axis[1]; dq[2] = axis[2];
      dq[3] = cos_core(fAngle * dt * R(0.5));
      /* pose quaternion q' = dq (x) q.  Bullet updates the inverse rotation,
       * m_baseQuat * (-axis, cos); the terms below are summed in that order. */
      nq[0] = ((o->q[3] * dq[0] + o->q[0] * dq[3]) - o->q[1] * dq[2]) + o->q[2] * dq[1];
      nq[1] = ((o->q[3] * dq[1] + o->q[1] * dq[3]) - o->q[2] * dq[0]) + o->q[0] * dq[2];
      nq[2] = ((o->q[3] * dq[2] + o->q[2] * dq[3]) - o->q[0] * dq[1]) + o->q[1] * dq[0];
      nq[3] = ((o->q[3] * dq[3] - o->q[0] * dq[0]) - o->q[1] * dq[1]) - o->q[2] * dq[2];
      l = R(1.0) / R_SQRT(nq[0] * nq[0] + nq[1] * nq[1] + nq[2] * nq[2] + nq[3] * nq[3]);
      for (i = 0; i < 4; ++i) o->q[i] = nq[i] * l;
    }
  }
  /* ---- 6. clearForcesAndTorques ---- */
  for (i = 0; i < 3; ++i) { st->pend_f[i] = 0; st->pend_t[i] = 0; }
}

/* p.applyExternalForce(cart, -1, (fx,fy,0), (0,0,0), frame) — bullet_cartpole.py:202-203,349-351 */
static void apply_cart_force(const bcp_config* cfg, ostate* st, real fx, real fy) {
  real F[3], rel[3], T[3], f_in[3];
  int i;
  f_in[0] = fx; f_in[1] = fy; f_in[2] = 0;
  if (cfg->force_frame == BCP_FRAME_LINK) {
    real Rm[9], z[3] = {0, 0, 0};
    quat_to_mat(st->cart.q, Rm);
    mat_mul_vec(Rm, f_in, F);
    mat_mul_vec(Rm, z, rel);
  } else {
    for (i = 0; i < 3; ++i) { F[i] = f_in[i]; rel[i] = R(0.0) - st->cart.x[i]; }
  }
  cross3(rel, F, T);
  for (i = 0; i < 3; ++i) { st->pend_f[i] += F[i]; st->pend_t[i] += T[i]; }
}

/* state_fields_of_pose_of x2 -> state[r] — bullet_cartpole.py:43-45,307-311 */
static void write_obs(const ostate* st, float* obs_r) {
  int i;
  for (i = 0; i < 3; ++i) obs_r[i] = (float)st->cart.x[i];
  for (i = 0; i < 4; ++i) obs_r[3 + i] = (float)st->cart.q[i];
  for (i = 0; i < 3; ++i) obs_r[7 + i] = (float)st->pole.x[i];
  for (i = 0; i < 4; ++i) obs_r[10 + i] = (float)st->pole.q[i];
}

/* BulletCartpole.reset — bullet_cartpole.py:313-346 */
static void env_reset(const bcp_config* cfg, ostate* st, int64_t global_env, float* obs) {
  int i, k;
  real theta_s = 0, theta_c = 1;
  uint32_t rnd[4];
  st->steps = 0;                                         /* :315 */
  st->done = 0;                                          /* :316 */
  st->reset_left = 0;
  for (i = 0; i < 3; ++i) {                              /* :319-320 */
    st->cart.x[i] = (real)cfg->cart.init_pos[i];
    st->pole.x[i] = (real)cfg->pole.init_pos[i];
    st->cart.v[i] = st->cart.w[i] = st->pole.v[i] = st->pole.w[i] = 0;
    st->cart.q[i] = st->pole.q[i] = 0;
  }
  st->cart.q[3] = st->pole.q[3] = 1;
  if (!cfg->carry_over_reset) {
    for (i = 0; i < 3; ++i) { st->pend_f[i] = st->pend_t[i] = 0; st->mani[i].count = 0; }
  }
  for (k = 0; k < cfg->settle_steps; ++k) substep(cfg, st, 0); /* :326-327 */
  if (cfg->theta_mode == BCP_THETA_PER_RESET && cfg->random_theta) {
    env_rng(cfg, global_env, (uint32_t)st->episode, RNG_PUSH, 0, rnd);
    sincos_2pi(u24(rnd[0]), &theta_s, &theta_c);
  }
  for (k = 0; k < cfg->initial_force_steps; ++k) {       /* :329-332 */
    substep(cfg, st, 0);
    if (cfg->theta_mode == BCP_THETA_PER_SUBSTEP && cfg->random_theta) {
      env_rng(cfg, global_env, (uint32_t)st->episode, RNG_PUSH, (uint32_t)k, rnd);
      sincos_2pi(u24(rnd[0]), &theta_s, &theta_c);
    }
    apply_cart_force(cfg, st, (real)cfg->initial_force * theta_c, (real)cfg->initial_force * theta_s); /* :354-359 */
  }
  st->episode += 1;
  if (obs)
    for (k = 0; k < cfg->action_repeats; ++k) write_obs(st, obs + 14 * k); /* :338-339 */
}

/* the done rule of bullet_cartpole.py:243-257 evaluated trig-free */
static int done_reason(const bcp_config* cfg, const ostate* st) {
  int reason = BCP_REASON_NONE;
  if (cfg->termination_enabled) {
    const real x = st->pole.x[0], y = st->pole.x[1];
    const real* q = st->pole.q;
    if (R_FABS(x) > (real)cfg->pos_threshold || R_FABS(y) > (real)cfg->pos_threshold) {
      reason = BCP_REASON_POSITION;
    } else {
      const real sarg = R(-2.0) * (q[0] * q[2] - q[3] * q[1]);                       /* sin(pitch) */
      const real sr = R(2.0) * (q[1] * q[2] + q[3] * q[0]);                           /* roll = atan2(sr, cr) */
      const real cr = q[3] * q[3] - q[0] * q[0] - q[1] * q[1] + q[2] * q[2];
      const int pitch_out = R_FABS(sarg) > (real)cfg->sin_angle_threshold;
      const int roll_out = (cr > R(0.0)) ? (R_FABS(sr) > (real)cfg->tan_angle_threshold * cr)
                                         : (cr < R(0.0) || sr != R(0.0));
      if (roll_out || pitch_out) reason = BCP_REASON_ORIENTATION;
    }
  }
  if (st->steps >= cfg->max_episode_len) reason = BCP_REASON_LENGTH; /* :255-257 */
  return reason;
}

static void action_to_force(const bcp_config* cfg, const void* action, real* fx, real* fy) {
  if (cfg->discrete_actions) {
    const int a = *(const int32_t*)action;
    const real F = (real)cfg->action_force;
    *fx = 0; *fy = 0;
    if (a == 1) *fx = F;
    else if (a == 2) *fx = -F;
    else if (a == 3) *fy = F;
    else if (a == 4) *fy = -F;
  } else {
    /* numpy: float32 array * python float stays float32 (bullet_cartpole.py:201) */
    const float* a = (const float*)action;
    const float af = (float)cfg->action_force;
    *fx = (real)(a[0] * af);
    *fy = (real)(a[1] * af);
  }
}

/* teleport of reset(): p.resetBasePositionAndOrientation x2 — bullet_cartpole.py:315-320 */
static void reset_begin(const bcp_config* cfg, ostate* st) {
  int i;
  st->steps = 0;
  st->done = 0;
  for (i = 0; i < 3; ++i) {
    st->cart.x[i] = (real)cfg->cart.init_pos[i];
    st->pole.x[i] = (real)cfg->pole.init_pos[i];
    st->cart.v[i] = st->cart.w[i] = st->pole.v[i] = st->pole.w[i] = 0;
    st->cart.q[i] = st->pole.q[i] = 0;
  }
  st->cart.q[3] = st->pole.q[3] = 1;
  if (!cfg->carry_over_reset) {
    for (i = 0; i < 3; ++i) { st->pend_f[i] = st->pend_t[i] = 0; st->mani[i].count = 0; }
  }
}

#define BCPO_INFO_WORDS 10 /* per substep: iterations, rows, friction rows, contacts[3], SAT code[3], active-set hash; iterations = -1: no substep ran */
static void put_info(int32_t* o, const osubstep_info* in) {
  if (!in) { int i; o[0] = -1; for (i = 1; i < BCPO_INFO_WORDS; ++i) o[i] = 0; return; }
  o[0] = in->iterations; o[1] = in->rows; o[2] = in->friction_rows;
  o[3] = in->contacts[0]; o[4] = in->contacts[1]; o[5] = in->contacts[2];
  o[6] = in->code[0]; o[7] = in->code[1]; o[8] = in->code[2];
  o[9] = in->active_hash;
}

/* BulletCartpole.step — bullet_cartpole.py:178-275.
 * With cfg->auto_reset == BCP_AUTO_RESET_PIPELINED (no reference counterpart; include/bcp.h) a finished env
 * runs its reset (:313-346) inside the following calls, R*S substeps per call: the SAME statements as
 * env_reset above, cut at call boundaries (tests/test_oracle_pipelined.py holds the two forms together).
 * info: optional [R*S][BCPO_INFO_WORDS] instrumentation of every substep of the call. */
static void env_step(const bcp_config* cfg, ostate* st, int64_t global_env, const void* action, float* obs, float* reward,
                     uint8_t* done, uint8_t* reason_out, int32_t* info, const int32_t* force_it /* [R*S] or NULL */) {
  enum { IDLE, LIVE, RESETTING, FRESH } mode;
  const int total = cfg->settle_steps + cfg->initial_force_steps;
  int r, s, reason, k = 0;
  real fx = 0, fy = 0;
  if (!st->done) mode = LIVE;
  else if (cfg->auto_reset != BCP_AUTO_RESET_PIPELINED) mode = IDLE;
  else {
    mode = RESETTING;
    if (st->reset_left == 0) { /* the episode ended in an earlier call: the reset starts with the teleport */
      reset_begin(cfg, st);
      st->done = 1;
      st->reset_left = total;
      if (total == 0) { st->done = 0; st->episode += 1; mode = FRESH; }
    }
  }
  if (mode == IDLE) { /* :179-181: obs unchanged, reward 0, done */
    *reward = 0.0f;
    *done = 1;
    if (reason_out) *reason_out = BCP_REASON_NONE;
    if (info) for (k = 0; k < cfg->action_repeats * cfg->steps_per_repeat; ++k) put_info(info + k * BCPO_INFO_WORDS, 0);
    return;
  }
  if (mode == LIVE) action_to_force(cfg, action, &fx, &fy);
  for (r = 0; r < cfg->action_repeats; ++r) {          /* :197 */
    for (s = 0; s < cfg->steps_per_repeat; ++s, ++k) { /* :198 */
      osubstep_info inf;
      const int on = mode == LIVE || (mode == RESETTING && st->reset_left > 0);
      inf.force_it = force_it ? force_it[k] : 0;
      if (on) substep(cfg, st, (info || force_it) ? &inf : 0); /* :199 / :327 / :330 */
      if (info) put_info(info + k * BCPO_INFO_WORDS, on ? &inf : 0);
      if (mode == LIVE) {
        apply_cart_force(cfg, st, fx, fy);             /* :201-203 */
      } else if (mode == RESETTING && st->reset_left > 0) {
        const int push = total - st->reset_left - cfg->settle_steps; /* index of the push this substep belongs to, < 0 while settling */
        if (push >= 0) {                               /* :329-332 */
          real theta_s = 0, theta_c = 1;
          if (cfg->random_theta) {
            uint32_t rnd[4];
            env_rng(cfg, global_env, (uint32_t)st->episode, RNG_PUSH,
                    cfg->theta_mode == BCP_THETA_PER_SUBSTEP ? (uint32_t)push : 0u, rnd);
            sincos_2pi(u24(rnd[0]), &theta_s, &theta_c);
          }
          apply_cart_force(cfg, st, (real)cfg->initial_force * theta_c, (real)cfg->initial_force * theta_s);
        }
        if (--st->reset_left == 0) { st->done = 0; st->episode += 1; mode = FRESH; }
      }
    }
    if (mode == LIVE && obs) write_obs(st, obs + 14 * r); /* :237 */
  }
  if (mode == LIVE) {
    st->steps += 1;                                    /* :239 */
    reason = done_reason(cfg, st);
    if (reason != BCP_REASON_NONE) st->done = 1;
    *reward = 1.0f;                                    /* :260 */
  } else {
    *reward = 0.0f;
    if (mode == FRESH) {
      if (obs) for (r = 0; r < cfg->action_repeats; ++r) write_obs(st, obs + 14 * r); /* :338-339 */
      reason = BCP_REASON_FRESH;
    } else {
      reason = BCP_REASON_RESETTING;
    }
  }
  *done = (uint8_t)st->done;
  if (reason_out) *reason_out = (uint8_t)reason;
}

static void draw_action(const bcp_config* cfg, int64_t global_env, const ostate* st, uint32_t mask, int32_t* ai,
                        float af[2]) {
  uint32_t rnd[4];
  env_rng(cfg, global_env, (uint32_t)st->episode, RNG_ACTION, (uint32_t)st->steps, rnd);
  if (cfg->discrete_actions) {
    int n = 0, i, pick, seen = 0;
    for (i = 0; i < 5; ++i) n += (mask >> i) & 1u;
    pick = (int)(((uint64_t)rnd[0] * (uint64_t)n) >> 32);
    *ai = 0;
    for (i = 0; i < 5; ++i)
      if ((mask >> i) & 1u) {
        if (seen == pick) { *ai = i; break; }
        seen++;
      }
  } else {
    af[0] = 2.0f * ((float)(rnd[0] >> 8) * 5.9604644775390625e-08f) - 1.0f;
    af[1] = 2.0f * ((float)(rnd[1] >> 8) * 5.9604644775390625e-08f) - 1.0f;
  }
}

/* ======================================================================== */
/* exported entry points (ctypes; numpy buffers)                              */
/* ======================================================================== */
#if defined(_OPENMP)
#include <omp.h>
#endif
/* threads of the parallel loops below; <= 0: leave the OpenMP default (launchers such as
 * torchrun export OMP_NUM_THREADS=1, which would silently serialise the CPU baseline) */
int SFX(bcpo_set_threads)(int n) {
#if defined(_OPENMP)
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}

int SFX(bcpo_substep)(const bcp_config* cfg, double* state, int n_envs, int32_t* info_out /*[n][6] or NULL*/) {
  int e;
#pragma omp parallel for schedule(static)
  for (e = 0; e < n_envs; ++e) {
    ostate st;
    osubstep_info info;
    info.force_it = 0;
    load_state(state + (size_t)e * BCP_STATE_DOUBLES, &st);
    substep(cfg, &st, &info);
    store_state(&st, state + (size_t)e * BCP_STATE_DOUBLES);
    if (info_out) {
      int32_t* o = info_out + (size_t)e * 6;
      o[0] = info.iterations; o[1] = info.rows; o[2] = info.friction_rows;
      o[3] = info.contacts[0]; o[4] = info.contacts[1]; o[5] = info.contacts[2];
    }
  }
  return 0;
}

int SFX(bcpo_apply_force)(const bcp_config* cfg, double* state, int n_envs, const double* fxy) {
  int e;
  for (e = 0; e < n_envs; ++e) {
    ostate st;
    load_state(state + (size_t)e * BCP_STATE_DOUBLES, &st);
    apply_cart_force(cfg, &st, (real)fxy[2 * e], (real)fxy[2 * e + 1]);
    store_state(&st, state + (size_t)e * BCP_STATE_DOUBLES);
  }
  return 0;
}

int SFX(bcpo_reset)(const bcp_config* cfg, double* state, int n_envs, const uint8_t* mask, float* obs) {
  int e;
  const int R_ = cfg->action_repeats;
#pragma omp parallel for schedule(dynamic, 16)
  for (e = 0; e < n_envs; ++e) {
    ostate st;
    if (mask && !mask[e]) continue;
    load_state(state + (size_t)e * BCP_STATE_DOUBLES, &st);
    env_reset(cfg, &st, cfg->env_id_offset + e, obs ? obs + (size_t)e * 14 * R_ : 0);
    store_state(&st, state + (size_t)e * BCP_STATE_DOUBLES);
  }
  return 0;
}

static int step_batch(const bcp_config* cfg, double* state, int n_envs, const void* actions, float* obs, float* reward,
                      uint8_t* done, uint8_t* reason, int32_t* info, const int32_t* force_it) {
  int e;
  const int R_ = cfg->action_repeats, RS_ = cfg->action_repeats * cfg->steps_per_repeat;
#pragma omp parallel for schedule(dynamic, 64)
  for (e = 0; e < n_envs; ++e) {
    ostate st;
    const void* a = cfg->discrete_actions ? (const void*)((const int32_t*)actions + e)
                                          : (const void*)((const float*)actions + 2 * (size_t)e);
    load_state(state + (size_t)e * BCP_STATE_DOUBLES, &st);
    env_step(cfg, &st, cfg->env_id_offset + e, a, obs ? obs + (size_t)e * 14 * R_ : 0, &reward[e], &done[e],
             reason ? &reason[e] : 0, info ? info + (size_t)e * RS_ * BCPO_INFO_WORDS : 0, force_it ? force_it + (size_t)e * RS_ : 0);
    store_state(&st, state + (size_t)e * BCP_STATE_DOUBLES);
  }
  return 0;
}
int SFX(bcpo_step)(const bcp_config* cfg, double* state, int n_envs, const void* actions, float* obs, float* reward,
                   uint8_t* done, uint8_t* reason) {
  return step_batch(cfg, state, n_envs, actions, obs, reward, done, reason, 0, 0);
}
/* the same with info[n][R*S][10] filled (BCPO_INFO_WORDS) and, optionally, the PGS iteration count of every substep
 * imposed (force_it[n][R*S], entries <= 0: the residual rule decides) */
int SFX(bcpo_step_info)(const bcp_config* cfg, double* state, int n_envs, const void* actions, float* obs, float* reward,
                        uint8_t* done, uint8_t* reason, int32_t* info, const int32_t* force_it) {
  return step_batch(cfg, state, n_envs, actions, obs, reward, done, reason, info, force_it);
}

/* README.md:76-93 random-agent protocol; hist has max_episode_len+1 bins */
int SFX(bcpo_rollout_random)(const bcp_config* cfg, double* state, int n_envs, int n_steps, uint32_t action_mask,
                             int max_episodes, float* obs, int64_t* hist, int64_t* header, double* reward_sum) {
  int e;
  const int R_ = cfg->action_repeats;
  const int nb = cfg->max_episode_len + 1;
  int64_t episodes = 0, env_steps = 0, resets = 0;
  double rsum = 0;
#pragma omp parallel
  {
    int64_t* lh = (int64_t*)calloc((size_t)nb, sizeof(int64_t));
    int64_t lep = 0, lst = 0, lrs = 0;
    double lr = 0;
    int b;
#pragma omp for schedule(dynamic, 8)
    for (e = 0; e < n_envs; ++e) {
      ostate st;
      int t;
      float* o = obs ? obs + (size_t)e * 14 * R_ : 0;
      load_state(state + (size_t)e * BCP_STATE_DOUBLES, &st);
      for (t = 0; t < n_steps; ++t) {
        int32_t ai = 0;
        float af[2] = {0, 0}, rew;
        uint8_t dn, rs;
        if (st.done) {
          /* st.episode counts finished resets: the env has delivered its share of episodes */
          if (max_episodes > 0 && st.episode >= max_episodes) break;
          env_reset(cfg, &st, cfg->env_id_offset + e, o);
          lrs++;
        }
        draw_action(cfg, cfg->env_id_offset + e, &st, action_mask, &ai, af);
        env_step(cfg, &st, cfg->env_id_offset + e, cfg->discrete_actions ? (const void*)&ai : (const void*)af, o, &rew, &dn, &rs, 0, 0);
        lst++;
        lr += rew;
        if (dn) {
          lh[st.steps <= cfg->max_episode_len ? st.steps : cfg->max_episode_len]++;
          lep++;
        }
      }
      store_state(&st, state + (size_t)e * BCP_STATE_DOUBLES);
    }
#pragma omp critical
    {
      for (b = 0; b < nb; ++b) hist[b] += lh[b];
      episodes += lep; env_steps += lst; resets += lrs; rsum += lr;
    }
    free(lh);
  }
  header[0] += episodes;
  header[1] += env_steps;
  header[2] += env_steps * (int64_t)(cfg->action_repeats * cfg->steps_per_repeat) +
               resets * (int64_t)(cfg->settle_steps + cfg->initial_force_steps);
  header[3] += resets;
  *reward_sum += rsum;
  return 0;
}

/* bare narrowphase, for unit tests of the box-box detector:
 * in: pA[3], qA[4], hA[3], pB[3], qB[4], hB[3]; out: up to 4 x (n3, pb3, depth) */
int SFX(bcpo_box_box)(const double* in, double* out) {
  real p1[3], q1[4], h1[3], p2[3], q2[4], h2[3], R1[9], R2[9];
  ocontact cs[4];
  int i, k, n;
  for (i = 0; i < 3; ++i) { p1[i] = (real)in[i]; h1[i] = (real)in[7 + i]; p2[i] = (real)in[10 + i]; h2[i] = (real)in[17 + i]; }
  for (i = 0; i < 4; ++i) { q1[i] = (real)in[3 + i]; q2[i] = (real)in[13 + i]; }
  quat_to_mat(q1, R1);
  quat_to_mat(q2, R2);
  n = box_box(p1, R1, h1, p2, R2, h2, cs);
  for (k = 0; k < n; ++k) {
    for (i = 0; i < 3; ++i) { out[k * 7 + i] = cs[k].n[i]; out[k * 7 + 3 + i] = cs[k].pb[i]; }
    out[k * 7 + 6] = cs[k].depth;
  }
  return n;
}

/* elementary functions, exposed so tests can pin them against libm */
void SFX(bcpo_elem)(double x, double y, double* out) {
  real s, c;
  out[0] = sin_core((real)x);
  out[1] = cos_core((real)x);
  out[2] = det_atan2((real)y, (real)x);
  sincos_2pi((real)x, &s, &c);
  out[3] = s;
  out[4] = c;
}
void SFX(bcpo_philox)(uint64_t seed, int64_t env, uint32_t episode, uint32_t purpose, uint32_t index, uint32_t* out) {
  bcp_config cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.seed = seed;
  env_rng(&cfg, env, episode, purpose, index, out);
}
int SFX(bcpo_done_reason)(const bcp_config* cfg, const double* state) {
  ostate st;
  load_state(state, &st);
  return done_reason(cfg, &st);
}
void SFX(bcpo_debug)(int on) { g_debug = on; }
size_t SFX(bcpo_config_size)(void) { return sizeof(bcp_config); }
