// bcp_collide.cuh — narrowphase + persistent manifold for box pairs.
//
// Device-side statement of btBoxBoxDetector (ODE's dBoxBox2: 15-axis SAT,
// incident-face clipping, cull to 4) and of btPersistentManifold's cache logic
// (getCacheEntry / replace / sortCachedPoints / refreshContactPoints), the two
// Bullet components `p.stepSimulation()` (bullet_cartpole.py:199) runs for the
// pole-on-cart, cart-on-ground and pole-on-ground pairs.  [BULLET-RECALL]:
// bullet3 src/BulletCollision/CollisionDispatch/btBoxBoxDetector.cpp,
// NarrowPhaseCollision/btPersistentManifold.cpp, btManifoldResult.cpp.
#pragma once
#include "bcp_types.cuh"

namespace bcp {

template <typename T> struct Contact {
  T n[3];   // normalOnBInWorld
  T pb[3];  // point on B (world)
  T depth;  // signed distance, negative = penetration
};
template <typename T> struct Posed {
  T x[3];
  T R[9];
  T h[3];
};
template <typename T> struct WorldPt {
  T pa[3], pb[3];
};

#define BCP_COL(M, j, i) ((M)[(i)*3 + (j)])
template <typename T> BCP_HD T dot_cc(const T* A, int ja, const T* B, int jb) {
  return BCP_COL(A, ja, 0) * BCP_COL(B, jb, 0) + BCP_COL(A, ja, 1) * BCP_COL(B, jb, 1) + BCP_COL(A, ja, 2) * BCP_COL(B, jb, 2);
}
template <typename T> BCP_HD T dot_cv(const T* A, int ja, const T* v) {
  return BCP_COL(A, ja, 0) * v[0] + BCP_COL(A, ja, 1) * v[1] + BCP_COL(A, ja, 2) * v[2];
}
template <typename T> BCP_HD T dot_vc(const T* v, const T* A, int ja) {
  return v[0] * BCP_COL(A, ja, 0) + v[1] * BCP_COL(A, ja, 1) + v[2] * BCP_COL(A, ja, 2);
}

// intersectRectQuad2: clip the quad p[8] against the rectangle +-h; <= 8 points.
// `buffer` is the second ping-pong array (caller's workspace, see box_box).
template <typename T> BCP_HD_NOINLINE int intersect_rect_quad(const T h[2], T p[8], T ret[16], T buffer[16]) {
  int nq = 4, nr = 0;
  T* q = p;
  T* r = ret;
  for (int dir = 0; dir <= 1; ++dir) {
    for (int sign = -1; sign <= 1; sign += 2) {
      T* pq = q;
      T* pr = r;
      nr = 0;
      for (int i = nq; i > 0; --i) {
        if ((T)sign * pq[dir] < h[dir]) {
          pr[0] = pq[0];
          pr[1] = pq[1];
          pr += 2;
          nr++;
          if (nr & 8) { q = r; goto done; }
        }
        T* nextq = (i > 1) ? pq + 2 : q;
        if (((T)sign * pq[dir] < h[dir]) ^ ((T)sign * nextq[dir] < h[dir])) {
          pr[1 - dir] = pq[1 - dir] + (nextq[1 - dir] - pq[1 - dir]) / (nextq[dir] - pq[dir]) * ((T)sign * h[dir] - pq[dir]);
          pr[dir] = (T)sign * h[dir];
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
  if (q != ret)
    for (int i = 0; i < nr * 2; ++i) ret[i] = q[i];
  return nr;
}

// cullPoints2: choose m (<= 8) of the n clipped points, spread in angle, i0 first.
// Returns the chosen indices packed four bits each (j-th pick in bits 4j..4j+3); `A` is an
// 8-word scratch array from the caller's workspace, the "still available" flags are a bit mask.
template <typename T> BCP_HD_NOINLINE unsigned cull_points(int n, const T p[], int m, int i0, T A[8]) {
  T a, cx, cy, q;
  const float M__PI = 3.14159265f;  // Bullet's float literal
  if (n == 1) {
    cx = p[0];
    cy = p[1];
  } else if (n == 2) {
    cx = T(0.5) * (p[0] + p[2]);
    cy = T(0.5) * (p[1] + p[3]);
  } else {
    a = 0; cx = 0; cy = 0;
    for (int i = 0; i < (n - 1); ++i) {
      q = p[i * 2] * p[i * 2 + 3] - p[i * 2 + 2] * p[i * 2 + 1];
      a += q;
      cx += q * (p[i * 2] + p[i * 2 + 2]);
      cy += q * (p[i * 2 + 1] + p[i * 2 + 3]);
    }
    q = p[n * 2 - 2] * p[1] - p[0] * p[n * 2 - 1];
    if (babs(a + q) > Lim<T>::eps()) a = T(1.0) / (T(3.0) * (a + q));
    else a = Lim<T>::large();
    cx = a * (cx + q * (p[n * 2 - 2] + p[0]));
    cy = a * (cy + q * (p[n * 2 - 1] + p[1]));
  }
  for (int i = 0; i < n; ++i) A[i] = det_atan2(p[i * 2 + 1] - cy, p[i * 2] - cx);
  unsigned avail = ((1u << n) - 1u) & ~(1u << i0);
  unsigned picks = (unsigned)i0;
  const T A0 = A[i0];
  for (int j = 1; j < m; ++j) {
    T maxdiff = T(1e9), diff;
    a = (T)j * (T)(2 * M__PI / (float)m) + A0;
    if (a > (T)M__PI) a -= (T)(2 * M__PI);
    int pick = i0;
    for (int i = 0; i < n; ++i) {
      if ((avail >> i) & 1u) {
        diff = babs(A[i] - a);
        if (diff > (T)M__PI) diff = (T)(2 * M__PI) - diff;
        if (diff < maxdiff) {
          maxdiff = diff;
          pick = i;
        }
      }
    }
    avail &= ~(1u << pick);
    picks |= (unsigned)pick << (4 * j);
  }
  return picks;
}

template <typename T>
BCP_HD void line_closest_approach(const T pa[3], const T ua[3], const T pb[3], const T ub[3], T& alpha, T& beta) {
  T p[3];
  p[0] = pb[0] - pa[0];
  p[1] = pb[1] - pa[1];
  p[2] = pb[2] - pa[2];
  const T uaub = dot3(ua, ub);
  const T q1 = dot3(ua, p);
  const T q2 = -dot3(ub, p);
  T d = T(1.0) - uaub * uaub;
  if (d <= T(0.0001f)) {
    alpha = 0;
    beta = 0;
  } else {
    d = T(1.0) / d;
    alpha = (q1 + uaub * q2) * d;
    beta = (uaub * q1 + q2) * d;
  }
}

}  // namespace bcp
#include "bcp_boxbox.cuh"
#include "bcp_manifold.cuh"
