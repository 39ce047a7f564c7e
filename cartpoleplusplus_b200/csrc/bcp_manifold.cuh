// bcp_manifold.cuh — btPersistentManifold cache logic (included by bcp_collide.cuh).
// Every access to the four cached points uses a compile-time index (unrolled loops with guards,
// element selection by comparison instead of m.p[i] with a runtime i), so that a Manifold held
// in a local variable stays in registers on the device instead of the thread's stack.
#pragma once

namespace bcp {

// btPersistentManifold::sortCachedPoints with gContactCalcArea3Points = true:
// which of the 4 cached points a 5th one replaces (keep deepest, maximise area).
template <typename T> BCP_HD int sort_cached_points(const Manifold<T>& m, const MPoint<T>& pt) {
  int maxPenetrationIndex = -1;
  T maxPenetration = pt.dist;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (m.p[i].dist < maxPenetration) { maxPenetrationIndex = i; maxPenetration = m.p[i].dist; }
  T res[4] = {0, 0, 0, 0};
  // (a0 index, b1 index, b2 index) per candidate: a = pt - p[ia]; b = p[ib1] - p[ib2]
#define BCP_AREA(c, ia, ib1, ib2)                                                                          \
  if (maxPenetrationIndex != c) {                                                                          \
    T a[3], b[3], cr[3];                                                                                   \
    for (int k = 0; k < 3; ++k) { a[k] = pt.la[k] - m.p[ia].la[k]; b[k] = m.p[ib1].la[k] - m.p[ib2].la[k]; } \
    cross3(a, b, cr);                                                                                      \
    res[c] = dot3(cr, cr);                                                                                 \
  }
  BCP_AREA(0, 1, 3, 2)
  BCP_AREA(1, 0, 3, 2)
  BCP_AREA(2, 0, 3, 1)
  BCP_AREA(3, 0, 2, 1)
#undef BCP_AREA
  // btVector4::closestAxis4 == maxAxis4 of absolute values
  int best = -1;
  T maxVal = -Lim<T>::large();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const T v = babs(res[i]);
    if (v > maxVal) { best = i; if (i < 3) maxVal = v; }
  }
  return best;
}

// btManifoldResult::addContactPoint: cache lookup by local point on A, then
// replace (keeping the cached impulse) / append / evict.
template <typename T>
BCP_HD void manifold_add(Manifold<T>& m, const Posed<T>& A, const Posed<T>& B, const Contact<T>& c, T breaking) {
  if (c.depth > breaking) return;
  MPoint<T> np;
  T pointA[3], v[3];
  for (int i = 0; i < 3; ++i) pointA[i] = c.pb[i] + c.n[i] * c.depth;
  for (int i = 0; i < 3; ++i) v[i] = pointA[i] - A.x[i];
  matT_mul_vec(A.R, v, np.la);
  for (int i = 0; i < 3; ++i) v[i] = c.pb[i] - B.x[i];
  matT_mul_vec(B.R, v, np.lb);
  for (int i = 0; i < 3; ++i) np.n[i] = c.n[i];
  np.dist = c.depth;
  np.imp = 0;
  T shortest = breaking * breaking;
  int nearest = -1;
  T near_imp = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (i < m.count) {
      T d[3];
      d[0] = m.p[i].la[0] - np.la[0]; d[1] = m.p[i].la[1] - np.la[1]; d[2] = m.p[i].la[2] - np.la[2];
      const T dd = dot3(d, d);
      if (dd < shortest) { shortest = dd; nearest = i; near_imp = m.p[i].imp; }
    }
  }
  int insert;
  if (nearest >= 0) {
    np.imp = near_imp;
    insert = nearest;
  } else {
    insert = m.count;
    if (insert == 4) insert = sort_cached_points(m, np);
    else m.count++;
    if (insert < 0) insert = 0;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (i == insert) m.p[i] = np;
}

// btPersistentManifold::refreshContactPoints; also yields the world points the
// solver needs (positionWorldOnA/B).
template <typename T>
BCP_HD void manifold_refresh(Manifold<T>& m, const Posed<T>& A, const Posed<T>& B, T breaking, WorldPt<T> wp[4]) {
#pragma unroll
  for (int i = 3; i >= 0; --i) {
    if (i < m.count) {
      MPoint<T>& p = m.p[i];
      T d[3];
      mat_mul_vec(A.R, p.la, wp[i].pa);
      mat_mul_vec(B.R, p.lb, wp[i].pb);
      for (int k = 0; k < 3; ++k) { wp[i].pa[k] += A.x[k]; wp[i].pb[k] += B.x[k]; }
      for (int k = 0; k < 3; ++k) d[k] = wp[i].pa[k] - wp[i].pb[k];
      p.dist = dot3(d, p.n);
    }
  }
#pragma unroll
  for (int i = 3; i >= 0; --i) {
    if (i < m.count) {
      MPoint<T>& p = m.p[i];
      bool remove = false;
      if (!(p.dist <= breaking)) {
        remove = true;
      } else {
        T proj[3], diff[3];
        for (int k = 0; k < 3; ++k) proj[k] = wp[i].pa[k] - p.n[k] * p.dist;
        for (int k = 0; k < 3; ++k) diff[k] = wp[i].pb[k] - proj[k];
        if (dot3(diff, diff) > breaking * breaking) remove = true;
      }
      if (remove) {  // the last point takes the removed one's slot
        const int last = m.count - 1;
#pragma unroll
        for (int l = i + 1; l < 4; ++l)
          if (l == last) { m.p[i] = m.p[l]; wp[i] = wp[l]; }
        m.count--;
      }
    }
  }
}

// The same two operations on a manifold that lives in ADDRESSABLE memory (k_prep keeps the pair it works on in its
// shared-memory frame): plain run-time indices instead of guarded unrolled copies - the arithmetic, its order and every
// decision are those of manifold_add / manifold_refresh above, statement for statement (tests/test_host_kernels.py holds
// the two forms together bit for bit).  What changes is the instruction count: an insert is 11 stores instead of 44
// predicated moves, and the four points no longer occupy 44 registers.
template <typename T>
BCP_HD void manifold_add_mem(Manifold<T>& m, const Posed<T>& A, const Posed<T>& B, const Contact<T>& c, T breaking) {
  if (c.depth > breaking) return;
  MPoint<T> np;
  T pointA[3], v[3];
  for (int i = 0; i < 3; ++i) pointA[i] = c.pb[i] + c.n[i] * c.depth;
  for (int i = 0; i < 3; ++i) v[i] = pointA[i] - A.x[i];
  matT_mul_vec(A.R, v, np.la);
  for (int i = 0; i < 3; ++i) v[i] = c.pb[i] - B.x[i];
  matT_mul_vec(B.R, v, np.lb);
  for (int i = 0; i < 3; ++i) np.n[i] = c.n[i];
  np.dist = c.depth;
  np.imp = 0;
  T shortest = breaking * breaking;
  int nearest = -1;
  const int count = m.count;
  for (int i = 0; i < count; ++i) {
    T d[3];
    d[0] = m.p[i].la[0] - np.la[0]; d[1] = m.p[i].la[1] - np.la[1]; d[2] = m.p[i].la[2] - np.la[2];
    const T dd = dot3(d, d);
    if (dd < shortest) { shortest = dd; nearest = i; }
  }
  int insert;
  if (nearest >= 0) {
    np.imp = m.p[nearest].imp;
    insert = nearest;
  } else {
    insert = count;
    if (insert == 4) insert = sort_cached_points(m, np);
    else m.count = count + 1;
    if (insert < 0) insert = 0;
  }
  m.p[insert] = np;
}
template <typename T>
BCP_HD void manifold_refresh_mem(Manifold<T>& m, const Posed<T>& A, const Posed<T>& B, T breaking, WorldPt<T>* wp) {
  int count = m.count;
  for (int i = count - 1; i >= 0; --i) {
    MPoint<T>& p = m.p[i];
    T d[3];
    mat_mul_vec(A.R, p.la, wp[i].pa);
    mat_mul_vec(B.R, p.lb, wp[i].pb);
    for (int k = 0; k < 3; ++k) { wp[i].pa[k] += A.x[k]; wp[i].pb[k] += B.x[k]; }
    for (int k = 0; k < 3; ++k) d[k] = wp[i].pa[k] - wp[i].pb[k];
    p.dist = dot3(d, p.n);
  }
  for (int i = count - 1; i >= 0; --i) {
    if (i >= count) continue;  // (cannot happen: count only shrinks below the points already visited)
    const MPoint<T>& p = m.p[i];
    bool remove = false;
    if (!(p.dist <= breaking)) {
      remove = true;
    } else {
      T proj[3], diff[3];
      for (int k = 0; k < 3; ++k) proj[k] = wp[i].pa[k] - p.n[k] * p.dist;
      for (int k = 0; k < 3; ++k) diff[k] = wp[i].pb[k] - proj[k];
      if (dot3(diff, diff) > breaking * breaking) remove = true;
    }
    if (remove) {  // the last point takes the removed one's slot
      const int last = count - 1;
      if (last > i) { m.p[i] = m.p[last]; wp[i] = wp[last]; }
      count--;
    }
  }
  m.count = count;
}

}  // namespace bcp
