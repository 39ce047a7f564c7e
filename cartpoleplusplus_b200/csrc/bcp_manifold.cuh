// bcp_manifold.cuh — btPersistentManifold cache logic (included by bcp_collide.cuh).
#pragma once

namespace bcp {

// btPersistentManifold::sortCachedPoints with gContactCalcArea3Points = true:
// which of the 4 cached points a 5th one replaces (keep deepest, maximise area).
template <typename T> BCP_HD int sort_cached_points(const Manifold<T>& m, const MPoint<T>& pt) {
  int maxPenetrationIndex = -1;
  T maxPenetration = pt.dist;
  for (int i = 0; i < 4; ++i)
    if (m.p[i].dist < maxPenetration) { maxPenetrationIndex = i; maxPenetration = m.p[i].dist; }
  T res[4] = {0, 0, 0, 0};
  // (a0 index, b1 index, b2 index) per candidate: a = pt - p[ia]; b = p[ib1] - p[ib2]
  const int ia[4] = {1, 0, 0, 0}, ib1[4] = {3, 3, 3, 2}, ib2[4] = {2, 2, 1, 1};
  for (int c = 0; c < 4; ++c) {
    if (maxPenetrationIndex != c) {
      T a[3], b[3], cr[3];
      for (int k = 0; k < 3; ++k) { a[k] = pt.la[k] - m.p[ia[c]].la[k]; b[k] = m.p[ib1[c]].la[k] - m.p[ib2[c]].la[k]; }
      cross3(a, b, cr);
      res[c] = dot3(cr, cr);
    }
  }
  // btVector4::closestAxis4 == maxAxis4 of absolute values
  int best = -1;
  T maxVal = -Lim<T>::large();
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
  for (int i = 0; i < m.count; ++i) {
    T d[3];
    d[0] = m.p[i].la[0] - np.la[0]; d[1] = m.p[i].la[1] - np.la[1]; d[2] = m.p[i].la[2] - np.la[2];
    const T dd = dot3(d, d);
    if (dd < shortest) { shortest = dd; nearest = i; }
  }
  if (nearest >= 0) {
    np.imp = m.p[nearest].imp;
    m.p[nearest] = np;
  } else {
    int insert = m.count;
    if (insert == 4) insert = sort_cached_points(m, np);
    else m.count++;
    if (insert < 0) insert = 0;
    m.p[insert] = np;
  }
}

// btPersistentManifold::refreshContactPoints; also yields the world points the
// solver needs (positionWorldOnA/B).
template <typename T>
BCP_HD void manifold_refresh(Manifold<T>& m, const Posed<T>& A, const Posed<T>& B, T breaking, WorldPt<T> wp[4]) {
  for (int i = m.count - 1; i >= 0; --i) {
    MPoint<T>& p = m.p[i];
    T d[3];
    mat_mul_vec(A.R, p.la, wp[i].pa);
    mat_mul_vec(B.R, p.lb, wp[i].pb);
    for (int k = 0; k < 3; ++k) { wp[i].pa[k] += A.x[k]; wp[i].pb[k] += B.x[k]; }
    for (int k = 0; k < 3; ++k) d[k] = wp[i].pa[k] - wp[i].pb[k];
    p.dist = dot3(d, p.n);
  }
  for (int i = m.count - 1; i >= 0; --i) {
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
    if (remove) {
      const int last = m.count - 1;
      if (i != last) { m.p[i] = m.p[last]; wp[i] = wp[last]; }
      m.count--;
    }
  }
}

}  // namespace bcp
