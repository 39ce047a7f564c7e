// bcp_state.cuh — env state in HBM and its register image.
//
// Layout: one RECORD per env, `fdata[e * NF + f]` for the NF float fields (f follows the
// canonical record of include/bcp.h: bodies, pending force/torque, 3 x 4 manifold points)
// and `idata[e * NI + f]` for the NI integer fields.  Records are 16-byte aligned and are
// moved with 16-byte accesses (float4 / double2): a lane reads its env's 128-byte body
// block and only the manifold blocks that hold points.  Record-per-env (rather than
// field-major arrays) is what lets the split pipeline visit envs in ANY order - it regroups
// them every step so that a warp's 32 envs are in the same contact regime - without
// scattering every access over 32 sectors.
#pragma once
#include "bcp_types.cuh"

namespace bcp {

enum { NF = BCP_S_MANI + 3 * BCP_MANI_STRIDE, NI = 8 };  // 164 scalars + cnt[3], steps, done, episode, reset_left, mode
enum { I_CNT = 0, I_STEPS = 3, I_DONE = 4, I_EPISODE = 5, I_RESET_LEFT = 6,
       I_MODE = 7 };  // MODE_* of the call in flight (split pipeline only; not part of the canonical record)

// 16-byte copies between a record and a local array; n is a multiple of 16 / sizeof(T)
template <typename T> BCP_HD void copy16(const T* __restrict__ src, T* __restrict__ dst, int n) {
  for (int i = 0; i < n; ++i) dst[i] = src[i];
}
#if defined(__CUDA_ARCH__)
template <> __device__ __forceinline__ void copy16<float>(const float* __restrict__ src, float* __restrict__ dst, int n) {
#pragma unroll
  for (int i = 0; i < n; i += 4) {
    const float4 v = *reinterpret_cast<const float4*>(src + i);
    dst[i] = v.x; dst[i + 1] = v.y; dst[i + 2] = v.z; dst[i + 3] = v.w;
  }
}
template <> __device__ __forceinline__ void copy16<double>(const double* __restrict__ src, double* __restrict__ dst, int n) {
#pragma unroll
  for (int i = 0; i < n; i += 2) {
    const double2 v = *reinterpret_cast<const double2*>(src + i);
    dst[i] = v.x; dst[i + 1] = v.y;
  }
}
#endif
template <typename T> BCP_HD void store16(T* __restrict__ dst, const T* __restrict__ src, int n) {
  for (int i = 0; i < n; ++i) dst[i] = src[i];
}
#if defined(__CUDA_ARCH__)
template <> __device__ __forceinline__ void store16<float>(float* __restrict__ dst, const float* __restrict__ src, int n) {
#pragma unroll
  for (int i = 0; i < n; i += 4) *reinterpret_cast<float4*>(dst + i) = make_float4(src[i], src[i + 1], src[i + 2], src[i + 3]);
}
template <> __device__ __forceinline__ void store16<double>(double* __restrict__ dst, const double* __restrict__ src, int n) {
#pragma unroll
  for (int i = 0; i < n; i += 2) *reinterpret_cast<double2*>(dst + i) = make_double2(src[i], src[i + 1]);
}
#endif

template <typename T> BCP_HD void unpack_body(const T* b, Body<T>& o) {
  for (int i = 0; i < 3; ++i) { o.x[i] = b[i]; o.v[i] = b[7 + i]; o.w[i] = b[10 + i]; }
  for (int i = 0; i < 4; ++i) o.q[i] = b[3 + i];
}
template <typename T> BCP_HD void pack_body(const Body<T>& o, T* b) {
  for (int i = 0; i < 3; ++i) { b[i] = o.x[i]; b[7 + i] = o.v[i]; b[10 + i] = o.w[i]; }
  for (int i = 0; i < 4; ++i) b[3 + i] = o.q[i];
}

// A record is read and written in pieces: the CORE (bodies, pending force, counters, the three
// point counts) and, separately, the point block of each manifold - the split pipeline streams
// a pair's points in right before it works on the pair and out right after (short reuse
// distance in L1), and leaves the blocks it does not touch where they are.
template <typename T> BCP_HD void load_env_core(const T* f, const int* id, size_t e, Env<T>& st) {
  const T* r = f + e * NF;
  const int* ir = id + e * NI;
  int iv[NI];
#if defined(__CUDA_ARCH__)
  {
    const int4 a = *reinterpret_cast<const int4*>(ir), b = *reinterpret_cast<const int4*>(ir + 4);
    iv[0] = a.x; iv[1] = a.y; iv[2] = a.z; iv[3] = a.w; iv[4] = b.x; iv[5] = b.y; iv[6] = b.z; iv[7] = b.w;
  }
#else
  for (int i = 0; i < NI; ++i) iv[i] = ir[i];
#endif
  T head[BCP_S_MANI];
  copy16(r, head, BCP_S_MANI);
  unpack_body(head + BCP_S_CART, st.cart);
  unpack_body(head + BCP_S_POLE, st.pole);
  for (int i = 0; i < 3; ++i) { st.pend_f[i] = head[BCP_S_PEND_F + i]; st.pend_t[i] = head[BCP_S_PEND_T + i]; }
  for (int m = 0; m < 3; ++m) st.mani[m].count = iv[I_CNT + m];
  st.steps = iv[I_STEPS];
  st.done = iv[I_DONE];
  st.episode = iv[I_EPISODE];
  st.reset_left = iv[I_RESET_LEFT];
}
// points 0..mf.count-1 of pair m (mf.count is set by the caller)
template <typename T> BCP_HD void load_manifold(const T* f, size_t e, int m, Manifold<T>& mf) {
  if (mf.count <= 0) return;
  T blk[BCP_MANI_STRIDE];
  copy16(f + e * NF + BCP_S_MANI + m * BCP_MANI_STRIDE, blk, BCP_MANI_STRIDE);
  for (int k = 0; k < 4; ++k) {
    if (k < mf.count) {
      const T* q = blk + k * BCP_MANI_POINT;
      MPoint<T>& p = mf.p[k];
      for (int i = 0; i < 3; ++i) { p.la[i] = q[i]; p.lb[i] = q[3 + i]; p.n[i] = q[6 + i]; }
      p.dist = q[9];
      p.imp = q[10];
    }
  }
}
template <typename T> BCP_HD void store_manifold(T* f, size_t e, int m, const Manifold<T>& mf) {
  if (mf.count <= 0) return;
  T blk[BCP_MANI_STRIDE];
  for (int k = 0; k < 4; ++k) {
    T* q = blk + k * BCP_MANI_POINT;
    const MPoint<T>& p = mf.p[k];
    const bool live = k < mf.count;
    for (int i = 0; i < 3; ++i) { q[i] = live ? p.la[i] : T(0); q[3 + i] = live ? p.lb[i] : T(0); q[6 + i] = live ? p.n[i] : T(0); }
    q[9] = live ? p.dist : T(0);
    q[10] = live ? p.imp : T(0);
  }
  store16(f + e * NF + BCP_S_MANI + m * BCP_MANI_STRIDE, blk, BCP_MANI_STRIDE);
}
// The same for a manifold that lives in addressable memory (k_prep's shared-memory frame): the four points of Manifold<T> are
// 44 contiguous scalars in the record block's own order (la3, lb3, n3, dist, imp per point), so the block moves as eleven
// 16-byte pieces without a staging array and without per-field unpacking.  Points past `count` carry whatever the block held
// on the way in (never read) and go out as zeros, like store_manifold.
static_assert(sizeof(MPoint<float>) == BCP_MANI_POINT * sizeof(float) && sizeof(MPoint<double>) == BCP_MANI_POINT * sizeof(double),
              "MPoint is the record's point layout");
template <typename T> BCP_HD void load_manifold_mem(const T* f, size_t e, int m, Manifold<T>& mf) {
  if (mf.count <= 0) return;
  const T* src = f + e * NF + BCP_S_MANI + m * BCP_MANI_STRIDE;
  T* dst = &mf.p[0].la[0];
  constexpr int W = 16 / (int)sizeof(T);
  for (int i = 0; i < BCP_MANI_STRIDE; i += W) {
    T v[W];
    copy16(src + i, v, W);
    for (int k = 0; k < W; ++k) dst[i + k] = v[k];
  }
}
template <typename T> BCP_HD void store_manifold_mem(T* f, size_t e, int m, const Manifold<T>& mf) {
  if (mf.count <= 0) return;
  T* dst = f + e * NF + BCP_S_MANI + m * BCP_MANI_STRIDE;
  const T* src = &mf.p[0].la[0];
  const int live = mf.count * BCP_MANI_POINT;
  constexpr int W = 16 / (int)sizeof(T);
  for (int i = 0; i < BCP_MANI_STRIDE; i += W) {
    T v[W];
    for (int k = 0; k < W; ++k) v[k] = (i + k < live) ? src[i + k] : T(0);
    store16(dst + i, v, W);
  }
}
template <typename T> BCP_HD void store_env_core(T* f, int* id, size_t e, const Env<T>& st) {
  T* r = f + e * NF;
  int* ir = id + e * NI;
  T head[BCP_S_MANI];
  pack_body(st.cart, head + BCP_S_CART);
  pack_body(st.pole, head + BCP_S_POLE);
  for (int i = 0; i < 3; ++i) { head[BCP_S_PEND_F + i] = st.pend_f[i]; head[BCP_S_PEND_T + i] = st.pend_t[i]; }
  store16(r, head, BCP_S_MANI);
  // the mode word (I_MODE) belongs to the split pipeline and is written by it
  ir[I_CNT] = st.mani[0].count; ir[I_CNT + 1] = st.mani[1].count; ir[I_CNT + 2] = st.mani[2].count;
  ir[I_STEPS] = st.steps; ir[I_DONE] = st.done; ir[I_EPISODE] = st.episode; ir[I_RESET_LEFT] = st.reset_left;
}

// the whole record.  `N` is unused (kept so that call sites read the same for any layout)
template <typename T>
BCP_HD void load_env(const T* __restrict__ f, const int* __restrict__ id, size_t N, size_t e, Env<T>& st) {
  (void)N;
  load_env_core(f, id, e, st);
  for (int m = 0; m < 3; ++m) load_manifold(f, e, m, st.mani[m]);
}
template <typename T>
BCP_HD void store_env(T* __restrict__ f, int* __restrict__ id, size_t N, size_t e, const Env<T>& st) {
  (void)N;
  store_env_core(f, id, e, st);
  for (int m = 0; m < 3; ++m) store_manifold(f, e, m, st.mani[m]);
}

// canonical record (doubles, include/bcp.h) <-> register image
template <typename T> BCP_HD void env_from_canonical(const double* s, Env<T>& st) {
  alignas(16) T tmp[NF];
  for (int i = 0; i < NF; ++i) tmp[i] = (T)s[i];
  int cnt[3];
  for (int m = 0; m < 3; ++m) cnt[m] = (int)s[BCP_S_MANI_CNT + m];
  alignas(16) const int ids[NI] = {cnt[0], cnt[1], cnt[2], (int)s[BCP_S_STEPS], (int)s[BCP_S_DONE], (int)s[BCP_S_EPISODE], (int)s[BCP_S_RESET_LEFT], 0};
  load_env<T>(tmp, ids, 1, 0, st);
}
template <typename T> BCP_HD void env_to_canonical(const Env<T>& st, double* s) {
  alignas(16) T tmp[NF];
  alignas(16) int ids[NI];
  for (int i = 0; i < NF; ++i) tmp[i] = 0;
  store_env<T>(tmp, ids, 1, 0, st);
  for (int i = 0; i < NF; ++i) s[i] = (double)tmp[i];
  for (int m = 0; m < 3; ++m) s[BCP_S_MANI_CNT + m] = ids[I_CNT + m];
  s[BCP_S_STEPS] = ids[I_STEPS];
  s[BCP_S_DONE] = ids[I_DONE];
  s[BCP_S_EPISODE] = ids[I_EPISODE];
  s[BCP_S_RESET_LEFT] = ids[I_RESET_LEFT];
  for (int i = BCP_S_RESET_LEFT + 1; i < BCP_STATE_DOUBLES; ++i) s[i] = 0.0;
}

}  // namespace bcp
