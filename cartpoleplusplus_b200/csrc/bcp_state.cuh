// bcp_state.cuh — structure-of-arrays env state in HBM and its register image.
//
// Layout: `fdata[f * N + e]` for the NF float fields (f follows the canonical record
// of include/bcp.h: bodies, pending force/torque, 3 x 4 manifold points) and
// `idata[f * N + e]` for the NI integer fields.  Consecutive lanes of a warp own
// consecutive envs, so every field access is one fully coalesced 128/256-byte
// transaction; manifold points beyond the live count are neither read nor written.
#pragma once
#include "bcp_types.cuh"

namespace bcp {

enum { NF = BCP_S_MANI + 3 * BCP_MANI_STRIDE, NI = 8 };  // 164 scalars + cnt[3], steps, done, episode, reset_left, mode
enum { I_CNT = 0, I_STEPS = 3, I_DONE = 4, I_EPISODE = 5, I_RESET_LEFT = 6,
       I_MODE = 7 };  // MODE_* of the call in flight (split pipeline only; not part of the canonical record)

template <typename T> BCP_HD void load_body(const T* __restrict__ f, size_t N, size_t e, int base, Body<T>& b) {
  for (int i = 0; i < 3; ++i) b.x[i] = f[(size_t)(base + i) * N + e];
  for (int i = 0; i < 4; ++i) b.q[i] = f[(size_t)(base + 3 + i) * N + e];
  for (int i = 0; i < 3; ++i) b.v[i] = f[(size_t)(base + 7 + i) * N + e];
  for (int i = 0; i < 3; ++i) b.w[i] = f[(size_t)(base + 10 + i) * N + e];
}
template <typename T> BCP_HD void store_body(T* __restrict__ f, size_t N, size_t e, int base, const Body<T>& b) {
  for (int i = 0; i < 3; ++i) f[(size_t)(base + i) * N + e] = b.x[i];
  for (int i = 0; i < 4; ++i) f[(size_t)(base + 3 + i) * N + e] = b.q[i];
  for (int i = 0; i < 3; ++i) f[(size_t)(base + 7 + i) * N + e] = b.v[i];
  for (int i = 0; i < 3; ++i) f[(size_t)(base + 10 + i) * N + e] = b.w[i];
}

template <typename T>
BCP_HD void load_env(const T* __restrict__ f, const int* __restrict__ id, size_t N, size_t e, Env<T>& st) {
  load_body(f, N, e, BCP_S_CART, st.cart);
  load_body(f, N, e, BCP_S_POLE, st.pole);
  for (int i = 0; i < 3; ++i) { st.pend_f[i] = f[(size_t)(BCP_S_PEND_F + i) * N + e]; st.pend_t[i] = f[(size_t)(BCP_S_PEND_T + i) * N + e]; }
  for (int m = 0; m < 3; ++m) {
    st.mani[m].count = id[(size_t)(I_CNT + m) * N + e];
    for (int k = 0; k < 4; ++k) {
      if (k < st.mani[m].count) {
        const int base = BCP_S_MANI + m * BCP_MANI_STRIDE + k * BCP_MANI_POINT;
        MPoint<T>& p = st.mani[m].p[k];
        for (int i = 0; i < 3; ++i) { p.la[i] = f[(size_t)(base + i) * N + e]; p.lb[i] = f[(size_t)(base + 3 + i) * N + e]; p.n[i] = f[(size_t)(base + 6 + i) * N + e]; }
        p.dist = f[(size_t)(base + 9) * N + e];
        p.imp = f[(size_t)(base + 10) * N + e];
      }
    }
  }
  st.steps = id[(size_t)I_STEPS * N + e];
  st.done = id[(size_t)I_DONE * N + e];
  st.episode = id[(size_t)I_EPISODE * N + e];
  st.reset_left = id[(size_t)I_RESET_LEFT * N + e];
}
template <typename T>
BCP_HD void store_env(T* __restrict__ f, int* __restrict__ id, size_t N, size_t e, const Env<T>& st) {
  store_body(f, N, e, BCP_S_CART, st.cart);
  store_body(f, N, e, BCP_S_POLE, st.pole);
  for (int i = 0; i < 3; ++i) { f[(size_t)(BCP_S_PEND_F + i) * N + e] = st.pend_f[i]; f[(size_t)(BCP_S_PEND_T + i) * N + e] = st.pend_t[i]; }
  for (int m = 0; m < 3; ++m) {
    id[(size_t)(I_CNT + m) * N + e] = st.mani[m].count;
    for (int k = 0; k < 4; ++k) {
      if (k < st.mani[m].count) {
        const int base = BCP_S_MANI + m * BCP_MANI_STRIDE + k * BCP_MANI_POINT;
        const MPoint<T>& p = st.mani[m].p[k];
        for (int i = 0; i < 3; ++i) { f[(size_t)(base + i) * N + e] = p.la[i]; f[(size_t)(base + 3 + i) * N + e] = p.lb[i]; f[(size_t)(base + 6 + i) * N + e] = p.n[i]; }
        f[(size_t)(base + 9) * N + e] = p.dist;
        f[(size_t)(base + 10) * N + e] = p.imp;
      }
    }
  }
  id[(size_t)I_STEPS * N + e] = st.steps;
  id[(size_t)I_DONE * N + e] = st.done;
  id[(size_t)I_EPISODE * N + e] = st.episode;
  id[(size_t)I_RESET_LEFT * N + e] = st.reset_left;
}

// canonical record (doubles, include/bcp.h) <-> register image
template <typename T> BCP_HD void env_from_canonical(const double* s, Env<T>& st) {
  T tmp[NF];
  for (int i = 0; i < NF; ++i) tmp[i] = (T)s[i];
  int cnt[3];
  for (int m = 0; m < 3; ++m) cnt[m] = (int)s[BCP_S_MANI_CNT + m];
  const int ids[NI] = {cnt[0], cnt[1], cnt[2], (int)s[BCP_S_STEPS], (int)s[BCP_S_DONE], (int)s[BCP_S_EPISODE], (int)s[BCP_S_RESET_LEFT], 0};
  load_env<T>(tmp, ids, 1, 0, st);
}
template <typename T> BCP_HD void env_to_canonical(const Env<T>& st, double* s) {
  T tmp[NF];
  int ids[NI];
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
