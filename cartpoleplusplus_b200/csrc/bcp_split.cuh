// bcp_split.cuh — the split pipeline of the fast path: one call of every env as
//   k_prep(setup 0) k_solve k_prep(finish 0, setup 1) k_solve ... k_prep(finish RS-1, close)
//
// Why: the projected Gauss-Seidel solve dominates a substep and its length is per-env data
// (3..50 iterations, the early exit of btMultiBodyConstraintSolver).  With one env per lane
// for the whole substep a warp runs as long as its slowest lane; measured on B200 the sweep
// then averages 4 of 32 lanes busy (profiles/).  Here the solve is cut loose from the lane
// that owns the env:
//   k_prep  one env per lane, SoA state, coalesced: finishes the previous substep
//           (velocities += deltas, integrate, force application / reset progress, pose capture)
//           and sets up the next (forward dynamics, narrowphase, manifold caches, contact rows).
//           Rows go to a per-env SOLVE RECORD in global memory (L2-resident) and the env is
//           queued, predicted-long solves first.
//   k_solve persistent warps; every lane pulls envs off the queue, copies the record's live
//           rows into its shared-memory column and iterates; a lane whose env has converged
//           retires it and pulls the next, so all 32 lanes sweep on every iteration.
// The arithmetic of an env is exactly substep_fast's (same functions, same order).
#pragma once
#include <cstring>
#include "bcp_fast.cuh"
#include "bcp_state.cuh"

namespace bcp {

// ---- solve record (global memory, one per env, 16-byte aligned) ----
// words 0..15: header; then the three pairs' rows, padded to 12 / 16 words per row
enum { REC_C_GC = 0, REC_C_CP = 1, REC_C_GP = 2, REC_ITERS = 3, REC_DVEL = 4, REC_HEADER = 16 };
enum { RW1_PAD = 12, RW2_PAD = 16 };

template <typename T> BCP_HD RowLayout record_layout(const Params<T>& P) {
  RowLayout L;
  int at = REC_HEADER;
  for (int m = 0; m < 3; ++m) {
    L.base[m] = at;
    L.w[m] = (m == BCP_PAIR_CART_POLE) ? (int)RW2_PAD : (int)RW1_PAD;
    at += 4 * P.rows_per_contact[m] * L.w[m];
  }
  return L;
}
template <typename T> BCP_HD int record_words(const Params<T>& P) {
  const RowLayout L = record_layout(P);
  return L.base[2] + 4 * P.rows_per_contact[2] * L.w[2];
}
BCP_HD int as_int(float x) {
#if defined(__CUDA_ARCH__)
  return __float_as_int(x);
#else
  int i; memcpy(&i, &x, 4); return i;
#endif
}
BCP_HD float as_real(int i, float) {
#if defined(__CUDA_ARCH__)
  return __int_as_float(i);
#else
  float x; memcpy(&x, &i, 4); return x;
#endif
}
BCP_HD int as_int(double x) { return (int)x; }
BCP_HD double as_real(int i, double) { return (double)i; }

// setup half of a substep for one env: rows + header into `rec`, the unconstrained
// velocities parked in the state (angular in the BODY frame until split_finish).
template <typename T, bool FRICTION>
BCP_HD void split_setup(const Params<T>& P, Env<T>& st, T* rec, bool on) {
  FBody<T> cart, pole;
  const Col<T> col{rec, 1};
  const RowLayout L = record_layout(P);
  fast_setup<T, FRICTION>(P, st, col, L, cart, pole, on, nullptr);
  if (on) {
    rec[REC_C_GC] = as_real(st.mani[BCP_PAIR_GROUND_CART].count, T(0));
    rec[REC_C_CP] = as_real(st.mani[BCP_PAIR_CART_POLE].count, T(0));
    rec[REC_C_GP] = as_real(st.mani[BCP_PAIR_GROUND_POLE].count, T(0));
    for (int i = 0; i < 3; ++i) {  // warm-start deltas (zero unless warmstart_factor > 0)
      rec[REC_DVEL + i] = cart.dwl[i]; rec[REC_DVEL + 3 + i] = cart.dv[i];
      rec[REC_DVEL + 6 + i] = pole.dwl[i]; rec[REC_DVEL + 9 + i] = pole.dv[i];
      st.cart.w[i] = cart.wl[i]; st.cart.v[i] = cart.v[i];
      st.pole.w[i] = pole.wl[i]; st.pole.v[i] = pole.v[i];
      st.pend_f[i] = 0; st.pend_t[i] = 0;  // clearForcesAndTorques
    }
  }
}

// record -> compact shared-memory column (live rows only); returns the contact counts
template <typename T>
BCP_HD void record_to_column(const Params<T>& P, const T* rec, const Col<T>& col, int cnt[3]) {
  const RowLayout R = record_layout(P), C = compact_layout(P);
  cnt[0] = as_int(rec[REC_C_GC]); cnt[1] = as_int(rec[REC_C_CP]); cnt[2] = as_int(rec[REC_C_GP]);
  for (int m = 0; m < 3; ++m) {
    const int rows = cnt[m] * P.rows_per_contact[m];
    for (int r = 0; r < rows; ++r) {
      const T* src = rec + R.base[m] + r * R.w[m];
      const int dst = C.base[m] + r * C.w[m];
      for (int w = 0; w < C.w[m]; ++w) col[dst + w] = src[w];
    }
  }
}
// solve results -> record: iteration count, velocity deltas, applied normal impulses
template <typename T>
BCP_HD void column_to_record(const Params<T>& P, T* rec, const Col<T>& col, const int cnt[3], int iters,
                             const FBody<T>& cart, const FBody<T>& pole) {
  const RowLayout R = record_layout(P), C = compact_layout(P);
  rec[REC_ITERS] = as_real(iters, T(0));
  for (int i = 0; i < 3; ++i) {
    rec[REC_DVEL + i] = cart.dwl[i]; rec[REC_DVEL + 3 + i] = cart.dv[i];
    rec[REC_DVEL + 6 + i] = pole.dwl[i]; rec[REC_DVEL + 9 + i] = pole.dv[i];
  }
  for (int m = 0; m < 3; ++m) {
    const int ap = (m == BCP_PAIR_CART_POLE) ? (int)R2_APPL : (int)R1_APPL;
    const int rpc = P.rows_per_contact[m];
    for (int j = 0; j < cnt[m]; ++j) rec[R.base[m] + j * rpc * R.w[m] + ap] = col[C.base[m] + j * rpc * C.w[m] + ap];
  }
}
template <typename T> BCP_HD void solver_bodies(const Params<T>& P, const T* rec, FBody<T>& cart, FBody<T>& pole) {
  for (int i = 0; i < 3; ++i) {
    cart.dwl[i] = rec[REC_DVEL + i]; cart.dv[i] = rec[REC_DVEL + 3 + i];
    pole.dwl[i] = rec[REC_DVEL + 6 + i]; pole.dv[i] = rec[REC_DVEL + 9 + i];
    cart.invI[i] = T(1.0) / P.cart.I[i];
    pole.invI[i] = T(1.0) / P.pole.I[i];
  }
  cart.invM = P.cart.inv_mass;
  pole.invM = P.pole.inv_mass;
}

// finish half of a substep: velocities += deltas, integrate, impulses -> manifold cache
template <typename T> BCP_HD void split_finish(const Params<T>& P, Env<T>& st, const T* rec) {
  FBody<T> cart, pole;
  solver_bodies(P, rec, cart, pole);
  quat_to_mat(st.cart.q, cart.R);
  quat_to_mat(st.pole.q, pole.R);
  for (int i = 0; i < 3; ++i) {
    cart.wl[i] = st.cart.w[i]; cart.v[i] = st.cart.v[i];
    pole.wl[i] = st.pole.w[i]; pole.v[i] = st.pole.v[i];
  }
  const RowLayout R = record_layout(P);
  for (int m = 0; m < 3; ++m) {
    const int ap = (m == BCP_PAIR_CART_POLE) ? (int)R2_APPL : (int)R1_APPL;
    for (int j = 0; j < st.mani[m].count; ++j) st.mani[m].p[j].imp = rec[R.base[m] + j * P.rows_per_contact[m] * R.w[m] + ap];
  }
  fbody_finish(P, cart, st.cart);
  fbody_finish(P, pole, st.pole);
}

// host-side composition (tests): the whole substep through a record, one env
template <typename T, bool FRICTION>
BCP_HD void substep_split_reference(const Params<T>& P, Env<T>& st, T* rec, const Col<T>& col, SubstepInfo* info) {
  split_setup<T, FRICTION>(P, st, rec, true);
  int cnt[3];
  record_to_column(P, rec, col, cnt);
  FBody<T> cart, pole;
  solver_bodies(P, rec, cart, pole);
  int it = 0;
  for (int k = 0; k < P.solver_iterations; ++k) {
    const T lsr = fast_iteration<T, FRICTION>(P, col, cnt[0], cnt[1], cnt[2], cart, pole);
    it = k + 1;
    if (lsr <= P.residual_thr || k >= P.solver_iterations - 1) break;
  }
  column_to_record(P, rec, col, cnt, it, cart, pole);
  split_finish(P, st, rec);
  if (info) {
    info->iterations = it;
    info->rows = cnt[0] + cnt[1] + cnt[2];
    info->friction_rows = 0;
    for (int m = 0; m < 3; ++m) { info->contacts[m] = cnt[m]; if (FRICTION) info->friction_rows += cnt[m] * (P.rows_per_contact[m] - 1); }
  }
}

#if defined(__CUDACC__)
// ---- queue: long-predicted solves from the front, the rest from the back ----
enum { Q_LONG = 0, Q_SHORT = 1, Q_FETCH = 2, Q_WORDS = 4 };
constexpr int kLongIters = 16;   // a previous solve this long queues the env in the long bin
constexpr int kFetchChunk = 32;  // queue entries a warp takes per atomic

struct SplitIO {
  const void* actions;
  float* obs;
  float* reward;
  uint8_t* done;
  uint8_t* reason;
};

// s_fin: substep to finish (-1: none, this is the first kernel of the call)
// s_set: substep to set up  (-1: none, this is the last kernel of the call)
template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock, 4) k_prep(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id,
                                                 T* __restrict__ recs, int rec_words, int* __restrict__ queue,
                                                 int* __restrict__ qc_set, SplitIO io, Stats stats, int s_fin, int s_set) {
  const int N = P.num_envs;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = e < N;
  const int64_t genv = P.env_id_offset + e;
  const int width = 14 * P.R;
  T* rec = recs + (size_t)(valid ? e : 0) * rec_words;
  float* obs = (io.obs && valid) ? io.obs + (size_t)e * width : nullptr;
  Env<T> st;
  st.done = 1;
  st.reset_left = 0;
  int mode = MODE_IDLE;
  if (valid) {
    load_env<T>(f, id, N, e, st);
    mode = id[(size_t)I_MODE * N + e];
  }
  if (s_fin < 0) mode = call_begin(P, st, valid);
  if (s_fin >= 0 && valid) {
    if (substep_on(mode, st)) split_finish(P, st, rec);
    int a = 0;
    float ax = 0.f, ay = 0.f;
    if (mode == MODE_LIVE) {
      if (P.discrete) a = reinterpret_cast<const int32_t*>(io.actions)[e];
      else {
        const float2 av = reinterpret_cast<const float2*>(io.actions)[e];
        ax = av.x; ay = av.y;
      }
    }
    T fx, fy;
    action_force(P, a, ax, ay, fx, fy);
    post_substep(P, st, genv, mode, fx, fy);
    if ((s_fin + 1) % P.S == 0 && mode == MODE_LIVE && obs) write_obs(st, obs + 14 * (s_fin / P.S));  // :237
  }
  if (s_set >= 0) {
    const bool on = valid && substep_on(mode, st);
    const int last_iters = on ? as_int(rec[REC_ITERS]) : 0;
    split_setup<T, FRICTION>(P, st, rec, on);
    // queue the env: one atomic per warp and bin
    const bool is_long = on && last_iters >= kLongIters;
    const unsigned ml = __ballot_sync(0xffffffffu, is_long), ms = __ballot_sync(0xffffffffu, on && !is_long);
    const int lane = threadIdx.x & 31;
    int bl = 0, bs = 0;
    if (lane == 0) {
      if (ml) bl = atomicAdd(&qc_set[Q_LONG], __popc(ml));
      if (ms) bs = atomicAdd(&qc_set[Q_SHORT], __popc(ms));
    }
    bl = __shfl_sync(0xffffffffu, bl, 0);
    bs = __shfl_sync(0xffffffffu, bs, 0);
    const unsigned lt = (1u << lane) - 1;
    if (is_long) queue[bl + __popc(ml & lt)] = e;
    else if (on) queue[N - 1 - (bs + __popc(ms & lt))] = e;
  } else {  // the call closes: steps / done / reward (:239-260), outputs, episode statistics
    bool live = false;
    const int rs = call_end(P, st, mode, obs, live);
    if (valid) {
      if (io.reward) io.reward[e] = live ? 1.0f : 0.0f;
      if (io.done) io.done[e] = (uint8_t)st.done;
      if (io.reason) io.reason[e] = (uint8_t)rs;
      if (live && st.done && stats.hist) {
        atomicAdd(&stats.hist[st.steps <= P.max_len ? st.steps : P.max_len], 1ull);
        atomicAdd(&stats.header[0], 1ull);
      }
    }
    if (stats.header) {
      const unsigned w = __popc(__ballot_sync(0xffffffffu, live));
      if ((threadIdx.x & 31) == 0 && w) atomicAdd(&stats.header[1], (unsigned long long)w);
    }
  }
  if (valid) {
    store_env<T>(f, id, N, e, st);
    id[(size_t)I_MODE * N + e] = mode;
  }
}

// persistent pooled solve: every lane keeps pulling envs until the queue is dry.
// Shared memory: one column per lane, lane-major with an odd stride ([lane][col_stride]), so
// that both access patterns are conflict-free - 32 lanes reading the same word of their own
// columns (the sweep) and 32 lanes writing consecutive words of ONE column (the cooperative
// copy of a record's rows when a lane takes a new env).
constexpr int kFetchMin = 6;  // idle lanes that trigger a retire + fetch round

template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock) k_solve(const __grid_constant__ Params<T> P, T* __restrict__ recs, int rec_words,
                                                  const int* __restrict__ queue, int* __restrict__ qc, int col_stride) {
  const unsigned FULL = 0xffffffffu;
  T* const smem = reinterpret_cast<T*>(bcp_smem);
  const int lane = threadIdx.x & 31, warp_base = threadIdx.x - lane;
  Col<T> col;
  col.base = smem + (size_t)threadIdx.x * col_stride;
  col.stride = 1;
  const RowLayout RL = record_layout(P), CL = compact_layout(P);
  const int N = P.num_envs;
  const int n_long = qc[Q_LONG], total = n_long + qc[Q_SHORT];
  const unsigned lt = (1u << lane) - 1;
  int w_next = 0, w_end = 0;  // the warp's slice of the queue (uniform)
  bool exhausted = total == 0;
  FBody<T> cart, pole;
  int cnt[3] = {0, 0, 0};
  int it = 0;
  T* rec = nullptr;
  bool busy = false;
  while (true) {
    const unsigned idle = __ballot_sync(FULL, !busy);
    const bool any_busy = idle != FULL;
    if (idle && (!any_busy || (!exhausted && __popc(idle) >= kFetchMin))) {
      if (!busy && rec) {  // retire: iteration count, velocity deltas, applied normal impulses
        column_to_record(P, rec, col, cnt, it, cart, pole);
        rec = nullptr;
        cnt[0] = cnt[1] = cnt[2] = 0;
      }
      unsigned need = idle;
      int idx = -1;
      while (need && !exhausted) {
        const int avail = w_end - w_next;
        const int rank = __popc(need & lt);
        const bool mine = ((need >> lane) & 1u) && rank < avail;
        if (mine) idx = w_next + rank;
        const unsigned got = __ballot_sync(FULL, mine);
        need &= ~got;
        w_next += __popc(got);
        if (!need) break;
        int b = 0;
        if (lane == 0) b = atomicAdd(&qc[Q_FETCH], kFetchChunk);
        b = __shfl_sync(FULL, b, 0);
        if (b >= total) exhausted = true;
        else { w_next = b; w_end = min(b + kFetchChunk, total); }
      }
      int env = -1;
      if (idx >= 0) {
        const int pos = idx < n_long ? idx : N - 1 - (idx - n_long);
        env = queue[pos];
        rec = recs + (size_t)env * rec_words;
        cnt[0] = as_int(rec[REC_C_GC]); cnt[1] = as_int(rec[REC_C_CP]); cnt[2] = as_int(rec[REC_C_GP]);
        solver_bodies(P, rec, cart, pole);
        it = 0;
        busy = true;
      }
      // the warp copies the live rows of every newly taken env into its lane's column
      unsigned taken = __ballot_sync(FULL, idx >= 0);
      while (taken) {
        const int src = __ffs(taken) - 1;
        taken &= taken - 1;
        const int e_src = __shfl_sync(FULL, env, src);
        const T* r_src = recs + (size_t)e_src * rec_words;
        T* c_dst = smem + (size_t)(warp_base + src) * col_stride;
#pragma unroll
        for (int m = 0; m < 3; ++m) {
          const int wc = CL.w[m], wp = RL.w[m];
          const int nw = __shfl_sync(FULL, cnt[m], src) * P.rows_per_contact[m] * wc;
          for (int i = lane; i < nw; i += 32) {
            const int r = i / wc, w = i - r * wc;
            c_dst[CL.base[m] + i] = r_src[RL.base[m] + r * wp + w];
          }
        }
      }
      __syncwarp();
    }
    if (!__any_sync(FULL, busy)) break;
    const T lsr = fast_iteration<T, FRICTION>(P, col, busy ? cnt[0] : 0, busy ? cnt[1] : 0, busy ? cnt[2] : 0, cart, pole);
    if (busy) {
      it++;
      if (lsr <= P.residual_thr || it >= P.solver_iterations) busy = false;
    }
  }
}
#endif  // __CUDACC__

}  // namespace bcp
