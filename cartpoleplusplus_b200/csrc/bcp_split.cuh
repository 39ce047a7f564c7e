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
//           Rows go to a per-env SOLVE RECORD in global memory (L2-resident; row format of
//           bcp_solve.cuh) and the env is queued, predicted-long solves first.
//   k_solve persistent warps; every lane pulls envs off the queue, copies the record's live
//           rows into its shared-memory column and iterates; a lane whose env has converged
//           retires it and pulls the next, so all 32 lanes sweep on every iteration.
// Collision, manifolds, forward dynamics and integration are substep_fast's functions; the
// rows are the same rows in the solver's 16-byte format (bcp_solve.cuh).
#pragma once
#include <cstring>
#include "bcp_lambda.cuh"
#include "bcp_state.cuh"

namespace bcp {

// ---- solve record (global memory, one per env, 16-byte aligned) ----
// words 0..15 header, 16..27 applied normal impulses per slot (warm start in, result out),
// 28..31 pad, then the row area in the layout of solve_layout()
enum { REC_C_GC = 0, REC_C_CP = 1, REC_C_GP = 2, REC_ITERS = 3, REC_DVEL = 4, REC_APN = 16, REC_ROWS = 32 };

template <typename T> BCP_HD int record_words(const Params<T>& P) { return REC_ROWS + solve_layout(P).words; }
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
BCP_HD int slot_of_pair(int m) { return m == BCP_PAIR_GROUND_CART ? (int)SLOT_GC : (m == BCP_PAIR_CART_POLE ? (int)SLOT_CP : (int)SLOT_GP); }

// setup half of a substep for one env: rows + header into `rec`, the unconstrained
// velocities parked in the state (angular in the BODY frame until split_finish).
// The setup in pieces, so that k_prep can stream the manifolds through it one pair at a time:
// setup_begin (poses, forward dynamics), setup_pair x 3 (narrowphase, manifold cache, rows),
// setup_end (record header, velocities parked in the state).
template <typename V> BCP_HD V pick3(const V* a, int m) { return m == 0 ? a[0] : (m == 1 ? a[1] : a[2]); }
// The context is in three parts on purpose: the posed boxes and the narrowphase workspace are
// handed to box_box by address (k_prep keeps them in the thread's shared-memory frame), while
// the bodies and the impulse slots are only ever touched with compile-time offsets, so they
// stay in registers.
template <typename T> struct SetupFrame {  // what box_box reads and scribbles on, and the manifold of the pair being worked on
  Posed<T> cbox, pbox;
  T ws[kBoxWs];      // dBoxBox2's workspace; its <= 4 contacts come back in the first words; once they are in the manifold
                     // the pair's world points (WorldPt x 4 = 24 words) take their place
  Manifold<T> mf;    // in addressable memory: run-time indices (manifold_add_mem / manifold_refresh_mem), no 44 registers
};
static_assert(sizeof(WorldPt<float>) * 4 <= sizeof(float) * kBoxWs, "world points alias the narrowphase workspace");
template <typename T> struct SetupCtx {
  FBody<T> cart, pole;
  T apn_all[N_SLOTS];
};
template <typename T> BCP_HD void ground_box(const Params<T>& P, Posed<T>& g) {
  for (int i = 0; i < 3; ++i) { g.x[i] = P.ground.x0[i]; g.h[i] = P.ground.h[i]; }
  for (int i = 0; i < 9; ++i) g.R[i] = (i % 4 == 0) ? T(1.0) : T(0.0);
}
template <typename T> BCP_HD void setup_begin(const Params<T>& P, const Env<T>& st, SetupFrame<T>& F, SetupCtx<T>& C) {
  Posed<T> cb, pb;  // registers first: the frame only ever sees stores
  pose_box(st.cart, P.cart, cb);
  pose_box(st.pole, P.pole, pb);
  const T zero3[3] = {0, 0, 0};
  fbody_begin(P, st.cart, P.cart, cb.R, st.pend_f, st.pend_t, C.cart);
  fbody_begin(P, st.pole, P.pole, pb.R, zero3, zero3, C.pole);
  F.cbox = cb;
  F.pbox = pb;
  for (int k = 0; k < N_SLOTS; ++k) C.apn_all[k] = T(0.0);
}
// `mf` is the frame's own manifold (F.mf): loaded by the caller, updated here in place
template <typename T, bool FRICTION>
BCP_HD void setup_pair(const Params<T>& P, int m, const Posed<T>& gbox, SetupFrame<T>& F, SetupCtx<T>& C,
                       T* rec, const SolveLayout& L) {
  Manifold<T>& mf = F.mf;
  const Posed<T>& Am = (m == BCP_PAIR_CART_POLE) ? F.cbox : gbox;
  const Posed<T>& Bm = (m == BCP_PAIR_GROUND_CART) ? F.cbox : F.pbox;
  const int nc = box_box(Am.x, Am.R, Am.h, Bm.x, Bm.R, Bm.h, F.ws);
  const Contact<T>* cs = box_contacts(F.ws);
  const Posed<T> A = Am, B = Bm;  // by value: read once, then registers
  // per-pair constants picked by comparison: an array in constant memory indexed by a register
  // is a slow, scoreboarded load
  const T brk = pick3(P.breaking, m);
  const int base = pick3(L.base, m), fbase = pick3(L.fbase, m), rpc = pick3(L.rpc, m);
  for (int k = 0; k < nc; ++k) manifold_add_mem(mf, A, B, cs[k], brk);
  WorldPt<T>* wp = reinterpret_cast<WorldPt<T>*>(F.ws);  // the contacts are in the manifold now
  manifold_refresh_mem(mf, A, B, brk, wp);
  T* rows = rec + REC_ROWS + base;
  T* frows = rec + REC_ROWS + (fbase >= 0 ? fbase : 0);
  if (m == BCP_PAIR_GROUND_CART) {
    put_rows<T, FRICTION, false>(P, rows, frows, C.apn_all + slot_of_pair(BCP_PAIR_GROUND_CART), BCP_PAIR_GROUND_CART, rpc, mf, wp, C.cart, C.cart);
  } else if (m == BCP_PAIR_CART_POLE) {
    put_rows<T, FRICTION, true>(P, rows, frows, C.apn_all + slot_of_pair(BCP_PAIR_CART_POLE), BCP_PAIR_CART_POLE, rpc, mf, wp, C.cart, C.pole);
  } else {
    put_rows<T, FRICTION, false>(P, rows, frows, C.apn_all + slot_of_pair(BCP_PAIR_GROUND_POLE), BCP_PAIR_GROUND_POLE, rpc, mf, wp, C.pole, C.pole);
  }
}
template <typename T> BCP_HD void setup_end(const Params<T>& P, Env<T>& st, const SetupCtx<T>& C, T* rec) {
  rec[REC_C_GC] = as_real(st.mani[BCP_PAIR_GROUND_CART].count, T(0));
  rec[REC_C_CP] = as_real(st.mani[BCP_PAIR_CART_POLE].count, T(0));
  rec[REC_C_GP] = as_real(st.mani[BCP_PAIR_GROUND_POLE].count, T(0));
  T d[12];
  for (int i = 0; i < 3; ++i) {  // warm-start deltas
    d[i] = C.cart.dwl[i]; d[3 + i] = C.cart.dv[i];
    d[6 + i] = C.pole.dwl[i]; d[9 + i] = C.pole.dv[i];
    st.cart.w[i] = C.cart.wl[i]; st.cart.v[i] = C.cart.v[i];
    st.pole.w[i] = C.pole.wl[i]; st.pole.v[i] = C.pole.v[i];
    st.pend_f[i] = 0; st.pend_t[i] = 0;  // clearForcesAndTorques
  }
  for (int k = 0; k < 3; ++k) { st4(rec + REC_DVEL + 4 * k, d + 4 * k); st4(rec + REC_APN + 4 * k, C.apn_all + 4 * k); }
}
// the whole setup on a fully loaded env (host composition below, tests)
template <typename T, bool FRICTION>
BCP_HD void split_setup(const Params<T>& P, Env<T>& st, T* rec, bool on) {
  if (!on) return;
  SetupCtx<T> C;
  SetupFrame<T> F;
  Posed<T> gbox;
  ground_box(P, gbox);
  const SolveLayout L = solve_layout(P);
  setup_begin(P, st, F, C);
  for (int m = 0; m < 3; ++m) {
    F.mf = st.mani[m];
    setup_pair<T, FRICTION>(P, m, gbox, F, C, rec, L);
    st.mani[m] = F.mf;
  }
  setup_end(P, st, C, rec);
}

// what a lane needs to start a solve: contact counts, warm-start deltas and impulses
// sI = sqrt(invI), rI = sqrt(I) per body axis: in and out of the solver's scaled angular deltas
template <typename T> struct SolveScale {
  T cart_sI[3], cart_rI[3], pole_sI[3], pole_rI[3];
};
template <typename T> BCP_HD void solve_consts(const Params<T>& P, SolveRegs<T>& R, SolveScale<T>& S) {
  for (int i = 0; i < 3; ++i) {
    S.cart_sI[i] = bsqrt(T(1.0) / P.cart.I[i]); S.cart_rI[i] = T(1.0) / S.cart_sI[i];
    S.pole_sI[i] = bsqrt(T(1.0) / P.pole.I[i]); S.pole_rI[i] = T(1.0) / S.pole_sI[i];
  }
  R.cart.invM = P.cart.inv_mass;
  R.pole.invM = P.pole.inv_mass;
}
template <typename T> BCP_HD void solve_load(const T* rec, const SolveScale<T>& S, int cnt[3], SolveRegs<T>& R) {
  T h[4], d[12], a[12];
  ld4(rec, h);
  for (int k = 0; k < 3; ++k) { ld4(rec + REC_DVEL + 4 * k, d + 4 * k); ld4(rec + REC_APN + 4 * k, a + 4 * k); }
  cnt[0] = as_int(h[REC_C_GC]); cnt[1] = as_int(h[REC_C_CP]); cnt[2] = as_int(h[REC_C_GP]);
  for (int i = 0; i < 3; ++i) {
    R.cart.wt[i] = d[i] * S.cart_rI[i]; R.cart.dv[i] = d[3 + i];
    R.pole.wt[i] = d[6 + i] * S.pole_rI[i]; R.pole.dv[i] = d[9 + i];
  }
#pragma unroll
  for (int s = 0; s < N_SLOTS; ++s) {
    const int m = s >> 2, j = s & 3;  // SLOT_* = pair * 4
    R.apn[s] = (j < cnt[m]) ? a[s] : T(0.0);
    R.apf[s][0] = 0;
    R.apf[s][1] = 0;
  }
}
// solve results -> record: iteration count, velocity deltas, applied normal impulses (slots
// past the contact count hold 0 and are never read back)
template <typename T> BCP_HD void solve_end(T* rec, const SolveScale<T>& S, int iters, const SolveRegs<T>& R) {
  rec[REC_ITERS] = as_real(iters, T(0));
  T d[12];
  for (int i = 0; i < 3; ++i) {
    d[i] = R.cart.wt[i] * S.cart_sI[i]; d[3 + i] = R.cart.dv[i];
    d[6 + i] = R.pole.wt[i] * S.pole_sI[i]; d[9 + i] = R.pole.dv[i];
  }
  for (int k = 0; k < 3; ++k) { st4(rec + REC_DVEL + 4 * k, d + 4 * k); st4(rec + REC_APN + 4 * k, R.apn + 4 * k); }
}

// finish half of a substep: velocities += deltas, integrate (finish_bodies), impulses ->
// manifold cache (split_finish adds that for a fully loaded env; k_prep does it per pair)
template <typename T> BCP_HD void finish_bodies(const Params<T>& P, Env<T>& st, const T* dvel) {  // dvel: the record's REC_DVEL block
  FBody<T> cart, pole;
  for (int i = 0; i < 3; ++i) {
    cart.dwl[i] = dvel[i]; cart.dv[i] = dvel[3 + i];
    pole.dwl[i] = dvel[6 + i]; pole.dv[i] = dvel[9 + i];
    cart.wl[i] = st.cart.w[i]; cart.v[i] = st.cart.v[i];
    pole.wl[i] = st.pole.w[i]; pole.v[i] = st.pole.v[i];
  }
  quat_to_mat(st.cart.q, cart.R);
  quat_to_mat(st.pole.q, pole.R);
  fbody_finish(P, cart, st.cart);
  fbody_finish(P, pole, st.pole);
}
template <typename T> BCP_HD void split_finish(const Params<T>& P, Env<T>& st, const T* rec) {
  for (int m = 0; m < 3; ++m)
    for (int j = 0; j < st.mani[m].count; ++j) st.mani[m].p[j].imp = rec[REC_APN + slot_of_pair(m) + j];
  finish_bodies(P, st, rec + REC_DVEL);
}

// host-side composition (tests): the whole substep through a record, one env.  Slots the env
// does not use must be zero rows, as in k_solve's columns.
template <typename T, bool FRICTION>
BCP_HD void substep_split_reference(const Params<T>& P, Env<T>& st, T* rec, SubstepInfo* info) {
  const SolveLayout L = solve_layout(P);
  for (int i = 0; i < L.words; ++i) rec[REC_ROWS + i] = 0;
  split_setup<T, FRICTION>(P, st, rec, true);
  int cnt[3];
  SolveRegs<T> R;
  SolveScale<T> S;
  solve_consts(P, R, S);
  solve_load(rec, S, cnt, R);
  int it = 0;
  const int lmode = lam_config_ok(P) ? lam_mode_of(cnt[0], cnt[1], cnt[2]) : (int)LAM_NONE;
  if (lmode != LAM_NONE && !getenv("BCP_NO_LAMBDA")) {  // the impulse-space iteration of k_solve_lam (bcp_lambda.cuh)
    LamSys<T> Sy;
    const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
    if (lmode == LAM_STACK) lam_build_stack<false>(P, rec + REC_ROWS, L, R.apn, Sy);
    else lam_build_split<false, FRICTION>(P, rec + REC_ROWS, L, R.apn, Sy);
    for (int k = 0; k < P.solver_iterations; ++k) {
      const T lsr = lmode == LAM_STACK ? lam_round_stack(P, cnt[BCP_PAIR_CART_POLE], cnt[BCP_PAIR_GROUND_CART], Sy)
                                       : lam_round_split<FRICTION>(P, fr, cnt[BCP_PAIR_GROUND_POLE], cnt[BCP_PAIR_GROUND_CART], Sy);
      it = k + 1;
      if (lsr <= P.residual_thr || k >= P.solver_iterations - 1) break;
    }
    if (lmode == LAM_STACK) lam_finish_stack<false>(P, rec + REC_ROWS, L, Sy, R);
    else lam_finish_split<false, FRICTION>(P, rec + REC_ROWS, L, Sy, R);
  } else if (lam_config_ok(P) && !getenv("BCP_NO_LAMBDA") && !getenv("BCP_NO_LAMGEN")) {  // k_solve_gen_once's iteration
    LamGen<T> Sy;
    const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
    lam_build_gen<false, FRICTION>(P, rec + REC_ROWS, L, R.apn, Sy);
    for (int k = 0; k < P.solver_iterations; ++k) {
      const T lsr = lam_round_gen<FRICTION>(P, fr, Sy);
      it = k + 1;
      if (lsr <= P.residual_thr || k >= P.solver_iterations - 1) break;
    }
    lam_finish_gen<false, FRICTION>(P, rec + REC_ROWS, L, Sy, R);
  } else {
    for (int k = 0; k < P.solver_iterations; ++k) {
      const T lsr = solve_iteration<T, FRICTION>(P, rec + REC_ROWS, L, cnt, R);
      it = k + 1;
      if (lsr <= P.residual_thr || k >= P.solver_iterations - 1) break;
    }
  }
  solve_end(rec, S, it, R);
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
  // layout of `obs`: floats between two envs and between two repeats of an env.  (14 R, 14): the (B, R, 2, 7) block of
  // bcp_step; (14, 14 B): repeat-major slabs [R][B][2][7], what bcp_step_host copies to the host slab by slab
  size_t env_stride, rep_stride;
  // bcp_rollout_random through this pipeline: > 0 parks an env that is done with `episode` >= max_episodes (it has delivered
  // its share of the protocol's episodes), and the closing kernel also accumulates the rollout's counters (substeps, resets,
  // reward sum - step() returns the rewards to its caller instead)
  int max_episodes, rollout;
};

// Visiting order of the envs.  The kernel that closes a call files every env under its contact
// regime (which of the three pairs hold points: 8 bins); the next call's k_prep launches walk
// the bins, so that the 32 envs of a warp - and the 32 queue entries a solver warp takes - run
// the same narrowphase branches and need the same row slots.  cnt == nullptr: identity order.
enum { N_BINS = 8 };
struct VisitOrder {
  const int* bins;  // [N_BINS][N] env ids
  const int* cnt;   // [N_BINS]
  int* next_bins;   // filled by the closing kernel
  int* next_cnt;
  int rev;          // walk the bins from the contact-rich regimes down (their blocks are the slow ones: first wave)
};
__device__ __forceinline__ int visit(const VisitOrder& vo, int i, int N) {
  if (!vo.cnt) return i < N ? i : -1;
  int at = i;
#pragma unroll
  for (int k = 0; k < N_BINS; ++k) {
    const int b = vo.rev ? N_BINS - 1 - k : k;
    const int c = vo.cnt[b];
    if (at < c) return vo.bins[(size_t)b * N + at];
    at -= c;
  }
  return -1;
}

// a finished solve's normal impulses into the cached points of pair m, in place in the record
// (the only thing a finish changes in a manifold)
// slot j of pair m out of the 12 applied impulses, and one of three counts, by comparison: an index that is only known at
// run time (m of the rolled pair loop) would put the whole array - and the Env record with it - into local memory
template <typename T> BCP_HD T apn_of(const T* apn, int m, int j) {
  return m == BCP_PAIR_GROUND_CART ? apn[SLOT_GC + j] : (m == BCP_PAIR_CART_POLE ? apn[SLOT_CP + j] : apn[SLOT_GP + j]);
}
template <typename T> BCP_HD int count_of(const Env<T>& st, int m) {
  return m == 0 ? st.mani[0].count : (m == 1 ? st.mani[1].count : st.mani[2].count);
}
template <typename T> BCP_HD void set_count(Env<T>& st, int m, int c) {
  if (m == 0) st.mani[0].count = c;
  else if (m == 1) st.mani[1].count = c;
  else st.mani[2].count = c;
}
template <typename T> BCP_HD void finish_impulses(T* f, size_t e, int m, int count, const T* fin_apn) {
  T* blk = f + e * NF + BCP_S_MANI + m * BCP_MANI_STRIDE;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (j < count) blk[j * BCP_MANI_POINT + 10] = apn_of(fin_apn, m, j);
}
template <typename T> BCP_HD int queue_meta(const Env<T>& st) {
  return st.mani[0].count | (st.mani[1].count << 8) | (st.mani[2].count << 16);
}
// s_fin: substep to finish (-1: none, this is the first kernel of the call)
// s_set: substep to set up  (-1: none, this is the last kernel of the call)
// Two resident blocks per SM: the register budget (up to 255) is what lets the manifold being
// worked on, its world points and most of the row arithmetic stay in registers; at four blocks
// per SM (128 registers) they spill to a stack that does not fit L1 (measured: 0.268 vs 0.245 ms
// per env.step at 65 536 envs, 0.87 vs 0.70 ms at 262 144).
#ifndef BCP_PREP_MINB
#define BCP_PREP_MINB 2
#endif
// k_prep's dynamic shared memory: the ground box, then one SetupFrame per thread
constexpr int kPrepGround = 16;
constexpr int kPrepFrame = (int)(sizeof(SetupFrame<float>) / sizeof(float)) | 1;
inline size_t prep_smem_bytes(int block) { return sizeof(float) * ((size_t)kPrepGround + (size_t)block * kPrepFrame); }
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// SETUP = false: the instance for the kernel that closes a call (s_set < 0).  It has no narrowphase, no manifold work and
// no rows to write, so it needs neither the shared-memory frame nor 240 registers: compiled without the set-up branch it
// runs at four blocks per SM instead of two (the slots of resident warps are what bounds the step).
template <typename T, bool FRICTION, bool SETUP = true>
__global__ void __launch_bounds__(kBlock, SETUP ? BCP_PREP_MINB : 4) k_prep(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id,
                                                 T* __restrict__ recs, int rec_words, int2* __restrict__ queue,
                                                 int* __restrict__ qc_set, SplitIO io, Stats stats, VisitOrder vo, int s_fin,
                                                 int s_set_arg, int use_lam) {
  const int s_set = SETUP ? s_set_arg : -1;
  const int N = P.num_envs;
  const int e_visit = visit(vo, blockIdx.x * blockDim.x + threadIdx.x, N);
  const bool valid = e_visit >= 0;
  const int e = valid ? e_visit : 0;
  const int64_t genv = P.env_id_offset + e;
  T* rec = recs + (size_t)(valid ? e : 0) * rec_words;
  float* obs = (io.obs && valid) ? io.obs + (size_t)e * io.env_stride : nullptr;
  Env<T> st;
  st.done = 1;
  st.reset_left = 0;
  int mode = MODE_IDLE;
  for (int m = 0; m < 3; ++m) st.mani[m].count = 0;
  // Everything this env will need from global memory whose address only depends on the env id
  // is requested here, before any of it is used and whether or not its phase will use it: the
  // state core, the previous solve's results in the record, the action.  One exposed DRAM
  // latency instead of one per dependent decision (a thread has nothing else to overlap it with
  // at two warps per scheduler).
  T fin_apn[N_SLOTS];  // applied normal impulses of the solve being finished -> manifold caches
  T fin_dvel[12];      // its velocity deltas
  int act_i = 0, last_iters = 0;
  float2 act_f = make_float2(0.f, 0.f);
  if (valid) {
    if (s_fin >= 0) {
      for (int k = 0; k < 3; ++k) { ld4(rec + REC_APN + 4 * k, fin_apn + 4 * k); ld4(rec + REC_DVEL + 4 * k, fin_dvel + 4 * k); }
      if (io.actions) {
        if (P.discrete) act_i = reinterpret_cast<const int32_t*>(io.actions)[e];
        else act_f = reinterpret_cast<const float2*>(io.actions)[e];
      }
    }
    if (s_set >= 0) last_iters = as_int(rec[REC_ITERS]);
    load_env_core<T>(f, id, e, st);  // the manifold points are streamed per pair below
    mode = id[(size_t)e * NI + I_MODE];
    if (s_set >= 0) {  // ... and their blocks are asked into L2 now
#pragma unroll
      for (int m = 0; m < 3; ++m)
        if (st.mani[m].count > 0) {
          const T* blk = f + (size_t)e * NF + BCP_S_MANI + m * BCP_MANI_STRIDE;
          prefetch_l2(blk);
          prefetch_l2(blk + BCP_MANI_STRIDE - 1);
        }
    }
  }
  if (s_fin < 0) {
    const bool idle_done = valid && st.done && st.reset_left == 0;
    const bool parked = io.max_episodes > 0 && idle_done && st.episode >= io.max_episodes;
    mode = call_begin(P, st, valid && !parked, false);
    if (P.settled && idle_done && mode != MODE_IDLE) {  // a reset began on the cached post-settle state: its manifold blocks too
#pragma unroll
      for (int m = 0; m < 3; ++m)
        if (st.mani[m].count > 0) {
          T blk[BCP_MANI_STRIDE];
          copy16(P.settled + BCP_S_MANI + m * BCP_MANI_STRIDE, blk, BCP_MANI_STRIDE);
          store16(f + (size_t)e * NF + BCP_S_MANI + m * BCP_MANI_STRIDE, blk, BCP_MANI_STRIDE);
        }
    }
  }
  const bool fin = s_fin >= 0 && valid && substep_on(mode, st);
  if (fin) finish_bodies(P, st, fin_dvel);
  if (s_fin >= 0 && valid) {
    int a = 0;
    float ax = 0.f, ay = 0.f;
    if (mode == MODE_LIVE) {
      a = act_i;
      ax = act_f.x; ay = act_f.y;
    }
    T fx, fy;
    action_force(P, a, ax, ay, fx, fy);
    post_substep(P, st, genv, mode, fx, fy);
    if ((s_fin + 1) % P.S == 0 && mode == MODE_LIVE && obs) write_obs(st, obs + io.rep_stride * (s_fin / P.S));  // :237
  }
  bool q_on = false, q_gen = false, q_long = false;  // the env's queue entry (written at the end of the kernel)
  unsigned ml = 0, ms = 0, mg = 0;
  int bl = 0, bs = 0, bg = 0;
  if (SETUP && s_set >= 0) {
    const bool on = valid && substep_on(mode, st);
    if (!on) last_iters = 0;
    SetupCtx<T> C;
    const SolveLayout L = solve_layout(P);
    // the thread's frame in shared memory ([thread][kPrepFrame] words, odd stride: conflict-free
    // when the lanes of a warp touch the same member); the ground box once per block
    T* const smem = reinterpret_cast<T*>(bcp_smem);
    Posed<T>& gbox = *reinterpret_cast<Posed<T>*>(smem);
    SetupFrame<T>& F = *reinterpret_cast<SetupFrame<T>*>(smem + kPrepGround + (size_t)threadIdx.x * kPrepFrame);
    if (threadIdx.x == 0) ground_box(P, gbox);
    __syncthreads();
    if (on) setup_begin(P, st, F, C);
#pragma unroll 1
    for (int m = 0; m < 3; ++m) {
      BCP_WARP_SYNC();
      if (on) {
        Manifold<T>& mf = F.mf;
        mf.count = count_of(st, m);
        load_manifold_mem<T>(f, e, m, mf);
        if (fin) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (j < mf.count) mf.p[j].imp = apn_of(fin_apn, m, j);
        }
        setup_pair<T, FRICTION>(P, m, gbox, F, C, rec, L);
        set_count(st, m, mf.count);
        store_manifold_mem<T>(f, e, m, mf);
      } else if (fin) {
        finish_impulses(f, e, m, count_of(st, m), fin_apn);
      }
    }
    BCP_WARP_SYNC();
    // queue the env: one atomic per warp and bin, ISSUED here and used at the very end of the kernel - the round trip of a
    // contended atomic (a microsecond) then runs beside the stores of setup_end and of the state core instead of stalling
    // the whole warp in front of them (it was 8 % of the kernel's stall samples).
    // With the impulse-space solver on the main queue, an env in a regime it does not cover goes to
    // the generic queue (second half of `queue`, second counter block)
    q_on = on;
    q_gen = on && use_lam &&
        lam_mode_of(st.mani[BCP_PAIR_GROUND_CART].count, st.mani[BCP_PAIR_CART_POLE].count, st.mani[BCP_PAIR_GROUND_POLE].count) == LAM_NONE;
    q_long = on && !q_gen && last_iters >= kLongIters;
    ml = __ballot_sync(0xffffffffu, q_long);
    ms = __ballot_sync(0xffffffffu, on && !q_long && !q_gen);
    mg = __ballot_sync(0xffffffffu, q_gen);
    if ((threadIdx.x & 31) == 0) {
      if (ml) bl = atomicAdd(&qc_set[Q_LONG], __popc(ml));
      if (ms) bs = atomicAdd(&qc_set[Q_SHORT], __popc(ms));
      if (mg) bg = atomicAdd(&qc_set[Q_WORDS + Q_SHORT], __popc(mg));
    }
    if (on) setup_end(P, st, C, rec);
  } else {  // the call closes: steps / done / reward (:239-260), outputs, episode statistics
    if (fin) {
#pragma unroll
      for (int m = 0; m < 3; ++m) finish_impulses(f, e, m, st.mani[m].count, fin_apn);
    }
    bool live = false;
    const int rs = call_end(P, st, mode, obs, live, io.rep_stride);
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
      if (io.rollout) {  // k_rollout's other counters: substeps of every env that took part in the call, finished resets, reward
        const unsigned on_n = __popc(__ballot_sync(0xffffffffu, valid && mode != MODE_IDLE));
        const unsigned fr_n = __popc(__ballot_sync(0xffffffffu, valid && rs == BCP_REASON_FRESH));
        if ((threadIdx.x & 31) == 0) {
          if (on_n) atomicAdd(&stats.header[2], (unsigned long long)on_n * (unsigned long long)(P.R * P.S));
          if (fr_n) atomicAdd(&stats.header[3], (unsigned long long)fr_n);
          if (w && stats.reward_sum) atomicAdd(stats.reward_sum, (double)w);
        }
      }
    }
    if (vo.next_cnt) {  // file the env under the contact regime its next call will run in
      int key = (st.mani[BCP_PAIR_GROUND_POLE].count > 0 ? 4 : 0) | (st.mani[BCP_PAIR_CART_POLE].count > 0 ? 2 : 0) |
                (st.mani[BCP_PAIR_GROUND_CART].count > 0 ? 1 : 0);
      if (st.done && st.reset_left == 0 && P.auto_reset == BCP_AUTO_RESET_PIPELINED) key = 3;  // teleported onto the cart
      const unsigned act = __ballot_sync(0xffffffffu, valid);
      if (valid) {
        const unsigned peers = __match_any_sync(act, key);
        const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(&vo.next_cnt[key], __popc(peers));
        base = __shfl_sync(peers, base, leader);
        vo.next_bins[(size_t)key * N + base + __popc(peers & ((1u << lane) - 1))] = e;
      }
    }
  }
  if (valid) {
    store_env_core<T>(f, id, e, st);
    id[(size_t)e * NI + I_MODE] = mode;
  }
  if (SETUP && s_set >= 0) {  // the queue entry, at the slots the atomics above handed out
    const int lane = threadIdx.x & 31;
    bl = __shfl_sync(0xffffffffu, bl, 0);
    bs = __shfl_sync(0xffffffffu, bs, 0);
    bg = __shfl_sync(0xffffffffu, bg, 0);
    const unsigned lt = (1u << lane) - 1;
    // entry = env id + its three contact counts (the solver sizes its copy without the record)
    const int2 entry = make_int2(e, queue_meta(st));
    if (q_long) queue[bl + __popc(ml & lt)] = entry;
    else if (q_gen) queue[N + N - 1 - (bg + __popc(mg & lt))] = entry;
    else if (q_on) queue[N - 1 - (bs + __popc(ms & lt))] = entry;
  }
}

// bcp_rollout_random through the split pipeline: the random agent's action of every env for the call that follows, the same
// draw as k_rollout's (Philox keyed by global env id, episode and step: draw_action, bcp_episode.cuh)
template <typename T>
__global__ void k_draw_actions(const __grid_constant__ Params<T> P, const int* __restrict__ id, void* __restrict__ actions, uint32_t mask) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= P.num_envs) return;
  Env<T> st;
  st.episode = id[(size_t)e * NI + I_EPISODE];
  st.steps = id[(size_t)e * NI + I_STEPS];
  int a = 0;
  float ax = 0.f, ay = 0.f;
  draw_action(P, P.env_id_offset + e, st, mask, a, ax, ay);
  if (P.discrete) reinterpret_cast<int32_t*>(actions)[e] = a;
  else reinterpret_cast<float2*>(actions)[e] = make_float2(ax, ay);
}

// persistent pooled solve: every lane keeps pulling envs until the queue is dry.
// Shared memory: one column per lane holding a copy of the env's whole solve record, lane-major
// ([lane][col_stride] words, col_stride / 4 odd): conflict-free for 32 lanes reading the same
// float4 of their own columns (the sweep).  A lane that takes an env does not wait for it: the
// warp issues the record -> column copy as 16-byte cp.async (zero-filled for the row slots the
// env does not use, so the sweep stays branch-free), runs one more iteration of the lanes that
// are busy, and only then starts the staged lanes.
constexpr int kFetchMin = 10;  // idle lanes that trigger a fetch round

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gmem_src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

constexpr int kMaxPieces = 4;  // 16-byte pieces of a record one lane copies (records <= 512 words)

// What k_solve and k_solve_lam share: a warp's slice of the queue, and the staged copy of the records it takes into the
// lanes' shared-memory columns.
//   init : this lane moves the 16-byte pieces lane, lane + 32, ... of every record its warp takes.  What block a piece
//          lies in never changes, so it is worked out once: piece k is live iff
//          pc_off[k] < count(pair pc_shift[k] / 8) * pc_per[k]   (header pieces: always).
//   take : queue entries for the idle lanes (one atomic per warp per kFetchChunk entries; long-predicted solves come
//          from the front of the queue, the rest from the back), then the whole warp issues the record -> column copies
//          as 16-byte cp.async, zero-filled for the row slots an env does not use, and does not wait for them.
//          Returns this lane's env (-1: none); `staged` = lanes whose copy is in flight.
template <typename T> struct PoolFeed {
  int pc_shift[kMaxPieces], pc_off[kMaxPieces], pc_per[kMaxPieces];
  int w_next, w_end;  // the warp's slice of the queue (uniform)
  int n_long, total, N, rec16;
  bool exhausted;
  __device__ __forceinline__ void init(const SolveLayout& L, int lane, int rec_words, const int* qc, int n_envs) {
    N = n_envs;
    n_long = qc[Q_LONG];
    total = n_long + qc[Q_SHORT];
    rec16 = (rec_words * (int)sizeof(T)) >> 4;
    const int hdr16 = (REC_ROWS * (int)sizeof(T)) >> 4;
#pragma unroll
    for (int k = 0; k < kMaxPieces; ++k) {
      const int i = lane + 32 * k;
      pc_shift[k] = 0; pc_off[k] = (i < hdr16) ? -1 : 0x7fffffff; pc_per[k] = 0;  // header piece / outside every block
#pragma unroll
      for (int m = 0; m < 3; ++m) {
        const int w = (m == BCP_PAIR_CART_POLE) ? (int)SW2 : (int)SW1;
        const int lo = ((REC_ROWS + L.base[m]) * (int)sizeof(T)) >> 4, per = (w * (int)sizeof(T)) >> 4;
        if (i >= lo && i < lo + 4 * per) { pc_shift[k] = 8 * m; pc_off[k] = i - lo; pc_per[k] = per; }
        if (L.fbase[m] >= 0) {
          const int flo = ((REC_ROWS + L.fbase[m]) * (int)sizeof(T)) >> 4, fper = ((L.rpc[m] - 1) * w * (int)sizeof(T)) >> 4;
          if (i >= flo && i < flo + 4 * fper) { pc_shift[k] = 8 * m; pc_off[k] = i - flo; pc_per[k] = fper; }
        }
      }
    }
    w_next = 0;
    w_end = 0;
    exhausted = total == 0;
  }
  __device__ __forceinline__ int take(unsigned idle, int lane, const int2* __restrict__ queue, int* __restrict__ qc, const T* __restrict__ recs,
                                      int rec_words, T* smem, int warp_base, int col_stride, unsigned& staged) {
    const unsigned FULL = 0xffffffffu, lt = (1u << lane) - 1;
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
    int env = -1, meta = 0;
    if (idx >= 0) {
      const int2 q = queue[idx < n_long ? idx : N - 1 - (idx - n_long)];
      env = q.x;
      meta = q.y;
    }
    staged = __ballot_sync(FULL, idx >= 0);
    stage(staged, lane, env, meta, recs, rec_words, smem, warp_base, col_stride);
    return env;
  }
  // the whole warp issues the record -> column copies of the lanes in `taken` (lane l: env / meta of its own record)
  __device__ __forceinline__ void stage(unsigned taken, int lane, int env, int meta, const T* __restrict__ recs, int rec_words, T* smem,
                                        int warp_base, int col_stride) {
    const unsigned FULL = 0xffffffffu;
    while (taken) {
      const int src = __ffs(taken) - 1;
      taken &= taken - 1;
      const int e_src = __shfl_sync(FULL, env, src), m_src = __shfl_sync(FULL, meta, src);
      const char* r_src = reinterpret_cast<const char*>(recs + (size_t)e_src * rec_words) + 16 * lane;
      char* c_dst = reinterpret_cast<char*>(smem + (size_t)(warp_base + src) * col_stride) + 16 * lane;
#pragma unroll
      for (int k = 0; k < kMaxPieces; ++k) {
        if (lane + 32 * k < rec16) {
          const bool live = pc_off[k] < (int)((m_src >> pc_shift[k]) & 0xff) * pc_per[k];
          cp_async16(c_dst + 512 * k, r_src + 512 * k, live ? 16 : 0);
        }
      }
    }
  }
};


template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock) k_solve(const __grid_constant__ Params<T> P, T* __restrict__ recs, int rec_words,
                                                  const int2* __restrict__ queue, int* __restrict__ qc, int col_stride,
                                                  int fetch_min) {
  const unsigned FULL = 0xffffffffu;
  T* const smem = reinterpret_cast<T*>(bcp_smem);
  const int lane = threadIdx.x & 31, warp_base = threadIdx.x - lane;
  T* const col = smem + (size_t)threadIdx.x * col_stride;
  const SolveLayout L = solve_layout(P);
  const int N = P.num_envs;
  PoolFeed<T> feed;  // the warp's slice of the queue and the record -> column copy (shared with k_solve_lam)
  feed.init(L, lane, rec_words, qc, N);
  SolveRegs<T> R;
  SolveScale<T> S;
  solve_consts(P, R, S);
  int cnt[3] = {0, 0, 0}, mx[3] = {0, 0, 0};
  int it = 0;
  T* rec = nullptr;
  unsigned busy_m = 0, staged_m = 0;  // lanes iterating / lanes whose copy is in flight (warp-uniform)
  const T res_thr = P.residual_thr;
  const int max_it = P.solver_iterations;
  while (true) {
    if (staged_m) {  // copies issued one iteration ago: start those lanes
      cp_async_wait_all();
      __syncwarp();
      if ((staged_m >> lane) & 1u) {
        solve_load(col, S, cnt, R);
        it = 0;
      }
      busy_m |= staged_m;
      staged_m = 0;
      const bool busy = (busy_m >> lane) & 1u;
#pragma unroll
      for (int m = 0; m < 3; ++m) mx[m] = __reduce_max_sync(FULL, busy ? cnt[m] : 0);
    }
    const unsigned idle = ~busy_m;
    if (idle && !feed.exhausted && (idle == FULL || __popc(idle) >= fetch_min)) {
      const int env = feed.take(idle, lane, queue, qc, recs, rec_words, smem, warp_base, col_stride, staged_m);
      if (env >= 0) rec = recs + (size_t)env * rec_words;
    }
    if (!busy_m) {
      if (!staged_m) break;
      continue;
    }
    const T lsr = solve_iteration<T, FRICTION, true>(P, col + REC_ROWS, L, mx, R);
    const bool busy = (busy_m >> lane) & 1u;
    bool conv = false;
    if (busy) {
      it++;
      conv = lsr <= res_thr || it >= max_it;
    }
    const unsigned cm = __ballot_sync(FULL, conv);
    if (cm) {
      if (conv) solve_end(rec, S, it, R);  // retire at once: the lane sweeps its stale column until refilled
      busy_m &= ~cm;
      const bool still = (busy_m >> lane) & 1u;
#pragma unroll
      for (int m = 0; m < 3; ++m) mx[m] = __reduce_max_sync(FULL, still ? cnt[m] : 0);
    }
  }
}
// k_solve in impulse space (bcp_lambda.cuh): same queue, same staged record -> column copies, same
// retire-and-refill; a lane builds its env's Delassus matrix when it starts the env and iterates on
// registers.  Serves the envs k_prep queued on the main queue when the configuration allows it
// (lam_config_ok); envs in a regime it does not cover go to the generic queue and k_solve.
template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock, 2) k_solve_lam(const __grid_constant__ Params<T> P, T* __restrict__ recs, int rec_words,
                                                  const int2* __restrict__ queue, int* __restrict__ qc, int col_stride,
                                                  int fetch_min, T* __restrict__ cont, int4* __restrict__ contm, int handoff_it) {
  const unsigned FULL = 0xffffffffu;
  T* const smem = reinterpret_cast<T*>(bcp_smem);
  const int lane = threadIdx.x & 31, warp_base = threadIdx.x - lane;
  T* const col = smem + (size_t)threadIdx.x * col_stride;
  const SolveLayout L = solve_layout(P);
  const int N = P.num_envs;
  const unsigned lt = (1u << lane) - 1;
  PoolFeed<T> feed;  // the warp's slice of the queue and the record -> column copy (shared with k_solve)
  feed.init(L, lane, rec_words, qc, N);
  SolveScale<T> S;
  {
    SolveRegs<T> R0;
    solve_consts(P, R0, S);
  }
  LamSys<T> Sy;  // the lane's env in impulse space (bcp_lambda.cuh)
  int lmode = LAM_NONE;
  const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
  int cnt[3] = {0, 0, 0}, mx[3] = {0, 0, 0};
  int it = 0;
  T* rec = nullptr;
  unsigned busy_m = 0, staged_m = 0;  // lanes iterating / lanes whose copy is in flight (warp-uniform)
  const T res_thr = P.residual_thr;
  const int max_it = P.solver_iterations;
  while (true) {
    if (staged_m) {  // copies issued one iteration ago: start those lanes
      cp_async_wait_all();
      __syncwarp();
      if ((staged_m >> lane) & 1u) {
        SolveRegs<T> R;
        solve_load(col, S, cnt, R);  // counts and warm-start impulses
        lmode = lam_mode_of(cnt[0], cnt[1], cnt[2]);
        if (lmode == LAM_STACK) lam_build_stack<true>(P, col + REC_ROWS, L, R.apn, Sy);
        else lam_build_split<true, FRICTION>(P, col + REC_ROWS, L, R.apn, Sy);
        it = 0;
      }
      busy_m |= staged_m;
      staged_m = 0;
      const bool busy = (busy_m >> lane) & 1u;
#pragma unroll
      for (int m = 0; m < 3; ++m) mx[m] = __reduce_max_sync(FULL, busy ? cnt[m] : 0);
    }
    const unsigned idle = ~busy_m;
    if (idle && !feed.exhausted && (idle == FULL || __popc(idle) >= fetch_min)) {
      const int env = feed.take(idle, lane, queue, qc, recs, rec_words, smem, warp_base, col_stride, staged_m);
      if (env >= 0) rec = recs + (size_t)env * rec_words;
    }
    if (!busy_m) {
      if (!staged_m) break;
      continue;
    }
    T lsr = 0;  // idle lanes iterate their stale system: registers only, nothing is written
    if (lmode == LAM_STACK) lsr = lam_round_stack(P, mx[BCP_PAIR_CART_POLE], mx[BCP_PAIR_GROUND_CART], Sy);
    else lsr = lam_round_split<FRICTION>(P, fr, mx[BCP_PAIR_GROUND_POLE], mx[BCP_PAIR_GROUND_CART], Sy);
    __syncwarp();
    const bool busy = (busy_m >> lane) & 1u;
    bool conv = false;
    if (busy) {
      it++;
      conv = lsr <= res_thr || it >= max_it;
    }
    // a solve still running after handoff_it iterations leaves the pool: its system goes to the
    // continuation buffer (STACK entries from the front, SPLIT from the back) and k_solve_tail
    // finishes it in a densely packed, regime-pure warp
    const bool hand = busy && !conv && handoff_it > 0 && it >= handoff_it;
    const unsigned hm = __ballot_sync(FULL, hand);
    if (hm) {
      const unsigned hs = __ballot_sync(FULL, hand && lmode == LAM_STACK), hp = hm & ~hs;
      int b0 = 0, b1 = 0;
      if (lane == 0) {
        if (hs) b0 = atomicAdd(&qc[2 * Q_WORDS + 0], __popc(hs));
        if (hp) b1 = atomicAdd(&qc[2 * Q_WORDS + 1], __popc(hp));
      }
      b0 = __shfl_sync(FULL, b0, 0);
      b1 = __shfl_sync(FULL, b1, 0);
      if (hand) {
        const int slot = lmode == LAM_STACK ? b0 + __popc(hs & lt) : N - 1 - (b1 + __popc(hp & lt));
        lam_store(cont, N, slot, Sy);
        contm[slot] = make_int4((int)((rec - recs) / rec_words), it, cnt[0] | (cnt[1] << 8) | (cnt[2] << 16), lmode);
      }
    }
    const unsigned cm = __ballot_sync(FULL, conv) | hm;
    if (cm) {
      if (conv) {  // retire at once: bodies' deltas from the impulses and the rows still in the column
        SolveRegs<T> R;
        R.cart.invM = P.cart.inv_mass;
        R.pole.invM = P.pole.inv_mass;
        if (lmode == LAM_STACK) lam_finish_stack<true>(P, col + REC_ROWS, L, Sy, R);
        else lam_finish_split<true, FRICTION>(P, col + REC_ROWS, L, Sy, R);
        solve_end(rec, S, it, R);
      }
      busy_m &= ~cm;
      const bool still = (busy_m >> lane) & 1u;
#pragma unroll
      for (int m = 0; m < 3; ++m) mx[m] = __reduce_max_sync(FULL, still ? cnt[m] : 0);
    }
  }
}
// k_solve_lam without the pool: warp w takes queue entries 32 w .. 32 w + 31, once.  With the hand-off to k_solve_tail a
// solve spends at most `handoff_it` iterations here, so there is no long tail for a refill to hide - and the launch has
// about one lane per queued env anyway -: the pool's slice atomics, refill rounds and per-round retire code were a third of
// k_solve_lam's instructions and 45 % of its stall samples (profiles/r2_k_solve_lam_by_source.txt).  Same staged
// record -> column copy, same per-lane arithmetic (build, rounds, finish, hand-off), so the results are the pool's bit for
// bit; a lane that converges early waits for its warp (at most handoff_it rounds) and all finishes run together.
template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock, 2) k_solve_lam_once(const __grid_constant__ Params<T> P, T* __restrict__ recs, int rec_words,
                                                       const int2* __restrict__ queue, int* __restrict__ qc, int col_stride,
                                                       T* __restrict__ cont, int4* __restrict__ contm, int handoff_it) {
  const unsigned FULL = 0xffffffffu;
  T* const smem = reinterpret_cast<T*>(bcp_smem);
  const int lane = threadIdx.x & 31, warp_base = threadIdx.x - lane;
  T* const col = smem + (size_t)threadIdx.x * col_stride;
  const SolveLayout L = solve_layout(P);
  const int N = P.num_envs;
  const unsigned lt = (1u << lane) - 1;
  PoolFeed<T> feed;  // its piece table and queue geometry; nothing is fetched through it
  feed.init(L, lane, rec_words, qc, N);
  SolveScale<T> S;
  {
    SolveRegs<T> R0;
    solve_consts(P, R0, S);
  }
  const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
  const T res_thr = P.residual_thr;
  const int max_it = P.solver_iterations;
  const int warps = gridDim.x * (blockDim.x >> 5);
  for (int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w * 32 < feed.total; w += warps) {
    const int idx = w * 32 + lane;
    const bool have = idx < feed.total;
    int env = 0, meta = 0;
    if (have) {
      const int2 q = queue[idx < feed.n_long ? idx : N - 1 - (idx - feed.n_long)];
      env = q.x;
      meta = q.y;
    }
    feed.stage(__ballot_sync(FULL, have), lane, env, meta, recs, rec_words, smem, warp_base, col_stride);
    cp_async_wait_all();
    __syncwarp();
    T* const rec = recs + (size_t)env * rec_words;
    LamSys<T> Sy;
    int lmode = LAM_NONE, it = 0;
    int cnt[3] = {0, 0, 0}, mx[3];
    if (have) {
      SolveRegs<T> R;
      solve_load(col, S, cnt, R);  // counts and warm-start impulses
      lmode = lam_mode_of(cnt[0], cnt[1], cnt[2]);
      if (lmode == LAM_STACK) lam_build_stack<true>(P, col + REC_ROWS, L, R.apn, Sy);
      else lam_build_split<true, FRICTION>(P, col + REC_ROWS, L, R.apn, Sy);
    }
#pragma unroll
    for (int m = 0; m < 3; ++m) mx[m] = __reduce_max_sync(FULL, have ? cnt[m] : 0);
    bool open = have, conv = false;
    while (__any_sync(FULL, open)) {
      if (open) {
        const T lsr = lmode == LAM_STACK ? lam_round_stack(P, mx[BCP_PAIR_CART_POLE], mx[BCP_PAIR_GROUND_CART], Sy)
                                         : lam_round_split<FRICTION>(P, fr, mx[BCP_PAIR_GROUND_POLE], mx[BCP_PAIR_GROUND_CART], Sy);
        it++;
        conv = lsr <= res_thr || it >= max_it;
        if (conv || (handoff_it > 0 && it >= handoff_it)) open = false;
      }
    }
    // still running after handoff_it iterations: the system goes to the continuation buffer (STACK entries from the
    // front, SPLIT from the back) and k_solve_tail
    const bool hand = have && !conv;
    const unsigned hm = __ballot_sync(FULL, hand);
    if (hm) {
      const unsigned hs = __ballot_sync(FULL, hand && lmode == LAM_STACK), hp = hm & ~hs;
      int b0 = 0, b1 = 0;
      if (lane == 0) {
        if (hs) b0 = atomicAdd(&qc[2 * Q_WORDS + 0], __popc(hs));
        if (hp) b1 = atomicAdd(&qc[2 * Q_WORDS + 1], __popc(hp));
      }
      b0 = __shfl_sync(FULL, b0, 0);
      b1 = __shfl_sync(FULL, b1, 0);
      if (hand) {
        const int slot = lmode == LAM_STACK ? b0 + __popc(hs & lt) : N - 1 - (b1 + __popc(hp & lt));
        lam_store(cont, N, slot, Sy);
        contm[slot] = make_int4(env, it, cnt[0] | (cnt[1] << 8) | (cnt[2] << 16), lmode);
      }
    }
    if (have && conv) {  // bodies' deltas from the impulses and the rows still in the column
      SolveRegs<T> R;
      R.cart.invM = P.cart.inv_mass;
      R.pole.invM = P.pole.inv_mass;
      if (lmode == LAM_STACK) lam_finish_stack<true>(P, col + REC_ROWS, L, Sy, R);
      else lam_finish_split<true, FRICTION>(P, col + REC_ROWS, L, Sy, R);
      solve_end(rec, S, it, R);
    }
    __syncwarp();  // the columns are reused by the warp's next chunk
  }
}
// The generic queue (pole touching cart AND ground: a few dozen envs per launch) in impulse space, one env per lane, one
// chunk of the queue per warp, to convergence (bcp_lambda.cuh, LamGen).  These solves mostly run all 50 iterations; row by
// row (k_solve) they were the longest dependent chain of the solve phase and the next k_prep waited for them: capping them
// at 30 iterations - wrong on purpose, as a measurement - shortened env.step by 7 %.
template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock, 1) k_solve_gen_once(const __grid_constant__ Params<T> P, T* __restrict__ recs, int rec_words,
                                                              const int2* __restrict__ queue, int* __restrict__ qc, int col_stride) {
  const unsigned FULL = 0xffffffffu;
  T* const smem = reinterpret_cast<T*>(bcp_smem);
  const int lane = threadIdx.x & 31, warp_base = threadIdx.x - lane;
  T* const col = smem + (size_t)threadIdx.x * col_stride;
  const SolveLayout L = solve_layout(P);
  const int N = P.num_envs;
  PoolFeed<T> feed;  // its piece table and queue geometry; nothing is fetched through it
  feed.init(L, lane, rec_words, qc, N);
  SolveScale<T> S;
  {
    SolveRegs<T> R0;
    solve_consts(P, R0, S);
  }
  const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
  const T res_thr = P.residual_thr;
  const int max_it = P.solver_iterations;
  const int warps = gridDim.x * (blockDim.x >> 5);
  for (int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w * 32 < feed.total; w += warps) {
    const int idx = w * 32 + lane;
    const bool have = idx < feed.total;
    int env = 0, meta = 0;
    if (have) {
      const int2 q = queue[idx < feed.n_long ? idx : N - 1 - (idx - feed.n_long)];
      env = q.x;
      meta = q.y;
    }
    feed.stage(__ballot_sync(FULL, have), lane, env, meta, recs, rec_words, smem, warp_base, col_stride);
    cp_async_wait_all();
    __syncwarp();
    LamGen<T> Sy;
    int it = 0;
    if (have) {
      int cnt[3];
      SolveRegs<T> R;
      solve_load(col, S, cnt, R);  // counts and warm-start impulses
      lam_build_gen<true, FRICTION>(P, col + REC_ROWS, L, R.apn, Sy);
    }
    bool open = have;
    while (__any_sync(FULL, open)) {
      if (open) {
        const T lsr = lam_round_gen<FRICTION>(P, fr, Sy);
        it++;
        if (lsr <= res_thr || it >= max_it) open = false;
      }
    }
    if (have) {
      SolveRegs<T> R;
      R.cart.invM = P.cart.inv_mass;
      R.pole.invM = P.pole.inv_mass;
      lam_finish_gen<true, FRICTION>(P, col + REC_ROWS, L, Sy, R);
      solve_end(recs + (size_t)env * rec_words, S, it, R);
    }
    __syncwarp();
  }
}
// The solves k_solve_lam handed off, one per lane, 32 consecutive continuation slots per warp (all
// of one regime): no shared memory, no refill - every lane here needs tens of iterations, and an
// iteration in impulse space is a short dependent chain.  Rows for the final body deltas come from
// the env's solve record.
template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock, 2) k_solve_tail(const __grid_constant__ Params<T> P, T* __restrict__ recs, int rec_words,
                                                           const int* __restrict__ qc, const T* __restrict__ cont,
                                                           const int4* __restrict__ contm) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int N = P.num_envs;
  const int n_stack = qc[2 * Q_WORDS + 0], n_split = qc[2 * Q_WORDS + 1];
  const int c_stack = (n_stack + 31) >> 5, chunks = c_stack + ((n_split + 31) >> 5);
  const SolveLayout L = solve_layout(P);
  const bool fr = FRICTION && L.fbase[BCP_PAIR_GROUND_POLE] >= 0;
  SolveScale<T> S;
  {
    SolveRegs<T> R0;
    solve_consts(P, R0, S);
  }
  const T res_thr = P.residual_thr;
  const int max_it = P.solver_iterations;
  const int warps = gridDim.x * (blockDim.x >> 5);
  for (int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < chunks; c += warps) {
    const bool stack = c < c_stack;
    const int k = stack ? c * 32 + lane : (c - c_stack) * 32 + lane;
    bool busy = k < (stack ? n_stack : n_split);
    const int slot = stack ? k : N - 1 - k;
    LamSys<T> Sy;
    int it = 0, meta = 0, env = 0;
    if (busy) {
      const int4 m = contm[slot];
      env = m.x; it = m.y; meta = m.z;
      lam_fetch(cont, N, slot, Sy);
    }
    int mx[3];
#pragma unroll
    for (int m = 0; m < 3; ++m) mx[m] = __reduce_max_sync(FULL, busy ? (meta >> (8 * m)) & 0xff : 0);
    bool open = busy;
    while (__any_sync(FULL, open)) {
      if (open) {  // a lane that is done keeps its impulses untouched until the whole warp is
        // all four slots of every block, whatever the warp's counts (unused slots are zero rows):
        // straight-line code, so the cart's rows schedule between the pole's
        const T lsr = stack ? lam_round_stack(P, 4, 4, Sy) : lam_round_split<FRICTION>(P, fr, 4, 4, Sy);
        it++;
        if (lsr <= res_thr || it >= max_it) open = false;
      }
    }
    if (busy) {  // all 32 at once: the record reads (one env per lane) overlap
      T* rec = recs + (size_t)env * rec_words;
      SolveRegs<T> R;
      R.cart.invM = P.cart.inv_mass;
      R.pole.invM = P.pole.inv_mass;
      if (stack) lam_finish_stack<false>(P, rec + REC_ROWS, L, Sy, R);
      else lam_finish_split<false, FRICTION>(P, rec + REC_ROWS, L, Sy, R);
      solve_end(rec, S, it, R);
    }
  }
}
#endif  // __CUDACC__

}  // namespace bcp
