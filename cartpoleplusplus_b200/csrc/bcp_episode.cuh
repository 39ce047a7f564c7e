// bcp_episode.cuh — BulletCartpole.reset / .step bookkeeping (included by bcp_env.cuh).
#pragma once

namespace bcp {

// teleport of reset(): p.resetBasePositionAndOrientation x2, bullet_cartpole.py:319-320
// (zeroes the base velocities; the pair caches and the pending force survive).
// With the settle cache (Params::settled; only in the --no-carry-over-reset mode, where the 100 settle substeps start from
// the same state in every env and are therefore the same computation) the env takes the cached post-settle state instead:
// bodies and point counts here, the points themselves in settled_points() (the split pipeline copies the blocks in HBM).
template <typename T> BCP_HD void reset_begin(const Params<T>& P, Env<T>& st) {
  st.steps = 0;  // :315
  st.done = 0;   // :316
  for (int i = 0; i < 3; ++i) {
    st.cart.x[i] = P.cart.x0[i];
    st.pole.x[i] = P.pole.x0[i];
    st.cart.v[i] = st.cart.w[i] = st.pole.v[i] = st.pole.w[i] = 0;
    st.cart.q[i] = st.pole.q[i] = 0;
  }
  st.cart.q[3] = st.pole.q[3] = 1;
  if (!P.carry_over) {
    for (int i = 0; i < 3; ++i) { st.pend_f[i] = st.pend_t[i] = 0; st.mani[i].count = 0; }
  }
  if (P.settled) {
    const T* c = P.settled + BCP_S_CART;
    const T* q = P.settled + BCP_S_POLE;
    for (int i = 0; i < 3; ++i) {
      st.cart.x[i] = c[i]; st.cart.v[i] = c[7 + i]; st.cart.w[i] = c[10 + i];
      st.pole.x[i] = q[i]; st.pole.v[i] = q[7 + i]; st.pole.w[i] = q[10 + i];
      st.pend_f[i] = P.settled[BCP_S_PEND_F + i]; st.pend_t[i] = P.settled[BCP_S_PEND_T + i];
      st.mani[i].count = P.settled_cnt[i];
    }
    for (int i = 0; i < 4; ++i) { st.cart.q[i] = c[3 + i]; st.pole.q[i] = q[3 + i]; }
  }
}
template <typename T> BCP_HD void settled_points(const Params<T>& P, Env<T>& st) {
  if (!P.settled) return;
  for (int m = 0; m < 3; ++m)
    for (int k = 0; k < 4; ++k) {
      const T* q = P.settled + BCP_S_MANI + m * BCP_MANI_STRIDE + k * BCP_MANI_POINT;
      MPoint<T>& p = st.mani[m].p[k];
      for (int i = 0; i < 3; ++i) { p.la[i] = q[i]; p.lb[i] = q[3 + i]; p.n[i] = q[6 + i]; }
      p.dist = q[9];
      p.imp = q[10];
    }
}
// substeps of a reset that are actually run: the settle loop is skipped when its result is cached
template <typename T> BCP_HD int reset_settle_steps(const Params<T>& P) { return P.settled ? 0 : P.settle_steps; }

// the k-th of the 30 initial pushes: bump_cart -> random_force_in_plane, :348-359
template <typename T> BCP_HD void push_force(const Params<T>& P, int64_t genv, int episode, int k, T& ts, T& tc) {
  if (!P.random_theta) { ts = 0; tc = 1; return; }
  const uint32_t idx = (P.theta_mode == BCP_THETA_PER_SUBSTEP) ? (uint32_t)k : 0u;
  const U4 r = env_rng(P.seed, genv, (uint32_t)episode, RNG_PUSH, idx);
  sincos_2pi(u24<T>(r.v[0]), ts, tc);
}

// The substep a per-env program runs: StepParity (bcp_env.cuh substep<>, bit-exact with the
// oracle in fp64) or StepFast (bcp_fast.cuh).  Programs below are templated on it.
// Every program is entered by ALL lanes of a warp and calls the substep the same number of
// times in each of them (`on` == false: the lane has no work; the fast path needs the whole
// warp at its votes).
template <typename T, bool FRICTION> struct StepParity {
  BCP_HD void operator()(const Params<T>& P, Env<T>& st, bool on) const {
    if (on) substep<T, FRICTION>(P, st, nullptr);
  }
};

// BulletCartpole.reset - bullet_cartpole.py:313-346
template <typename T, typename STEP>
BCP_HD void env_reset(const Params<T>& P, Env<T>& st, int64_t genv, const STEP& sub, bool on) {
  if (on) {
    reset_begin(P, st);
    settled_points(P, st);
    st.reset_left = 0;
  }
  for (int k = 0; k < reset_settle_steps(P); ++k) sub(P, st, on);  // :326-327
  T ts = 0, tc = 1;
  if (on && P.theta_mode == BCP_THETA_PER_RESET) push_force(P, genv, st.episode, 0, ts, tc);
  for (int k = 0; k < P.push_steps; ++k) {  // :329-332
    sub(P, st, on);
    if (on) {
      if (P.theta_mode == BCP_THETA_PER_SUBSTEP) push_force(P, genv, st.episode, k, ts, tc);
      apply_cart_force(P, st, P.initial_force * tc, P.initial_force * ts);
    }
  }
  if (on) st.episode += 1;
}

// action -> (fx, fy): discrete table (documented at :84,89; README.md:58) or
// continuous `action[0] * action_force` in float32 (:201)
template <typename T> BCP_HD void action_force(const Params<T>& P, int a_disc, float ax, float ay, T& fx, T& fy) {
  if (P.discrete) {
    fx = 0; fy = 0;
    if (a_disc == 1) fx = P.action_force;
    else if (a_disc == 2) fx = -P.action_force;
    else if (a_disc == 3) fy = P.action_force;
    else if (a_disc == 4) fy = -P.action_force;
  } else {
    fx = (T)(ax * P.action_force_f32);
    fy = (T)(ay * P.action_force_f32);
  }
}

// ---- one call of an env, in pieces ------------------------------------------------------
// BulletCartpole.step (bullet_cartpole.py:178-275) and the pipelined auto-reset machine
// (BCP_AUTO_RESET_PIPELINED) share one shape: decide what the env does during this call,
// run R*S substeps with a little work after each, close the call.  The pieces are used by
// the one-launch programs below and, one kernel per piece, by the split pipeline
// (bcp_split.cuh).
//   live env      : [stepSimulation; applyExternalForce] x S, pose capture, x R  (reward 1)
//   done env      : reference: nothing (:179-181).  Pipelined: enters its reset - teleport
//                   (:319-320) now, then the 100 settle and 30 push substeps (:326-332) R*S per
//                   call, so that a warp never waits on one lane's reset - and reports RESETTING
//   reset finishes: observation bootstrap (:338-339) at the end of that call, FRESH
enum { MODE_IDLE = 0, MODE_LIVE = 1, MODE_RESET = 2, MODE_FRESH = 3 };

// points: the caller holds the manifold points in `st` (one-launch programs); the split pipeline keeps them in HBM and copies
// the cached blocks itself when a reset begins (k_prep)
template <typename T> BCP_HD int call_begin(const Params<T>& P, Env<T>& st, bool valid, bool points = true) {
  if (!valid) return MODE_IDLE;
  if (!st.done) return MODE_LIVE;
  if (P.auto_reset != BCP_AUTO_RESET_PIPELINED) return MODE_IDLE;
  if (st.reset_left == 0) {  // episode ended in an earlier call: the reset starts
    reset_begin(P, st);
    if (points) settled_points(P, st);
    st.done = 1;
    st.reset_left = reset_settle_steps(P) + P.push_steps;
    if (st.reset_left == 0) { st.done = 0; st.episode += 1; return MODE_FRESH; }
  }
  return MODE_RESET;
}
template <typename T> BCP_HD bool substep_on(int mode, const Env<T>& st) {
  return mode == MODE_LIVE || (mode == MODE_RESET && st.reset_left > 0);
}
// after each p.stepSimulation(): the action force (:201-203) or the reset's push (:329-332)
template <typename T> BCP_HD void post_substep(const Params<T>& P, Env<T>& st, int64_t genv, int& mode, T fx, T fy) {
  if (mode == MODE_LIVE) {
    apply_cart_force(P, st, fx, fy);
  } else if (mode == MODE_RESET && st.reset_left > 0) {
    const int k = P.push_steps - st.reset_left;  // push index; < 0 while settling (reset_left counts settle + push substeps still to run)
    if (k >= 0) {
      T ts, tc;
      push_force(P, genv, st.episode, (P.theta_mode == BCP_THETA_PER_SUBSTEP) ? k : 0, ts, tc);
      apply_cart_force(P, st, P.initial_force * tc, P.initial_force * ts);
    }
    if (--st.reset_left == 0) { st.done = 0; st.episode += 1; mode = MODE_FRESH; }
  }
}
// steps / done / reward bookkeeping (:239-260); returns the reason code, `live` = reward 1
// rep_stride: floats between two repeats of one env in `obs` (14 in the (B, R, 2, 7) block; 14 * B in the repeat-major slabs
// bcp_step_host sends to the host)
template <typename T> BCP_HD int call_end(const Params<T>& P, Env<T>& st, int mode, float* obs, bool& live, size_t rep_stride = 14) {
  live = mode == MODE_LIVE;
  if (mode == MODE_LIVE) {
    st.steps += 1;  // :239
    const int reason = done_reason(P, st);
    if (reason != BCP_REASON_NONE) st.done = 1;
    return reason;
  }
  if (mode == MODE_FRESH) {
    if (obs) for (int r = 0; r < P.R; ++r) write_obs(st, obs + rep_stride * r);
    return BCP_REASON_FRESH;
  }
  return mode == MODE_RESET ? BCP_REASON_RESETTING : BCP_REASON_NONE;
}

// The whole call in one go.  `obs` points at this env's dense (R,2,7) block (or nullptr).
template <typename T, typename STEP>
BCP_HD int env_call(const Params<T>& P, Env<T>& st, int64_t genv, T fx, T fy, float* obs, const STEP& sub,
                    bool valid, bool& live) {
  int mode = call_begin(P, st, valid);
  for (int r = 0; r < P.R; ++r) {    // :197
    for (int s = 0; s < P.S; ++s) {  // :198
      sub(P, st, substep_on(mode, st));  // :199
      post_substep(P, st, genv, mode, fx, fy);
    }
    if (mode == MODE_LIVE && obs) write_obs(st, obs + 14 * r);  // :237
  }
  return call_end(P, st, mode, obs, live);
}

// random-agent action draw (README.md:76-93 protocols), Philox keyed by global env id
template <typename T>
BCP_HD void draw_action(const Params<T>& P, int64_t genv, const Env<T>& st, uint32_t mask, int& a_disc, float& ax, float& ay) {
  const U4 r = env_rng(P.seed, genv, (uint32_t)st.episode, RNG_ACTION, (uint32_t)st.steps);
  a_disc = 0; ax = 0; ay = 0;
  if (P.discrete) {
    int n = 0;
    for (int i = 0; i < 5; ++i) n += (mask >> i) & 1u;
    const int pick = (int)(((uint64_t)r.v[0] * (uint64_t)n) >> 32);
    int seen = 0;
    for (int i = 0; i < 5; ++i)
      if ((mask >> i) & 1u) {
        if (seen == pick) { a_disc = i; break; }
        seen++;
      }
  } else {
    ax = 2.0f * ((float)(r.v[0] >> 8) * 5.9604644775390625e-08f) - 1.0f;
    ay = 2.0f * ((float)(r.v[1] >> 8) * 5.9604644775390625e-08f) - 1.0f;
  }
}

}  // namespace bcp
