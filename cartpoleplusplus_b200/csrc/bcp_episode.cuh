// bcp_episode.cuh — BulletCartpole.reset / .step bookkeeping (included by bcp_env.cuh).
#pragma once

namespace bcp {

// teleport of reset(): p.resetBasePositionAndOrientation x2, bullet_cartpole.py:319-320
// (zeroes the base velocities; the pair caches and the pending force survive).
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
}

// the k-th of the 30 initial pushes: bump_cart -> random_force_in_plane, :348-359
template <typename T> BCP_HD void push_force(const Params<T>& P, int64_t genv, int episode, int k, T& ts, T& tc) {
  if (!P.random_theta) { ts = 0; tc = 1; return; }
  const uint32_t idx = (P.theta_mode == BCP_THETA_PER_SUBSTEP) ? (uint32_t)k : 0u;
  const U4 r = env_rng(P.seed, genv, (uint32_t)episode, RNG_PUSH, idx);
  sincos_2pi(u24<T>(r.v[0]), ts, tc);
}

// BulletCartpole.reset - bullet_cartpole.py:313-346
template <typename T, bool FRICTION>
BCP_HD void env_reset(const Params<T>& P, Env<T>& st, int64_t genv) {
  reset_begin(P, st);
  for (int k = 0; k < P.settle_steps; ++k) substep<T, FRICTION>(P, st, nullptr);  // :326-327
  T ts = 0, tc = 1;
  if (P.theta_mode == BCP_THETA_PER_RESET) push_force(P, genv, st.episode, 0, ts, tc);
  for (int k = 0; k < P.push_steps; ++k) {  // :329-332
    substep<T, FRICTION>(P, st, nullptr);
    if (P.theta_mode == BCP_THETA_PER_SUBSTEP) push_force(P, genv, st.episode, k, ts, tc);
    apply_cart_force(P, st, P.initial_force * tc, P.initial_force * ts);
  }
  st.episode += 1;
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

// BulletCartpole.step for a live env - bullet_cartpole.py:197-260.  `obs` points at
// this env's (R,2,7) block with element stride `obs_stride` floats between
// consecutive values (1 for a dense block; lets kernels stage through shared memory).
template <typename T, bool FRICTION>
BCP_HD int env_step(const Params<T>& P, Env<T>& st, T fx, T fy, float* obs) {
  for (int r = 0; r < P.R; ++r) {    // :197
    for (int s = 0; s < P.S; ++s) {  // :198
      substep<T, FRICTION>(P, st, nullptr);  // :199
      apply_cart_force(P, st, fx, fy);       // :201-203
    }
    if (obs) write_obs(st, obs + 14 * r);    // :237
  }
  st.steps += 1;  // :239
  const int reason = done_reason(P, st);
  if (reason != BCP_REASON_NONE) st.done = 1;
  return reason;
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
