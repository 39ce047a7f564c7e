// bcp_types.cuh — per-env state record and the scalar parameter block of the kernels.
#pragma once
#include "../../include/bcp.h"
#include <cmath>
#include "bcp_math.cuh"

namespace bcp {

// One dynamic box: pose (x, q = body->world, xyzw) and world-frame velocities.
template <typename T> struct Body {
  T x[3], q[4], v[3], w[3];
};
// btManifoldPoint subset that outlives a step.
template <typename T> struct MPoint {
  T la[3], lb[3], n[3], dist, imp;
};
// btPersistentManifold (<= 4 points) of one collision pair.
template <typename T> struct Manifold {
  int count;
  MPoint<T> p[4];
};
// Everything one environment carries between substeps.
template <typename T> struct Env {
  Body<T> cart, pole;
  T pend_f[3], pend_t[3];  // external force/torque waiting for the next stepSimulation
  Manifold<T> mani[3];     // BCP_PAIR_* order
  int steps, done, episode;
  int reset_left;  // pipelined auto-reset: reset substeps still to run (0 = not resetting)
};

template <typename T> struct BoxP {
  T mass, inv_mass, I[3], h[3], x0[3];
};

// bcp_config converted once on the host into the kernel's scalar type.
template <typename T> struct Params {
  BoxP<T> ground, cart, pole;
  T dt, gravity[3];
  T erp2, linear_slop, cfm_normal, cfm_friction, friction_erp;
  T restitution_thr, kl, ka, vmax, residual_thr, warmstart;
  T breaking[3], mu[3], rest[3];  // per pair
  T action_force, initial_force;
  float action_force_f32;
  T pos_thr, sin_thr, tan_thr;
  int solver_iterations, friction_dirs, cone_friction, manifold_order, gyro_term;
  int discrete, R, S, max_len, push_steps, settle_steps;
  int random_theta, theta_mode, force_frame, termination, carry_over, auto_reset;
  int num_envs;
  int64_t env_id_offset;
  uint64_t seed;
  bool any_friction;
  // fast-mode row column layout (bcp_fast.cuh): per pair, rows per contact (1 normal +
  // friction rows when the pair has friction) and the first word of the pair's rows
  int rows_per_contact[3], row_base[3], col_words;
  // --no-carry-over-reset with the settle cache (bcp_config.reset_settle_cache): the state every env is in after the
  // teleport and the 100 settle substeps - identical for all of them when reset() clears the pending force and the contact
  // caches - computed once per handle: one state record (layout of bcp_state.cuh: bodies, pending force, 3 x 4 manifold
  // points) and its three point counts, in device memory.  nullptr: every reset runs its settle substeps.
  const T* settled;
  const int* settled_cnt;
};

// ---- host-side conversion (plain host functions; never called from device code) ----
template <typename T> inline void fill_box(BoxP<T>& b, const bcp_body& s) {
  b.mass = (T)s.mass;
  b.inv_mass = (s.mass != 0.0) ? T(1.0) / (T)s.mass : T(0.0);  // computed in T, like the oracle
  for (int i = 0; i < 3; ++i) { b.I[i] = (T)s.inertia[i]; b.h[i] = (T)s.half_extents[i]; b.x0[i] = (T)s.init_pos[i]; }
}
inline double shape_breaking(const bcp_config& c, const bcp_body& b) {
  const double* h = b.half_extents;
  return c.contact_breaking_factor * (sqrt(4 * h[0] * h[0] + 4 * h[1] * h[1] + 4 * h[2] * h[2]) * 0.5);
}
template <typename T> inline Params<T> make_params(const bcp_config& c) {
  Params<T> p;
  fill_box(p.ground, c.ground);
  fill_box(p.cart, c.cart);
  fill_box(p.pole, c.pole);
  p.dt = (T)c.dt;
  for (int i = 0; i < 3; ++i) p.gravity[i] = (T)c.gravity[i];
  const T inv_dt = T(1.0) / p.dt;
  p.erp2 = (T)c.erp2;
  p.linear_slop = (T)c.linear_slop;
  p.cfm_normal = (T)c.global_cfm * inv_dt;
  p.cfm_friction = (T)c.friction_cfm * inv_dt;
  p.friction_erp = (T)c.friction_erp;
  p.restitution_thr = (T)c.restitution_velocity_threshold;
  p.kl = (T)c.linear_damping;
  p.ka = (T)c.angular_damping;
  p.vmax = (T)c.max_coordinate_velocity;
  p.residual_thr = (T)c.residual_threshold;
  p.warmstart = (T)c.warmstart_factor;
  const bcp_body* A[3] = {&c.ground, &c.cart, &c.ground};
  const bcp_body* B[3] = {&c.cart, &c.pole, &c.pole};
  p.any_friction = false;
  for (int m = 0; m < 3; ++m) {
    const double ta = shape_breaking(c, *A[m]), tb = shape_breaking(c, *B[m]);
    p.breaking[m] = (T)(ta < tb ? ta : tb);
    // the cart's friction against the ground may differ from its friction against the pole (compound cart, bcp.h)
    const double fb = (m == BCP_PAIR_GROUND_CART && c.cart_ground_friction >= 0.0) ? c.cart_ground_friction : B[m]->friction;
    double f = A[m]->friction * fb;
    if (f < -c.max_friction) f = -c.max_friction;
    if (f > c.max_friction) f = c.max_friction;
    p.mu[m] = (T)f;
    p.rest[m] = (T)(A[m]->restitution * B[m]->restitution);
    if (p.mu[m] != T(0.0)) p.any_friction = true;
  }
  p.action_force = (T)c.action_force;
  p.action_force_f32 = (float)c.action_force;
  p.initial_force = (T)c.initial_force;
  p.pos_thr = (T)c.pos_threshold;
  p.sin_thr = (T)c.sin_angle_threshold;
  p.tan_thr = (T)c.tan_angle_threshold;
  p.solver_iterations = c.solver_iterations;
  p.friction_dirs = c.friction_dirs;
  p.cone_friction = c.cone_friction;
  p.manifold_order = c.manifold_order;
  p.gyro_term = c.gyro_term;
  p.discrete = c.discrete_actions;
  p.R = c.action_repeats;
  p.S = c.steps_per_repeat;
  p.max_len = c.max_episode_len;
  p.push_steps = c.initial_force_steps;
  p.settle_steps = c.settle_steps;
  p.random_theta = c.random_theta;
  p.theta_mode = c.theta_mode;
  p.force_frame = c.force_frame;
  p.termination = c.termination_enabled;
  p.carry_over = c.carry_over_reset;
  p.auto_reset = c.auto_reset;
  p.num_envs = c.num_envs;
  p.env_id_offset = c.env_id_offset;
  p.seed = c.seed;
  p.settled = nullptr;
  p.settled_cnt = nullptr;
  p.col_words = 0;
  for (int m = 0; m < 3; ++m) {
    p.rows_per_contact[m] = (p.mu[m] != T(0.0)) ? 1 + c.friction_dirs : 1;
    p.row_base[m] = p.col_words;
    p.col_words += 4 * p.rows_per_contact[m] * (m == BCP_PAIR_CART_POLE ? 13 : 10);
  }
  return p;
}

}  // namespace bcp
