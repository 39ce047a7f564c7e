// Host build of the KERNEL physics headers (cartpoleplusplus_b200/csrc/*.cuh compiled
// by g++ with BCP_HD empty).  TEST TOOL ONLY: it lets the CPU-only test tier compare the
// exact per-env program the CUDA kernels run against the C oracle without a GPU.  It is
// not reachable from the package and is not a product fallback.
#include <cstring>
#include "../../cartpoleplusplus_b200/csrc/bcp_env.cuh"
#include "../../cartpoleplusplus_b200/csrc/bcp_state.cuh"

using namespace bcp;

template <typename T, bool F>
static void substep_all(const Params<T>& P, double* state, int n, int32_t* info_out) {
  for (int e = 0; e < n; ++e) {
    Env<T> st;
    SubstepInfo info;
    env_from_canonical<T>(state + (size_t)e * BCP_STATE_DOUBLES, st);
    substep<T, F>(P, st, &info);
    env_to_canonical<T>(st, state + (size_t)e * BCP_STATE_DOUBLES);
    if (info_out) {
      int32_t* o = info_out + (size_t)e * 6;
      o[0] = info.iterations; o[1] = info.rows; o[2] = info.friction_rows;
      o[3] = info.contacts[0]; o[4] = info.contacts[1]; o[5] = info.contacts[2];
    }
  }
}
template <typename T, bool F>
static void reset_all(const Params<T>& P, double* state, int n, const uint8_t* mask, float* obs) {
  for (int e = 0; e < n; ++e) {
    if (mask && !mask[e]) continue;
    Env<T> st;
    env_from_canonical<T>(state + (size_t)e * BCP_STATE_DOUBLES, st);
    env_reset<T, F>(P, st, P.env_id_offset + e);
    if (obs) for (int r = 0; r < P.R; ++r) write_obs(st, obs + ((size_t)e * P.R + r) * 14);
    env_to_canonical<T>(st, state + (size_t)e * BCP_STATE_DOUBLES);
  }
}
template <typename T, bool F>
static void step_all(const Params<T>& P, double* state, int n, const void* actions, float* obs, float* reward,
                     uint8_t* done, uint8_t* reason) {
  for (int e = 0; e < n; ++e) {
    Env<T> st;
    env_from_canonical<T>(state + (size_t)e * BCP_STATE_DOUBLES, st);
    if (st.done) {
      reward[e] = 0.0f; done[e] = 1; if (reason) reason[e] = 0;
      continue;
    }
    int a = 0; float ax = 0, ay = 0;
    if (P.discrete) a = ((const int32_t*)actions)[e];
    else { ax = ((const float*)actions)[2 * e]; ay = ((const float*)actions)[2 * e + 1]; }
    T fx, fy;
    action_force(P, a, ax, ay, fx, fy);
    const int rs = env_step<T, F>(P, st, fx, fy, obs ? obs + (size_t)e * P.R * 14 : nullptr);
    reward[e] = 1.0f; done[e] = (uint8_t)st.done; if (reason) reason[e] = (uint8_t)rs;
    env_to_canonical<T>(st, state + (size_t)e * BCP_STATE_DOUBLES);
  }
}

#define DISPATCH(T, call_true, call_false) \
  { const Params<T> P = make_params<T>(*cfg); if (P.any_friction) { call_true; } else { call_false; } }

extern "C" {
int hh_substep_f64(const bcp_config* cfg, double* s, int n, int32_t* info) { DISPATCH(double, (substep_all<double, true>(P, s, n, info)), (substep_all<double, false>(P, s, n, info))); return 0; }
int hh_substep_f32(const bcp_config* cfg, double* s, int n, int32_t* info) { DISPATCH(float, (substep_all<float, true>(P, s, n, info)), (substep_all<float, false>(P, s, n, info))); return 0; }
int hh_reset_f64(const bcp_config* cfg, double* s, int n, const uint8_t* m, float* obs) { DISPATCH(double, (reset_all<double, true>(P, s, n, m, obs)), (reset_all<double, false>(P, s, n, m, obs))); return 0; }
int hh_reset_f32(const bcp_config* cfg, double* s, int n, const uint8_t* m, float* obs) { DISPATCH(float, (reset_all<float, true>(P, s, n, m, obs)), (reset_all<float, false>(P, s, n, m, obs))); return 0; }
int hh_step_f64(const bcp_config* cfg, double* s, int n, const void* a, float* obs, float* rw, uint8_t* dn, uint8_t* rs) { DISPATCH(double, (step_all<double, true>(P, s, n, a, obs, rw, dn, rs)), (step_all<double, false>(P, s, n, a, obs, rw, dn, rs))); return 0; }
int hh_step_f32(const bcp_config* cfg, double* s, int n, const void* a, float* obs, float* rw, uint8_t* dn, uint8_t* rs) { DISPATCH(float, (step_all<float, true>(P, s, n, a, obs, rw, dn, rs)), (step_all<float, false>(P, s, n, a, obs, rw, dn, rs))); return 0; }
size_t hh_config_size(void) { return sizeof(bcp_config); }
}
