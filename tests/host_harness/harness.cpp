// Host build of the KERNEL physics headers (cartpoleplusplus_b200/csrc/*.cuh compiled
// by g++ with BCP_HD empty).  TEST TOOL ONLY: it lets the CPU-only test tier compare the
// exact per-env program the CUDA kernels run against the C oracle without a GPU.  It is
// not reachable from the package and is not a product fallback.
#include <cstring>
#include <vector>
#include "../../cartpoleplusplus_b200/csrc/bcp_split.cuh"
#include "../../cartpoleplusplus_b200/csrc/bcp_state.cuh"

using namespace bcp;

// path selector of the harness: 0 = parity substep<>, 1 = fast substep_fast<> (its shared-memory
// column is a plain host array here)
static int g_fast = 0;
template <typename T, bool F> struct HostStep {
  mutable std::vector<T> buf;
  int fast;
  mutable std::vector<T> rec;
  explicit HostStep(const Params<T>& P) : buf((size_t)P.col_words + 1), fast(g_fast), rec((size_t)record_words(P), T(0)) {}
  void operator()(const Params<T>& P, Env<T>& st, bool on) const { if (on) run(P, st, nullptr); }
  void run(const Params<T>& P, Env<T>& st, SubstepInfo* info) const {
    if (fast == 2) {  // the split pipeline's per-env pieces, composed
      substep_split_reference<T, F>(P, st, rec.data(), info);
    } else if (fast) {
      Col<T> col{buf.data(), 1};
      substep_fast<T, F>(P, st, col, true, info);
    } else {
      substep<T, F>(P, st, info);
    }
  }
};

template <typename T, bool F>
static void substep_all(const Params<T>& P, double* state, int n, int32_t* info_out) {
  const HostStep<T, F> sub(P);
  for (int e = 0; e < n; ++e) {
    Env<T> st;
    SubstepInfo info;
    env_from_canonical<T>(state + (size_t)e * BCP_STATE_DOUBLES, st);
    sub.run(P, st, &info);
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
  const HostStep<T, F> sub(P);
  for (int e = 0; e < n; ++e) {
    if (mask && !mask[e]) continue;
    Env<T> st;
    env_from_canonical<T>(state + (size_t)e * BCP_STATE_DOUBLES, st);
    env_reset(P, st, P.env_id_offset + e, sub, true);
    if (obs) for (int r = 0; r < P.R; ++r) write_obs(st, obs + ((size_t)e * P.R + r) * 14);
    env_to_canonical<T>(st, state + (size_t)e * BCP_STATE_DOUBLES);
  }
}
template <typename T, bool F>
static void step_all(const Params<T>& P, double* state, int n, const void* actions, float* obs, float* reward,
                     uint8_t* done, uint8_t* reason) {
  const HostStep<T, F> sub(P);
  for (int e = 0; e < n; ++e) {
    Env<T> st;
    env_from_canonical<T>(state + (size_t)e * BCP_STATE_DOUBLES, st);
    int a = 0; float ax = 0, ay = 0;
    if (P.discrete) a = ((const int32_t*)actions)[e];
    else { ax = ((const float*)actions)[2 * e]; ay = ((const float*)actions)[2 * e + 1]; }
    T fx, fy;
    action_force(P, a, ax, ay, fx, fy);
    float* o = obs ? obs + (size_t)e * P.R * 14 : nullptr;
    bool live = true;
    const int rs = env_call(P, st, P.env_id_offset + e, fx, fy, o, sub, true, live);
    reward[e] = live ? 1.0f : 0.0f; done[e] = (uint8_t)st.done; if (reason) reason[e] = (uint8_t)rs;
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
void hh_set_fast(int on) { g_fast = on; }
int hh_col_words(const bcp_config* cfg) { return make_params<float>(*cfg).col_words; }
}
