// Flop census of one p.stepSimulation() as the parity path states it (cartpoleplusplus_b200/csrc/bcp_env.cuh substep<>,
// the operation sequence that is bit-identical to the oracle's).  TEST / MEASUREMENT TOOL: the substep is instantiated
// with a scalar that counts every arithmetic operation it performs, attributed to the phase census_phase<> announces.
// Convention: add, subtract, multiply, divide, square root = 1 flop each (so an FMA is 2); comparisons, negation, abs,
// selects and moves = 0.  These are the ALGORITHM's flops - what a substep needs - not the instructions of any kernel.
#include <cmath>
#include <cstdint>
#include <cstring>
#include "../../cartpoleplusplus_b200/csrc/bcp_math.cuh"

namespace bcp {
static thread_local long long g_ops[CENSUS_PHASES];
static thread_local int g_phase = 0;
struct Cnt {
  double v;
  Cnt() = default;
  Cnt(double x) : v(x) {}
  Cnt(float x) : v(x) {}
  Cnt(int x) : v(x) {}
  Cnt(unsigned x) : v(x) {}
  explicit operator double() const { return v; }
  explicit operator float() const { return (float)v; }
  explicit operator int() const { return (int)v; }
};
#define CNT_BIN(op)                                                                          \
  inline Cnt operator op(Cnt a, Cnt b) { g_ops[g_phase]++; return Cnt(a.v op b.v); }        \
  inline Cnt& operator op##=(Cnt& a, Cnt b) { g_ops[g_phase]++; a.v op## = b.v; return a; }
CNT_BIN(+) CNT_BIN(-) CNT_BIN(*) CNT_BIN(/)
inline Cnt operator-(Cnt a) { return Cnt(-a.v); }
#define CNT_CMP(op) inline bool operator op(Cnt a, Cnt b) { return a.v op b.v; }
CNT_CMP(<) CNT_CMP(>) CNT_CMP(<=) CNT_CMP(>=) CNT_CMP(==) CNT_CMP(!=)
inline Cnt bsqrt(Cnt x) { g_ops[g_phase]++; return Cnt(std::sqrt(x.v)); }
inline Cnt babs(Cnt x) { return Cnt(std::fabs(x.v)); }
template <> struct Lim<Cnt> {
  static Cnt eps() { return Cnt(2.220446049250313e-16); }
  static Cnt large() { return Cnt(1e18); }
};
template <> inline void census_phase<Cnt>(int p) { g_phase = p; }
}  // namespace bcp

#include "../../cartpoleplusplus_b200/csrc/bcp_env.cuh"
#include "../../cartpoleplusplus_b200/csrc/bcp_state.cuh"

using namespace bcp;

template <bool F>
static void run(const Params<Cnt>& P, const double* state, int n, double* state_out, int64_t* out /* [n][PHASES + 4] */) {
  for (int e = 0; e < n; ++e) {
    Env<Cnt> st;
    SubstepInfo info;
    env_from_canonical<Cnt>(state + (size_t)e * BCP_STATE_DOUBLES, st);
    std::memset(g_ops, 0, sizeof(g_ops));
    substep<Cnt, F>(P, st, &info);
    int64_t* o = out + (size_t)e * (CENSUS_PHASES + 4);
    for (int k = 0; k < CENSUS_PHASES; ++k) o[k] = g_ops[k];
    o[CENSUS_PHASES] = info.contacts[0]; o[CENSUS_PHASES + 1] = info.contacts[1]; o[CENSUS_PHASES + 2] = info.contacts[2];
    o[CENSUS_PHASES + 3] = info.iterations;
    if (state_out) env_to_canonical<Cnt>(st, state_out + (size_t)e * BCP_STATE_DOUBLES);
  }
}

extern "C" int fc_substep(const bcp_config* cfg, const double* state, int n, double* state_out, int64_t* out) {
  const Params<Cnt> P = make_params<Cnt>(*cfg);
  if (P.any_friction) run<true>(P, state, n, state_out, out);
  else run<false>(P, state, n, state_out, out);
  return CENSUS_PHASES;
}
