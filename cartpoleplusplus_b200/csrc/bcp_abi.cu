// bcp_abi.cu — the extern "C" boundary of include/bcp.h over the CUDA kernels.
// No CPU implementation exists behind these entry points: without a CUDA device
// bcp_create fails and nothing else can be called.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/bcp.h"
#include "../../include/bcp_lqr.h"
#include "bcp_launch.cuh"
#include "bcp_state.cuh"

namespace {

thread_local std::string g_err;

int fail(const char* what, cudaError_t e = cudaSuccess) {
  g_err = what;
  if (e != cudaSuccess) {
    g_err += ": ";
    g_err += cudaGetErrorString(e);
  }
  return -1;
}
#define CK(call)                                  \
  do {                                            \
    cudaError_t _e = (call);                      \
    if (_e != cudaSuccess) return fail(#call, _e); \
  } while (0)

}  // namespace

namespace bcp {
int abi_fail(const char* what, cudaError_t e) { return fail(what, e); }  // the other translation units' way to bcp_last_error()
}

struct bcp_env {
  bcp_config cfg;
  bcp::Params<float> p32;
  bcp::Params<double> p64;
  void* fdata = nullptr;  // env records, [N][NF] in the handle's scalar type (bcp_state.cuh)
  int* idata = nullptr;   // [N][NI]
  unsigned long long* hist = nullptr;
  unsigned long long* header = nullptr;
  double* reward_sum = nullptr;
  int* list = nullptr;   // compacted env indices for masked resets
  int* count = nullptr;
  double* staging = nullptr;  // canonical records for get/set_state
  void* settled_f = nullptr;  // reset_settle_cache: the post-settle state record (Params::settled) ...
  int* settled_i = nullptr;   // ... and its integer fields (the three point counts first)
  // Split pipeline.  The envs of a handle run as `cohorts` independent pipelines over disjoint env ranges, each on its
  // own stream: k_prep and the solve phase are both latency-bound, and one cohort's solve tail (a few warps iterating)
  // leaves most issue slots to the other cohort's k_prep.  Every per-env array is env-major, so a cohort's state,
  // records and outputs are offset views of the handle's; queue, counters, bins and continuation buffers are its own.
  struct Cohort {
    int base = 0, n = 0;
    bcp::Params<float> P;
    bcp::SplitBuffers b = {nullptr, nullptr, nullptr, {nullptr, nullptr}, {nullptr, nullptr}, -1, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, -1, 0};
    cudaStream_t stream = nullptr;   // cohort 0 runs on the caller's stream
    cudaEvent_t ev_done = nullptr;
    std::vector<cudaEvent_t> ev_prep;  // bcp_step_host: after the k-th k_prep of the call
  };
  std::vector<Cohort> cohorts;
  cudaEvent_t ev_fork = nullptr;
  float* recs = nullptr;   // solve records of every env, [N][record words]
  int* trace = nullptr;    // bcp_solve_trace_*: [R*S][N][4]
  int use_split = 0;
  int use_lam = 0;
  // CUDA graphs of the plain bcp_step (no hook, no profiling, no trace): one per (output pointers, visiting-order parity).
  // The actions are copied into act_buf first, so the caller's action pointer is not part of the key.
  struct StepGraph {
    float* obs; float* reward; uint8_t* done; uint8_t* reason; int parity; cudaStream_t stream;
    cudaGraphExec_t exec; int launches;
  };
  std::vector<StepGraph> graphs;
  struct HostGraph {
    const void* act; float* obs; float* reward; uint8_t* done; uint8_t* reason; int parity;
    cudaGraphExec_t exec; int launches;
  };
  std::vector<HostGraph> host_graphs;
  int use_graph = 1;
  void* act_buf = nullptr;
  cudaStream_t gstream = nullptr;      // graphs are captured and launched here (the caller's stream may be the legacy default
  cudaEvent_t ev_gin = nullptr, ev_gout = nullptr;  // stream, which cannot capture), forked from / joined into the caller's
  // bcp_step_host's hook into the cohorts' launch sequences
  struct HookArg { bcp_env* h; int cohort; };
  void (*split_hook)(void* ctx, int k, cudaStream_t s) = nullptr;
  std::vector<HookArg> hook_args;
  // bcp_step_host: device-side actions / outputs / packed observation slabs, the copy stream
  struct HostStep {
    void* d_act = nullptr;
    unsigned char* d_out = nullptr;  // obs f32[N][R][14] | reward f32[N] | done u8[N] | reason u8[N]
    float* d_slab = nullptr;         // [R][N][14]
    cudaStream_t copy = nullptr;
    cudaEvent_t ev_ready = nullptr, ev_copied = nullptr, ev_h2d = nullptr;
    bool h2d_pending = false;        // the actions travel on the copy stream beside the first k_prep (which does not read them)
    float* h_obs = nullptr;          // this call's host slabs
    bool fired = false;
  } hs;
  bool profiling = false;
  std::vector<cudaEvent_t> prof_events;  // bracketing every kernel of bcp_step while profiling
  std::vector<int> prof_calls;           // events per bcp_step call
  int64_t launches = 0;
  bool f64() const { return cfg.dtype == BCP_DTYPE_F64; }
  bcp::StatsPtrs stats() const { return bcp::StatsPtrs{hist, header, reward_sum}; }
};

namespace {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

int validate(const bcp_config* c) {
  if (!c) return fail("bcp_create: null config");
  if (c->abi_version != BCP_ABI_VERSION) return fail("bcp_create: abi_version mismatch");
  if (c->num_envs < 1) return fail("bcp_create: num_envs must be >= 1");
  if (c->dtype != BCP_DTYPE_F32 && c->dtype != BCP_DTYPE_F64) return fail("bcp_create: bad dtype");
  if (c->action_repeats < 1 || c->steps_per_repeat < 1 || c->max_episode_len < 1)
    return fail("bcp_create: action_repeats, steps_per_repeat, max_episode_len must be >= 1");
  if (c->auto_reset != BCP_AUTO_RESET_OFF && c->auto_reset != BCP_AUTO_RESET_PIPELINED) return fail("bcp_create: bad auto_reset");
  if (c->friction_dirs != 1 && c->friction_dirs != 2) return fail("bcp_create: friction_dirs must be 1 or 2");
  if (c->solver_iterations < 1) return fail("bcp_create: solver_iterations must be >= 1");
  if (!(c->dt > 0)) return fail("bcp_create: dt must be > 0");
  if (!(c->cart.mass > 0) || !(c->pole.mass > 0)) return fail("bcp_create: cart and pole need positive mass");
  if (c->reset_settle_cache && c->carry_over_reset)
    return fail("bcp_create: reset_settle_cache needs carry_over_reset = 0 (with the pending force and the contact caches carried over, "
                "every env settles from a different state)");
  return 0;
}

}  // namespace

extern "C" {

int bcp_abi_version(void) { return BCP_ABI_VERSION; }
size_t bcp_config_size(void) { return sizeof(bcp_config); }
const char* bcp_last_error(void) { return g_err.c_str(); }

int bcp_create(const bcp_config* cfg, bcp_env** out) {
  if (!out) return fail("bcp_create: null out");
  *out = nullptr;
  if (validate(cfg)) return -1;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) return fail("bcp_create: no CUDA device (this library has no CPU path)", e);
  if (cfg->device_ordinal < 0 || cfg->device_ordinal >= ndev) return fail("bcp_create: device_ordinal out of range");
  bcp_env* h = new (std::nothrow) bcp_env();
  if (!h) return fail("bcp_create: out of host memory");
  h->cfg = *cfg;
  h->p32 = bcp::make_params<float>(*cfg);
  h->p64 = bcp::make_params<double>(*cfg);
  DeviceGuard g(cfg->device_ordinal);
  const size_t N = (size_t)cfg->num_envs;
  const size_t fbytes = h->f64() ? bcp::Launch<double>::state_bytes(cfg->num_envs) : bcp::Launch<float>::state_bytes(cfg->num_envs);
  const size_t nb = (size_t)cfg->max_episode_len + 1;
  cudaError_t err = cudaSuccess;
  auto A = [&](void** p, size_t bytes) {
    if (err == cudaSuccess) err = cudaMalloc(p, bytes);
    if (err == cudaSuccess) err = cudaMemset(*p, 0, bytes);
  };
  A(&h->fdata, fbytes);
  A((void**)&h->idata, (size_t)bcp::NI * N * sizeof(int));
  A((void**)&h->hist, nb * sizeof(unsigned long long));
  A((void**)&h->header, BCP_STATS_HEADER * sizeof(unsigned long long));
  A((void**)&h->reward_sum, sizeof(double));
  A((void**)&h->list, N * sizeof(int));
  A((void**)&h->count, sizeof(int));
  if (err == cudaSuccess) {  // no state until reset (bullet_cartpole.py:147-148): every env starts "done"
    std::vector<int> init(N * bcp::NI, 0);
    for (size_t e = 0; e < N; ++e) init[e * bcp::NI + bcp::I_DONE] = 1;
    err = cudaMemcpy(h->idata, init.data(), init.size() * sizeof(int), cudaMemcpyHostToDevice);
  }
  if (cfg->reset_settle_cache && err == cudaSuccess) {
    // the settle phase once, on a one-env scratch record, by the same reset program every env would run (teleport, caches and
    // pending force cleared, settle_steps substeps, no pushes); its result is what every later reset starts its pushes from
    const size_t fb1 = h->f64() ? bcp::Launch<double>::state_bytes(1) : bcp::Launch<float>::state_bytes(1);
    A(&h->settled_f, fb1);
    A((void**)&h->settled_i, (size_t)bcp::NI * sizeof(int));
    if (err == cudaSuccess) {
      if (h->f64()) {
        bcp::Params<double> P1 = h->p64;
        P1.num_envs = 1; P1.push_steps = 0;
        err = bcp::Launch<double>::reset(P1, h->settled_f, h->settled_i, nullptr, nullptr, 1, nullptr, 0);
      } else {
        bcp::Params<float> P1 = h->p32;
        P1.num_envs = 1; P1.push_steps = 0;
        err = bcp::Launch<float>::reset(P1, h->settled_f, h->settled_i, nullptr, nullptr, 1, nullptr, 0);
      }
      if (err == cudaSuccess) err = cudaDeviceSynchronize();
      h->p32.settled = static_cast<const float*>(h->settled_f);
      h->p64.settled = static_cast<const double*>(h->settled_f);
      h->p32.settled_cnt = h->p64.settled_cnt = h->settled_i + bcp::I_CNT;
    }
  }
  // fp32 with enough envs to fill the machine: env.step runs as the split pipeline
  // (BCP_SPLIT=0/1 overrides; the one-launch fast kernel is the alternative)
  {
    const char* sv = getenv("BCP_SPLIT");
    h->use_split = (!h->f64()) && (sv ? atoi(sv) != 0 : cfg->num_envs >= 4096);
    // the pooled solver implements the default row order and the two-direction friction cone
    if (cfg->manifold_order != 0 || cfg->friction_dirs != 2 || !cfg->cone_friction) h->use_split = 0;
    if (h->use_split) {
      const size_t rw = bcp::split_record_words(h->p32);
      A((void**)&h->recs, rw * N * sizeof(float));
      h->use_lam = bcp::split_uses_lambda(h->p32) ? 1 : 0;
      // two cohorts from 32 768 envs, four from 65 536, then one per 131 072, at most 8 (measured on B200 with the call as a CUDA
      // graph: 65 536 envs as 1 / 2 / 3 / 4 cohorts 0.355 / 0.311 / 0.307 / 0.302 ms per env.step; 131 072 as 2 or 4: 0.52;
      // 262 144 as 4: 1.00, as 8: 1.00; 1 M as 8: 3.9); BCP_COHORTS overrides
      const char* cv = getenv("BCP_COHORTS");
      int nc = 1;
      if (cv) nc = atoi(cv);
      else if (cfg->num_envs >= 32768) nc = cfg->num_envs < 65536 ? 2 : (cfg->num_envs <= 262144 ? 4 : (cfg->num_envs + 131071) / 131072);
      if (nc < 1) nc = 1;
      if (nc > 8) nc = 8;
      while (nc > 1 && cfg->num_envs / nc < 1024) nc--;
      h->cohorts.resize(nc);
      for (int c = 0; c < nc && err == cudaSuccess; ++c) {
        bcp_env::Cohort& co = h->cohorts[c];
        co.base = (int)((long long)cfg->num_envs * c / nc) & ~127;           // whole blocks, 16-byte aligned outputs
        const int end = (c == nc - 1) ? cfg->num_envs : ((int)((long long)cfg->num_envs * (c + 1) / nc) & ~127);
        co.n = end - co.base;
        co.P = h->p32;
        co.P.num_envs = co.n;
        co.P.env_id_offset = h->p32.env_id_offset + co.base;
        const size_t n = (size_t)co.n;
        co.b.recs = h->recs + (size_t)co.base * rw;
        co.b.lam = h->use_lam;
        co.b.trace_stride = cfg->num_envs;
        A((void**)&co.b.queue, n * 4 * sizeof(int));  // {env, contact counts} per entry; main queue, then the generic one
        A((void**)&co.b.qc, bcp::split_counter_words(co.P) * sizeof(int));
        A((void**)&co.b.cont, bcp::split_cont_words(co.P) * sizeof(float));
        A((void**)&co.b.contm, n * 4 * sizeof(int));
        for (int k = 0; k < 2; ++k) {
          A((void**)&co.b.bins[k], bcp::split_bins_words(co.P) * sizeof(int));
          A((void**)&co.b.cnts[k], 8 * sizeof(int));
        }
        if (c > 0 && err == cudaSuccess) {
          err = cudaStreamCreateWithFlags(&co.stream, cudaStreamNonBlocking);
          if (err == cudaSuccess) err = cudaEventCreateWithFlags(&co.ev_done, cudaEventDisableTiming);
        }
      }
      if (err == cudaSuccess) err = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
      A(&h->act_buf, N * 2 * sizeof(float));
      const char* gv = getenv("BCP_GRAPH");
      h->use_graph = gv ? atoi(gv) != 0 : 1;
      if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&h->gstream, cudaStreamNonBlocking);
      if (err == cudaSuccess) err = cudaEventCreateWithFlags(&h->ev_gin, cudaEventDisableTiming);
      if (err == cudaSuccess) err = cudaEventCreateWithFlags(&h->ev_gout, cudaEventDisableTiming);
    }
  }
  if (err != cudaSuccess) {
    bcp_destroy(h);
    return fail("bcp_create: device allocation failed", err);
  }
  *out = h;
  return 0;
}

int bcp_destroy(bcp_env* h) {
  if (!h) return 0;
  DeviceGuard g(h->cfg.device_ordinal);
  cudaFree(h->fdata); cudaFree(h->idata); cudaFree(h->hist); cudaFree(h->header);
  cudaFree(h->reward_sum); cudaFree(h->list); cudaFree(h->count); cudaFree(h->staging);
  cudaFree(h->settled_f); cudaFree(h->settled_i);
  if (h->hs.d_out) {
    cudaFree(h->hs.d_act); cudaFree(h->hs.d_out); cudaFree(h->hs.d_slab);
    cudaStreamDestroy(h->hs.copy); cudaEventDestroy(h->hs.ev_ready); cudaEventDestroy(h->hs.ev_copied); if (h->hs.ev_h2d) cudaEventDestroy(h->hs.ev_h2d);
  }
  for (bcp_env::Cohort& co : h->cohorts) {
    if (co.b.side) { cudaStreamDestroy(co.b.side); cudaEventDestroy(co.b.ev_prep); cudaEventDestroy(co.b.ev_gen); }
    cudaFree(co.b.queue); cudaFree(co.b.qc); cudaFree(co.b.cont); cudaFree(co.b.contm);
    for (int k = 0; k < 2; ++k) { cudaFree(co.b.bins[k]); cudaFree(co.b.cnts[k]); }
    if (co.stream) cudaStreamDestroy(co.stream);
    if (co.ev_done) cudaEventDestroy(co.ev_done);
    for (cudaEvent_t e : co.ev_prep) cudaEventDestroy(e);
  }
  for (bcp_env::StepGraph& g : h->graphs) cudaGraphExecDestroy(g.exec);
  for (bcp_env::HostGraph& g : h->host_graphs) cudaGraphExecDestroy(g.exec);
  cudaFree(h->act_buf);
  if (h->gstream) { cudaStreamDestroy(h->gstream); cudaEventDestroy(h->ev_gin); cudaEventDestroy(h->ev_gout); }
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  cudaFree(h->trace);
  cudaFree(h->recs);
  delete h;
  return 0;
}

int bcp_reset(bcp_env* h, const uint8_t* mask, float* obs_out, void* stream) {
  if (!h) return fail("bcp_reset: null handle");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  const int* list = nullptr;
  if (mask) {
    CK(bcp::launch_compact(mask, h->cfg.num_envs, h->list, h->count, s));
    h->launches++;
    list = h->list;
  }
  if (h->f64()) CK(bcp::Launch<double>::reset(h->p64, h->fdata, h->idata, list, h->count, h->cfg.num_envs, obs_out, s));
  else CK(bcp::Launch<float>::reset(h->p32, h->fdata, h->idata, list, h->count, h->cfg.num_envs, obs_out, s));
  h->launches++;
  return 0;
}

}  // extern "C"

// one call of every cohort, enqueued on `s` and the cohorts' own streams (forked from and joined back into `s`)
// slabs: `obs_out` is repeat-major, [R][num_envs][2][7] (bcp_step_host), instead of the (num_envs, R, 2, 7) block
static int enqueue_split(bcp_env* h, const void* actions, float* obs_out, float* reward_out, uint8_t* done_out, uint8_t* reason_out,
                         cudaStream_t s, int* launches, bool slabs = false) {
  const size_t env_stride = slabs ? 14 : (size_t)14 * h->cfg.action_repeats;
  const size_t rep_stride = slabs ? (size_t)14 * h->cfg.num_envs : 14;
  const size_t abytes = h->cfg.discrete_actions ? sizeof(int32_t) : 2 * sizeof(float);
  // while the kernels are being bracketed by events (bcp_profile_begin) the cohorts run one after the other on the
  // caller's stream, so that a kernel's time is its own
  const bool fork = h->cohorts.size() > 1 && !h->profiling;
  if (fork) CK(cudaEventRecord(h->ev_fork, s));
  for (size_t c = 0; c < h->cohorts.size(); ++c) {
    bcp_env::Cohort& co = h->cohorts[c];
    cudaStream_t cs = (fork && c > 0) ? co.stream : s;
    if (fork && c > 0) CK(cudaStreamWaitEvent(cs, h->ev_fork, 0));
    const size_t b = (size_t)co.base;
    const size_t ev0 = h->prof_events.size();
    co.b.trace = h->trace ? h->trace + b * 4 : nullptr;
    co.b.after_prep = h->split_hook;
    co.b.hook_ctx = h->split_hook ? &h->hook_args[c] : nullptr;
    CK(bcp::launch_step_split(co.P, (float*)h->fdata + b * bcp::NF, h->idata + b * bcp::NI, co.b,
                              static_cast<const char*>(actions) + b * abytes, obs_out ? obs_out + b * env_stride : nullptr,
                              reward_out ? reward_out + b : nullptr, done_out ? done_out + b : nullptr,
                              reason_out ? reason_out + b : nullptr, h->stats(), cs, launches,
                              h->profiling ? &h->prof_events : nullptr, env_stride, rep_stride));
    if (h->profiling) h->prof_calls.push_back((int)(h->prof_events.size() - ev0));
    if (fork && c > 0) {
      CK(cudaEventRecord(co.ev_done, cs));
      CK(cudaStreamWaitEvent(s, co.ev_done, 0));
    }
  }
  return 0;
}

extern "C" {

int bcp_step(bcp_env* h, const void* actions, float* obs_out, float* reward_out, uint8_t* done_out,
             uint8_t* reason_out, void* stream) {
  if (!h) return fail("bcp_step: null handle");
  if (!actions) return fail("bcp_step: null actions");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  if (!h->f64() && h->use_split) {
    // Steady state: the whole call - every cohort's memsets and kernels, the forks to the cohort and side streams and their
    // joins - is one CUDA graph launch (the host then spends ~10 us per env.step instead of one launch call per kernel).
    // Captured the first time a (outputs, stream, visiting-order parity) combination is seen after the handle's first call.
    // (a caller that is itself capturing `stream` gets the plain stream launches, which its capture records)
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
    const bool plain = h->use_graph && cap == cudaStreamCaptureStatusNone && !h->profiling && !h->trace && !h->split_hook &&
                       h->cohorts[0].b.cur >= 0 && h->cohorts[0].b.side_ready;
    if (plain) {
      const size_t abytes = (size_t)h->cfg.num_envs * (h->cfg.discrete_actions ? sizeof(int32_t) : 2 * sizeof(float));
      const int parity = h->cohorts[0].b.cur;
      bcp_env::StepGraph* g = nullptr;
      for (bcp_env::StepGraph& c : h->graphs)
        if (c.obs == obs_out && c.reward == reward_out && c.done == done_out && c.reason == reason_out && c.parity == parity) g = &c;
      if (!g && h->graphs.size() < 16) {
        cudaGraph_t graph = nullptr;
        int n = 0;
        cudaError_t e = cudaStreamBeginCapture(h->gstream, cudaStreamCaptureModeThreadLocal);
        int rc = 0;
        if (e == cudaSuccess) {
          rc = enqueue_split(h, h->act_buf, obs_out, reward_out, done_out, reason_out, h->gstream, &n);
          e = cudaStreamEndCapture(h->gstream, &graph);
        }
        cudaGraphExec_t exec = nullptr;
        if (e == cudaSuccess && rc == 0 && graph) e = cudaGraphInstantiate(&exec, graph, 0);
        if (graph) cudaGraphDestroy(graph);
        if (e != cudaSuccess || rc != 0 || !exec) {   // no graphs on this handle from now on; the direct path below still works
          cudaGetLastError();
          h->use_graph = 0;
          for (bcp_env::Cohort& co : h->cohorts) co.b.cur = parity;   // the captured enqueue flipped it
        } else {
          for (bcp_env::Cohort& co : h->cohorts) co.b.cur = parity;   // the capture ran the host code once: undo its flip
          h->graphs.push_back(bcp_env::StepGraph{obs_out, reward_out, done_out, reason_out, parity, h->gstream, exec, n});
          g = &h->graphs.back();
        }
      }
      if (g) {
        CK(cudaMemcpyAsync(h->act_buf, actions, abytes, cudaMemcpyDeviceToDevice, s));
        CK(cudaEventRecord(h->ev_gin, s));
        CK(cudaStreamWaitEvent(h->gstream, h->ev_gin, 0));
        CK(cudaGraphLaunch(g->exec, h->gstream));
        CK(cudaEventRecord(h->ev_gout, h->gstream));
        CK(cudaStreamWaitEvent(s, h->ev_gout, 0));
        for (bcp_env::Cohort& co : h->cohorts) co.b.cur = 1 - parity;
        h->launches += g->launches;
        return 0;
      }
    }
    int n = 0;
    if (enqueue_split(h, actions, obs_out, reward_out, done_out, reason_out, s, &n)) return -1;
    h->launches += n;
    return 0;
  }
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (h->profiling) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, s); }
  if (h->f64()) CK(bcp::Launch<double>::step(h->p64, h->fdata, h->idata, actions, obs_out, reward_out, done_out, reason_out, h->stats(), s));
  else CK(bcp::Launch<float>::step(h->p32, h->fdata, h->idata, actions, obs_out, reward_out, done_out, reason_out, h->stats(), s));
  if (h->profiling) { cudaEventRecord(e1, s); h->prof_events.push_back(e0); h->prof_events.push_back(e1); h->prof_calls.push_back(2); }
  h->launches++;
  return 0;
}

// ---- bcp_step_host: env.step with HOST buffers, copies overlapped with the kernels ----
namespace {
__global__ void k_pack_slab(const float* __restrict__ obs, float* __restrict__ slab, int n, int R, int r) {
  // obs f32[n][R][14] -> slab f32[n][14] (repeat r), one thread per float2
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * 7) return;
  const int e = i / 7, w = i - e * 7;
  reinterpret_cast<float2*>(slab)[i] = reinterpret_cast<const float2*>(obs + ((size_t)e * R + r) * 14)[w];
}
void pack_and_copy(bcp_env* h, int r, cudaStream_t s) {
  const int n = h->cfg.num_envs, R = h->cfg.action_repeats;
  float* slab = h->hs.d_slab + (size_t)r * n * 14;
  k_pack_slab<<<(n * 7 + 255) / 256, 256, 0, s>>>(reinterpret_cast<const float*>(h->hs.d_out), slab, n, R, r);
  cudaMemcpyAsync(h->hs.h_obs + (size_t)r * n * 14, slab, (size_t)n * 14 * sizeof(float), cudaMemcpyDeviceToHost, s);
  h->launches++;
}
// after the k-th k_prep of a cohort's call has been enqueued on its stream `s`: repeat r = k / S - 1 is complete for
// that cohort when k is a multiple of S.  The cohorts are enqueued one after the other, so the LAST cohort's hook is the
// point at which every cohort's event for k exists: the copy stream waits for all of them, packs the slab and copies it.
void host_step_hook(void* ctx, int k, cudaStream_t s) {
  bcp_env::HookArg* a = static_cast<bcp_env::HookArg*>(ctx);
  bcp_env* h = a->h;
  const int R = h->cfg.action_repeats, S = h->cfg.steps_per_repeat;
  const bool last = a->cohort == (int)h->cohorts.size() - 1;
  if (k == 0 && h->hs.h2d_pending) {
    cudaStreamWaitEvent(s, h->hs.ev_h2d, 0);  // every later kernel of the call reads the actions
    if (last) h->hs.h2d_pending = false;
  }
  if (k <= 0 || k >= R * S || k % S) return;  // the last repeat leaves with the call's other outputs
  bcp_env::Cohort& co = h->cohorts[a->cohort];
  while ((int)co.ev_prep.size() <= k) {
    cudaEvent_t e;
    cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    co.ev_prep.push_back(e);
  }
  cudaEventRecord(co.ev_prep[k], s);
  if (!last) return;
  for (bcp_env::Cohort& c : h->cohorts) cudaStreamWaitEvent(h->hs.copy, c.ev_prep[k], 0);
  {  // the kernels wrote repeat r straight into its slab: no pack kernel, just the copy
    const size_t n = (size_t)h->cfg.num_envs, r = (size_t)(k / S - 1);
    cudaMemcpyAsync(h->hs.h_obs + r * n * 14, h->hs.d_slab + r * n * 14, n * 14 * sizeof(float), cudaMemcpyDeviceToHost, h->hs.copy);
  }
  cudaEventRecord(h->hs.ev_copied, h->hs.copy);
  h->hs.fired = true;
}
}  // namespace

}  // extern "C"

// the host-buffer call, enqueued on `s` (and the copy / cohort / side streams forked from it and joined back)
static int host_step_enqueue(bcp_env* h, const void* h_actions, float* h_obs_rm, float* h_reward, uint8_t* h_done, uint8_t* h_reason,
                             cudaStream_t s) {
  const size_t n = (size_t)h->cfg.num_envs, R = (size_t)h->cfg.action_repeats;
  const size_t ob = n * R * 14 * sizeof(float), rb = n * sizeof(float);
  const size_t ab = n * (h->cfg.discrete_actions ? sizeof(int32_t) : 2 * sizeof(float));
  float* d_obs = reinterpret_cast<float*>(h->hs.d_out);
  float* d_rew = reinterpret_cast<float*>(h->hs.d_out + ob);
  uint8_t* d_done = h->hs.d_out + ob + rb;
  uint8_t* d_reason = d_done + n;
  const bool split = !h->f64() && h->use_split;
  if (split) {  // the first kernel of a split call (collision pass and rows of substep 0) does not depend on the actions
    CK(cudaEventRecord(h->hs.ev_ready, s));
    CK(cudaStreamWaitEvent(h->hs.copy, h->hs.ev_ready, 0));  // whatever read d_act before is done
    CK(cudaMemcpyAsync(h->hs.d_act, h_actions, ab, cudaMemcpyHostToDevice, h->hs.copy));
    CK(cudaEventRecord(h->hs.ev_h2d, h->hs.copy));
    h->hs.h2d_pending = true;
  } else {
    CK(cudaMemcpyAsync(h->hs.d_act, h_actions, ab, cudaMemcpyHostToDevice, s));
  }
  h->hs.h_obs = h_obs_rm;
  h->hs.fired = false;
  if (split) {
    // Split pipeline: k_prep writes the observation of repeat r straight into slab r of the device block
    // [R slabs | reward | done | reason]; slab r leaves as soon as every cohort's k_prep that completes it is done, and the last
    // slab leaves together with reward, done and reason - ONE copy when the caller's buffers are laid out the same way.
    float* s_obs = h->hs.d_slab;
    float* s_rew = h->hs.d_slab + n * R * 14;
    uint8_t* s_done = reinterpret_cast<uint8_t*>(s_rew + n);
    uint8_t* s_reason = s_done + n;
    h->hook_args.resize(h->cohorts.size());
    for (size_t c = 0; c < h->cohorts.size(); ++c) h->hook_args[c] = bcp_env::HookArg{h, (int)c};
    h->split_hook = host_step_hook;
    int nl = 0;
    const int rc = enqueue_split(h, h->hs.d_act, s_obs, s_rew, s_done, s_reason, s, &nl, true);
    h->split_hook = nullptr;
    h->launches += nl;
    if (h->hs.h2d_pending) {  // the call failed before its first kernel
      cudaStreamWaitEvent(s, h->hs.ev_h2d, 0);
      h->hs.h2d_pending = false;
    }
    if (rc) return rc;
    const size_t r0 = h->hs.fired ? R - 1 : 0;  // repeats that did not leave during the call
    float* h_from = h_obs_rm + r0 * n * 14;
    const float* d_from = s_obs + r0 * n * 14;
    const size_t obs_bytes = (R - r0) * n * 14 * sizeof(float);
    if (reinterpret_cast<uint8_t*>(h_obs_rm) + ob == reinterpret_cast<uint8_t*>(h_reward) && reinterpret_cast<uint8_t*>(h_reward) + rb == h_done &&
        h_done + n == h_reason) {
      CK(cudaMemcpyAsync(h_from, d_from, obs_bytes + rb + 2 * n, cudaMemcpyDeviceToHost, s));
    } else {
      CK(cudaMemcpyAsync(h_from, d_from, obs_bytes, cudaMemcpyDeviceToHost, s));
      CK(cudaMemcpyAsync(h_reward, s_rew, rb, cudaMemcpyDeviceToHost, s));
      CK(cudaMemcpyAsync(h_done, s_done, n, cudaMemcpyDeviceToHost, s));
      CK(cudaMemcpyAsync(h_reason, s_reason, n, cudaMemcpyDeviceToHost, s));
    }
    if (h->hs.fired) CK(cudaStreamWaitEvent(s, h->hs.ev_copied, 0));  // a sync on `stream` covers the early slabs too
    CK(cudaGetLastError());
    return 0;
  }
  // one-launch pipelines (fp64, small batches): bcp_step into the (B, R, 2, 7) block, then pack and copy every repeat
  const int rc = bcp_step(h, h->hs.d_act, d_obs, d_rew, d_done, d_reason, (void*)s);
  if (rc) return rc;
  for (int r = 0; r < (int)R; ++r) pack_and_copy(h, r, s);
  if (reinterpret_cast<uint8_t*>(h_reward) + rb == h_done && h_done + n == h_reason) {
    CK(cudaMemcpyAsync(h_reward, d_rew, rb + 2 * n, cudaMemcpyDeviceToHost, s));  // one transfer when the caller's buffers are adjacent
  } else {
    CK(cudaMemcpyAsync(h_reward, d_rew, rb, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(h_done, d_done, n, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(h_reason, d_reason, n, cudaMemcpyDeviceToHost, s));
  }
  CK(cudaGetLastError());
  return 0;
}

extern "C" {

int bcp_step_host(bcp_env* h, const void* h_actions, float* h_obs_rm, float* h_reward, uint8_t* h_done, uint8_t* h_reason,
                  void* stream) {
  if (!h) return fail("bcp_step_host: null handle");
  if (!h_actions || !h_obs_rm || !h_reward || !h_done || !h_reason) return fail("bcp_step_host: null argument");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t n = (size_t)h->cfg.num_envs, R = (size_t)h->cfg.action_repeats;
  const size_t ob = n * R * 14 * sizeof(float), rb = n * sizeof(float);
  const size_t ab = n * (h->cfg.discrete_actions ? sizeof(int32_t) : 2 * sizeof(float));
  if (!h->hs.d_out) {
    CK(cudaMalloc(&h->hs.d_act, ab));
    CK(cudaMalloc((void**)&h->hs.d_out, ob + rb + 2 * n));
    CK(cudaMalloc((void**)&h->hs.d_slab, ob + rb + 2 * n));  // [R slabs | reward | done | reason]
    // an env that is mid-reset keeps its previous observation in the output block (bcp_step writes nothing there): start from zeros like the caller's own buffers
    CK(cudaMemsetAsync(h->hs.d_out, 0, ob + rb + 2 * n, s));
    CK(cudaMemsetAsync(h->hs.d_slab, 0, ob + rb + 2 * n, s));
    {
      int lo = 0, hi = 0;
      cudaDeviceGetStreamPriorityRange(&lo, &hi);
      CK(cudaStreamCreateWithPriority(&h->hs.copy, cudaStreamNonBlocking, hi));  // its pack kernel goes first when blocks retire
    }
    CK(cudaEventCreateWithFlags(&h->hs.ev_ready, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->hs.ev_copied, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->hs.ev_h2d, cudaEventDisableTiming));
  }
  // Steady state on the split pipeline: the whole host call - H2D of the actions, every cohort's kernels, the slab packs
  // and the D2H copies into the caller's (pinned, pointer-stable) buffers - is one CUDA graph per (buffers, parity).
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
  const bool can_graph = h->use_graph && cap == cudaStreamCaptureStatusNone && !h->f64() && h->use_split && !h->profiling && !h->trace &&
                         h->cohorts[0].b.cur >= 0 && h->cohorts[0].b.side_ready;
  if (can_graph) {
    const int parity = h->cohorts[0].b.cur;
    bcp_env::HostGraph* hg = nullptr;
    for (bcp_env::HostGraph& c : h->host_graphs)
      if (c.act == h_actions && c.obs == h_obs_rm && c.reward == h_reward && c.done == h_done && c.reason == h_reason && c.parity == parity) hg = &c;
    if (!hg && h->host_graphs.size() < 16) {
      cudaGraph_t graph = nullptr;
      const int64_t l0 = h->launches;
      cudaError_t e = cudaStreamBeginCapture(h->gstream, cudaStreamCaptureModeThreadLocal);
      int rc = 0;
      if (e == cudaSuccess) {
        rc = host_step_enqueue(h, h_actions, h_obs_rm, h_reward, h_done, h_reason, h->gstream);
        e = cudaStreamEndCapture(h->gstream, &graph);
      }
      const int nl = (int)(h->launches - l0);
      h->launches = l0;
      cudaGraphExec_t exec = nullptr;
      if (e == cudaSuccess && rc == 0 && graph) e = cudaGraphInstantiate(&exec, graph, 0);
      if (graph) cudaGraphDestroy(graph);
      for (bcp_env::Cohort& co : h->cohorts) co.b.cur = parity;   // the captured enqueue flipped it
      if (e != cudaSuccess || rc != 0 || !exec) {
        cudaGetLastError();
        h->use_graph = 0;
      } else {
        h->host_graphs.push_back(bcp_env::HostGraph{h_actions, h_obs_rm, h_reward, h_done, h_reason, parity, exec, nl});
        hg = &h->host_graphs.back();
      }
    }
    if (hg) {
      CK(cudaEventRecord(h->ev_gin, s));
      CK(cudaStreamWaitEvent(h->gstream, h->ev_gin, 0));
      CK(cudaGraphLaunch(hg->exec, h->gstream));
      CK(cudaEventRecord(h->ev_gout, h->gstream));
      CK(cudaStreamWaitEvent(s, h->ev_gout, 0));
      for (bcp_env::Cohort& co : h->cohorts) co.b.cur = 1 - parity;
      h->launches += hg->launches;
      return 0;
    }
  }
  return host_step_enqueue(h, h_actions, h_obs_rm, h_reward, h_done, h_reason, s);
}

int bcp_step_host_finish(bcp_env* h, float* h_obs_rm, const uint8_t* h_reason, void* stream) {
  if (!h || !h_obs_rm || !h_reason) return fail("bcp_step_host_finish: null argument");
  DeviceGuard g(h->cfg.device_ordinal);
  CK(cudaStreamSynchronize((cudaStream_t)stream));
  // An env whose pipelined reset ended in this call (FRESH) gets its reset observation in every
  // repeat slot when the call closes (bullet_cartpole.py:338-339) - after the early slabs left.
  const size_t n = (size_t)h->cfg.num_envs;
  const int R = h->cfg.action_repeats;
  if (R > 1 && h->cfg.auto_reset == BCP_AUTO_RESET_PIPELINED) {
    const float* last = h_obs_rm + (size_t)(R - 1) * n * 14;
    auto patch = [&](size_t e) {
      for (int r = 0; r < R - 1; ++r) memcpy(h_obs_rm + ((size_t)r * n + e) * 14, last + e * 14, 14 * sizeof(float));
    };
    size_t e = 0;
    for (; e + 8 <= n; e += 8) {  // eight reason bytes at a time: any equal to BCP_REASON_FRESH?
      uint64_t w;
      memcpy(&w, h_reason + e, 8);
      w ^= 0x0505050505050505ull;
      if ((w - 0x0101010101010101ull) & ~w & 0x8080808080808080ull)
        for (size_t k = e; k < e + 8; ++k) if (h_reason[k] == BCP_REASON_FRESH) patch(k);
    }
    for (; e < n; ++e) if (h_reason[e] == BCP_REASON_FRESH) patch(e);
  }
  return 0;
}

int bcp_substep(bcp_env* h, void* stream) {
  if (!h) return fail("bcp_substep: null handle");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  if (h->f64()) CK(bcp::Launch<double>::substep(h->p64, h->fdata, h->idata, s));
  else CK(bcp::Launch<float>::substep(h->p32, h->fdata, h->idata, s));
  h->launches++;
  return 0;
}

int bcp_rollout_random(bcp_env* h, int32_t n_steps, uint32_t action_set_mask, int32_t max_episodes, float* obs_out,
                       void* stream) {
  if (!h) return fail("bcp_rollout_random: null handle");
  if (n_steps < 0 || max_episodes < 0) return fail("bcp_rollout_random: n_calls < 0 or max_episodes < 0");
  if (h->cfg.discrete_actions && (action_set_mask & 0x1fu) == 0) return fail("bcp_rollout_random: empty action set");
  if (n_steps == 0) return 0;
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  bcp::Params<double> p64 = h->p64;
  bcp::Params<float> p32 = h->p32;
  p64.auto_reset = p32.auto_reset = BCP_AUTO_RESET_PIPELINED;  // a rollout resets its envs in place
  const char* rv = getenv("BCP_ROLLOUT_SPLIT");  // read per call: tests compare the two paths in one process
  const int roll_split = rv ? atoi(rv) : 1;
  if (!h->f64() && h->use_split && roll_split && !h->profiling) {
    // the same n_steps calls of the pipelined machine as k_rollout runs in one launch, through the split pipeline (three times
    // the env-steps/s at these batch sizes): per call, the agent's actions (k_draw_actions), then the pipeline's kernels
    const uint32_t mask = action_set_mask & 0x1fu;
    int rc = 0;
    for (bcp_env::Cohort& co : h->cohorts) {
      co.b.roll_max_episodes = max_episodes;
      co.b.rollout = 1;
      co.P.auto_reset = BCP_AUTO_RESET_PIPELINED;
    }
    for (int t = 0; t < n_steps && rc == 0; ++t) {
      if (bcp::launch_draw_actions(p32, h->idata, h->act_buf, mask, s) != cudaSuccess) { rc = -1; break; }
      int n = 1;
      rc = enqueue_split(h, h->act_buf, obs_out, nullptr, nullptr, nullptr, s, &n);
      h->launches += n;
    }
    for (bcp_env::Cohort& co : h->cohorts) {
      co.b.roll_max_episodes = 0;
      co.b.rollout = 0;
      co.P.auto_reset = h->p32.auto_reset;
    }
    if (rc) return fail("bcp_rollout_random: split pipeline launch", cudaGetLastError());
    return 0;
  }
  if (h->f64()) CK(bcp::Launch<double>::rollout(p64, h->fdata, h->idata, n_steps, action_set_mask & 0x1fu, max_episodes, obs_out, h->stats(), s));
  else CK(bcp::Launch<float>::rollout(p32, h->fdata, h->idata, n_steps, action_set_mask & 0x1fu, max_episodes, obs_out, h->stats(), s));
  h->launches++;
  return 0;
}

int bcp_stats(bcp_env* h, int64_t* hist, int64_t* header, double* reward_sum, void* stream) {
  if (!h) return fail("bcp_stats: null handle");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t nb = (size_t)h->cfg.max_episode_len + 1;
  if (hist) CK(cudaMemcpyAsync(hist, h->hist, nb * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  if (header) CK(cudaMemcpyAsync(header, h->header, BCP_STATS_HEADER * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  if (reward_sum) CK(cudaMemcpyAsync(reward_sum, h->reward_sum, sizeof(double), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  return 0;
}

int bcp_stats_clear(bcp_env* h, void* stream) {
  if (!h) return fail("bcp_stats_clear: null handle");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  CK(cudaMemsetAsync(h->hist, 0, ((size_t)h->cfg.max_episode_len + 1) * sizeof(unsigned long long), s));
  CK(cudaMemsetAsync(h->header, 0, BCP_STATS_HEADER * sizeof(unsigned long long), s));
  CK(cudaMemsetAsync(h->reward_sum, 0, sizeof(double), s));
  return 0;
}

static int ensure_staging(bcp_env* h) {
  if (!h->staging) CK(cudaMalloc((void**)&h->staging, (size_t)h->cfg.num_envs * BCP_STATE_DOUBLES * sizeof(double)));
  return 0;
}

int bcp_get_state(bcp_env* h, double* host_buf, size_t n_doubles, void* stream) {
  if (!h || !host_buf) return fail("bcp_get_state: null argument");
  const size_t need = (size_t)h->cfg.num_envs * BCP_STATE_DOUBLES;
  if (n_doubles < need) return fail("bcp_get_state: buffer too small");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  if (ensure_staging(h)) return -1;
  if (h->f64()) CK(bcp::Launch<double>::export_state(h->cfg.num_envs, h->fdata, h->idata, h->staging, s));
  else CK(bcp::Launch<float>::export_state(h->cfg.num_envs, h->fdata, h->idata, h->staging, s));
  h->launches++;
  CK(cudaMemcpyAsync(host_buf, h->staging, need * sizeof(double), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  return 0;
}

int bcp_set_state(bcp_env* h, const double* host_buf, size_t n_doubles, void* stream) {
  if (!h || !host_buf) return fail("bcp_set_state: null argument");
  const size_t need = (size_t)h->cfg.num_envs * BCP_STATE_DOUBLES;
  if (n_doubles < need) return fail("bcp_set_state: buffer too small");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  if (ensure_staging(h)) return -1;
  CK(cudaMemcpyAsync(h->staging, host_buf, need * sizeof(double), cudaMemcpyHostToDevice, s));
  if (h->f64()) CK(bcp::Launch<double>::import_state(h->cfg.num_envs, h->fdata, h->idata, h->staging, s));
  else CK(bcp::Launch<float>::import_state(h->cfg.num_envs, h->fdata, h->idata, h->staging, s));
  h->launches++;
  CK(cudaStreamSynchronize(s));
  return 0;
}

int bcp_bench_fma_peak(int device_ordinal, int use_f64, int iters, double* tflops_out) {
  if (!tflops_out || iters < 1) return fail("bcp_bench_fma_peak: bad argument");
  DeviceGuard g(device_ordinal);
  CK(bcp::launch_fma_peak(use_f64, iters, tflops_out));
  return 0;
}

int64_t bcp_launch_count(const bcp_env* h) { return h ? h->launches : 0; }

int bcp_solver_kind(const bcp_env* env) {
  if (!env) return fail("bcp_solver_kind: null handle");
  if (env->f64() || !env->use_split || !env->recs) return BCP_SOLVER_ONE_LAUNCH;
  return env->use_lam ? BCP_SOLVER_SPLIT_IMPULSE : BCP_SOLVER_SPLIT_ROW;
}

int bcp_profile_begin(bcp_env* h) {
  if (!h) return fail("bcp_profile_begin: null handle");
  for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
  h->prof_events.clear();
  h->prof_calls.clear();
  h->profiling = true;
  return 0;
}

int bcp_profile_end(bcp_env* h, double* out4) {
  if (!h || !out4) return fail("bcp_profile_end: null argument");
  DeviceGuard g(h->cfg.device_ordinal);
  h->profiling = false;
  double ms[2] = {0, 0};
  long long cnt[2] = {0, 0};
  size_t at = 0;
  for (int n : h->prof_calls) {
    // n events bracket n-1 kernels: k_prep, k_solve, k_prep, ... (split) or one kernel
    for (int k = 0; k + 1 < n; ++k) {
      CK(cudaEventSynchronize(h->prof_events[at + k + 1]));
      float t = 0;
      CK(cudaEventElapsedTime(&t, h->prof_events[at + k], h->prof_events[at + k + 1]));
      ms[k & 1] += t;
      cnt[k & 1]++;
    }
    at += n;
  }
  for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
  h->prof_events.clear();
  h->prof_calls.clear();
  out4[0] = ms[0]; out4[1] = ms[1]; out4[2] = (double)cnt[0]; out4[3] = (double)cnt[1];
  return 0;
}

int bcp_last_solve_info(bcp_env* h, int32_t* host_out, size_t n_ints, void* stream) {
  if (!h || !host_out) return fail("bcp_last_solve_info: null argument");
  if (h->f64() || !h->use_split || !h->recs) return fail("bcp_last_solve_info: the handle does not run the split pipeline");
  if (n_ints != 4 * (size_t)h->cfg.num_envs) return fail("bcp_last_solve_info: n_ints must be 4 * num_envs");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  // the first four words of every record: three contact counts and the iteration count (bit patterns of ints)
  const size_t pitch = bcp::split_record_words(h->p32) * sizeof(float);
  CK(cudaMemcpy2DAsync(host_out, 4 * sizeof(int32_t), h->recs, pitch, 4 * sizeof(int32_t), (size_t)h->cfg.num_envs,
                       cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  return 0;
}


int bcp_solve_trace_enable(bcp_env* h, int on) {
  if (!h) return fail("bcp_solve_trace_enable: null handle");
  if (h->f64() || !h->use_split || !h->recs) return fail("bcp_solve_trace_enable: the handle does not run the split pipeline");
  DeviceGuard g(h->cfg.device_ordinal);
  if (on && !h->trace) {
    const size_t bytes = (size_t)h->cfg.action_repeats * h->cfg.steps_per_repeat * h->cfg.num_envs * 4 * sizeof(int);
    CK(cudaMalloc((void**)&h->trace, bytes));
    CK(cudaMemset(h->trace, 0, bytes));
  } else if (!on && h->trace) {
    CK(cudaDeviceSynchronize());
    CK(cudaFree(h->trace));
    h->trace = nullptr;
  }
  return 0;
}

int bcp_solve_trace_read(bcp_env* h, int32_t* host_out, size_t n_ints, void* stream) {
  if (!h || !host_out) return fail("bcp_solve_trace_read: null argument");
  if (!h->trace) return fail("bcp_solve_trace_read: tracing is not enabled on this handle");
  const size_t need = (size_t)h->cfg.action_repeats * h->cfg.steps_per_repeat * h->cfg.num_envs * 4;
  if (n_ints != need) return fail("bcp_solve_trace_read: n_ints must be 4 * R * S * num_envs");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  CK(cudaMemcpyAsync(host_out, h->trace, need * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  return 0;
}

}  // extern "C"

// ---- the fork's 8-state observation and LQR feedback (include/bcp_lqr.h) ----
namespace {

// pybullet's getEulerFromQuaternion (q = x,y,z,w): roll and pitch only, with its gimbal-lock branches
template <typename T> __device__ void roll_pitch(const T* q, T& roll, T& pitch) {
  const T sqx = q[0] * q[0], sqy = q[1] * q[1], sqz = q[2] * q[2], squ = q[3] * q[3];
  const T sarg = T(-2) * (q[0] * q[2] - q[3] * q[1]);
  const T half_pi = T(1.5707963267948966);
  if (sarg <= T(-0.99999)) { roll = 0; pitch = -half_pi; }
  else if (sarg >= T(0.99999)) { roll = 0; pitch = half_pi; }
  else { roll = atan2(T(2) * (q[1] * q[2] + q[3] * q[0]), squ - sqx - sqy + sqz); pitch = asin(sarg); }
}

template <typename T> __global__ void k_lqr(int N, const T* __restrict__ f, bcp_lqr_args a) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= N) return;
  const T* b = f + (size_t)e * bcp::NF + BCP_S_POLE;  // x[3] q[4] v[3] w[3]
  T q[4] = {b[3], b[4], b[5], b[6]}, roll, pitch;
  roll_pitch(q, roll, pitch);
  float st[BCP_LQR_STATE];
  st[0] = (float)(b[0] - (a.x0 ? (T)a.x0[e] : T(0)));
  st[1] = (float)b[7];
  st[2] = (float)b[1];
  st[3] = (float)b[8];
  st[4] = (float)roll;
  st[5] = (float)b[10];
  st[6] = (float)pitch;
  st[7] = (float)b[11];
  if (a.state_out) {
    float4* o = reinterpret_cast<float4*>(a.state_out + (size_t)e * BCP_LQR_STATE);
    o[0] = make_float4(st[0], st[1], st[2], st[3]);
    o[1] = make_float4(st[4], st[5], st[6], st[7]);
  }
  if (a.done_out) {
    uint8_t d = BCP_LQR_RUNNING;
    if (fabsf(st[0]) > a.position_threshold || fabsf(st[2]) > a.position_threshold) d = BCP_LQR_POSITION;
    else if (fabsf(st[4]) > a.angle_threshold || fabsf(st[6]) > a.angle_threshold) d = BCP_LQR_ANGLE;
    a.done_out[e] = d;
  }
  if (a.action_out) {
    float u[BCP_LQR_ACTIONS] = {0.f, 0.f};
    if (a.gains) {
      const float* K = a.gains + (a.gains_per_env ? (size_t)e * BCP_LQR_ACTIONS * BCP_LQR_STATE : 0);
      for (int i = 0; i < BCP_LQR_ACTIONS; ++i) {
        // numpy's dot of float64 operands in the fork; here the f32 state against f32 gains, summed in order
        float acc = 0.f;
        for (int j = 0; j < BCP_LQR_STATE; ++j)  // no FMA contraction: the oracle's numpy float32 sum, bit for bit
          acc = __fadd_rn(acc, __fmul_rn(K[i * BCP_LQR_STATE + j], st[j] - (a.zero_point ? a.zero_point[j] : 0.f)));
        u[i] = -acc;
      }
    }
    if (a.disturbance) {
      const float* d = a.disturbance + (a.disturbance_per_env ? (size_t)e * BCP_LQR_ACTIONS : 0);
      u[0] += d[0];
      u[1] += d[1];
    }
    *reinterpret_cast<float2*>(a.action_out + (size_t)e * BCP_LQR_ACTIONS) = make_float2(u[0], u[1]);
  }
}

}  // namespace

extern "C" {

size_t bcp_lqr_args_size(void) { return sizeof(bcp_lqr_args); }

int bcp_lqr_control(bcp_env* h, const bcp_lqr_args* args, void* stream) {
  if (!h || !args) return fail("bcp_lqr_control: null argument");
  if (!args->state_out && !args->action_out && !args->done_out) return fail("bcp_lqr_control: no output requested");
  if (args->done_out && !(args->position_threshold > 0.f && args->angle_threshold > 0.f))
    return fail("bcp_lqr_control: thresholds must be positive");
  DeviceGuard g(h->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  const int N = h->cfg.num_envs, block = 128, grid = (N + block - 1) / block;
  if (h->f64()) k_lqr<double><<<grid, block, 0, s>>>(N, (const double*)h->fdata, *args);
  else k_lqr<float><<<grid, block, 0, s>>>(N, (const float*)h->fdata, *args);
  CK(cudaGetLastError());
  ++h->launches;
  return 0;
}

}  // extern "C"
