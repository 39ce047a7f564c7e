// bcp_launch.cuh — host-side launch interface, one explicit instantiation per scalar
// (bcp_f32.cu: fast mode, FMA contraction on; bcp_f64.cu: parity mode, -fmad=false).
#pragma once
#include <cuda_runtime.h>
#include <vector>
#include "bcp_types.cuh"

namespace bcp {

struct StatsPtrs {
  unsigned long long* hist;
  unsigned long long* header;
  double* reward_sum;
};

template <typename T> struct Launch {
  static size_t state_bytes(int n_envs);  // bytes of the float-field SoA block
  static cudaError_t step(const Params<T>& P, void* f, int* id, const void* actions, float* obs, float* reward,
                          uint8_t* done, uint8_t* reason, StatsPtrs stats, cudaStream_t s);
  // list == nullptr: every env; else the first *count entries of list
  static cudaError_t reset(const Params<T>& P, void* f, int* id, const int* list, const int* count, int max_count,
                           float* obs, cudaStream_t s);
  static cudaError_t substep(const Params<T>& P, void* f, int* id, cudaStream_t s);
  static cudaError_t rollout(const Params<T>& P, void* f, int* id, int n_calls, uint32_t action_mask,
                             int max_episodes, float* obs, StatsPtrs stats, cudaStream_t s);
  static cudaError_t export_state(int n, const void* f, const int* id, double* out, cudaStream_t s);
  static cudaError_t import_state(int n, void* f, int* id, const double* in, cudaStream_t s);
};

// Split pipeline of the fast path (bcp_split.cuh), fp32 only: one call of every env as
// 2*R*S+1 launches.  `recs` holds num_envs solve records of split_record_words() floats,
// `queue` 2 * num_envs ints, `qc` split_counter_words() ints.
struct SplitBuffers {
  float* recs;
  int* queue;
  int* qc;
  int* bins[2];  // visiting order, double buffered: [8][num_envs] env ids each
  int* cnts[2];  // [8] each
  int cur;       // which of the two the next call reads; -1: no order yet (identity)
  float* cont;   // continuation buffer of the impulse-space solver: [LAM_WORDS][num_envs]
  int* contm;    // int4 per slot: env, iterations so far, contact counts, regime
  // the generic queue's k_solve runs beside the impulse-space kernels on a high-priority side stream
  cudaStream_t side;
  cudaEvent_t ev_prep, ev_gen;
  // optional: called right after the k-th k_prep of a call has been enqueued on `s` (bcp_step_host
  // starts the device->host copy of an observation slab as soon as its repeat is complete)
  void (*after_prep)(void* ctx, int k, cudaStream_t s);
  void* hook_ctx;
  // optional (bcp_solve_trace_enable): [R*S][num_envs][4] ints, the first four words of every solve record
  // (three contact counts, PGS iterations) copied out after the solve phase of each substep of a call
  int* trace;
  int trace_stride;  // envs per substep block of `trace` (the handle's env count: a cohort's buffers are views)
  int lam;           // impulse-space solve phase on the main queue: decided once per handle (-1: from this view's size)
  int side_ready;    // the lazily created side stream / kernel attributes exist (or are not needed): the call can be captured
  int roll_max_episodes, rollout;  // set by bcp_rollout_random around its calls (SplitIO::max_episodes / rollout)
};
// the random agent's actions of the next call into `actions` (bcp_rollout_random through the split pipeline)
cudaError_t launch_draw_actions(const Params<float>& P, const int* id, void* actions, uint32_t mask, cudaStream_t s);
size_t split_cont_words(const Params<float>& P);
bool split_uses_lambda(const Params<float>& P);  // impulse-space solve phase for this configuration / batch size?
size_t split_bins_words(const Params<float>& P);
size_t split_record_words(const Params<float>& P);
size_t split_counter_words(const Params<float>& P);
// `prof` (optional): an event is appended before the first kernel and after every kernel, so
// consecutive events bracket k_prep, k_solve, k_prep, ... of the call.
// obs_env_stride / obs_rep_stride (floats; 0 = the (B, R, 2, 7) block: 14 R and 14): where env e, repeat r goes in `obs`.
cudaError_t launch_step_split(const Params<float>& P, float* f, int* id, SplitBuffers& b, const void* actions, float* obs,
                              float* reward, uint8_t* done, uint8_t* reason, StatsPtrs stats, cudaStream_t s,
                              int* launches, std::vector<cudaEvent_t>* prof, size_t obs_env_stride = 0, size_t obs_rep_stride = 0);

cudaError_t launch_fma_peak(int use_f64, int iters, double* tflops);
cudaError_t launch_compact(const uint8_t* flag, int n, int* list, int* count, cudaStream_t s);

}  // namespace bcp
