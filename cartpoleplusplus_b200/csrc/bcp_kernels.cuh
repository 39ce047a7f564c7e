// bcp_kernels.cuh — the __global__ kernels, templated on scalar, friction and path.
//
// Mapping: one env per lane, 32 consecutive envs per warp ("env group"), SoA state so
// every state access is a full-width coalesced transaction.  An env's whole env.step
// (R x S substeps) runs inside one launch; HBM sees one state read and one state write
// per call.  FAST = true is the fast path of bcp_fast.cuh: the constraint rows of the
// block's envs live in dynamic shared memory, one word-major column per env
// ([col_words][blockDim], bank = lane, conflict-free).  FAST = false is the parity path
// (bit-exact with the oracle in fp64).  Observations are either staged per block in shared
// memory and streamed out as float4, or (stage_obs = 0) stored directly as float2.
#pragma once
#include <cuda_runtime.h>
#include "bcp_fast.cuh"
#include "bcp_state.cuh"

namespace bcp {

constexpr int kBlock = 128;

struct Stats {
  unsigned long long* hist;    // [max_len + 1]
  unsigned long long* header;  // episodes, env_steps, substeps, resets
  double* reward_sum;
};

extern __shared__ __align__(16) unsigned char bcp_smem[];

template <typename T, bool FRICTION, bool FAST> struct Path;
template <typename T, bool FRICTION> struct Path<T, FRICTION, false> {
  typedef StepParity<T, FRICTION> Step;
  static __device__ __forceinline__ Step make(const Params<T>&) { return Step(); }
  static __device__ __forceinline__ size_t col_bytes(const Params<T>&) { return 0; }
};
template <typename T, bool FRICTION> struct Path<T, FRICTION, true> {
  typedef StepFast<T, FRICTION> Step;
  static __device__ __forceinline__ Step make(const Params<T>&) {
    Step s;
    s.col.base = reinterpret_cast<T*>(bcp_smem) + threadIdx.x;
    s.col.stride = blockDim.x;
    return s;
  }
  static __device__ __forceinline__ size_t col_bytes(const Params<T>& P) {
    return (size_t)P.col_words * blockDim.x * sizeof(T);
  }
};

// straight float4 copy-out of a block's staged observations (sobs is the dense layout)
__device__ __forceinline__ void flush_obs(const float* sobs, float* __restrict__ out, int total) {
  const int nvec = total >> 2;  // block base is 16-byte aligned (blockDim*width*4 bytes per block)
  for (int v = threadIdx.x; v < nvec; v += blockDim.x)
    reinterpret_cast<float4*>(out)[v] = reinterpret_cast<const float4*>(sobs)[v];
  for (int i = (nvec << 2) + threadIdx.x; i < total; i += blockDim.x) out[i] = sobs[i];
}

// BulletCartpole.step for every env of the shard (bullet_cartpole.py:178-275); with
// P.auto_reset == BCP_AUTO_RESET_PIPELINED one call of env_advance per env instead.
template <typename T, bool FRICTION, bool FAST>
__global__ void __launch_bounds__(kBlock) k_step(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id,
                                                 const void* __restrict__ actions, float* __restrict__ obs,
                                                 float* __restrict__ reward, uint8_t* __restrict__ done,
                                                 uint8_t* __restrict__ reason, Stats stats, int stage_obs) {
  typedef Path<T, FRICTION, FAST> PathT;
  const typename PathT::Step sub = PathT::make(P);
  float* sobs = reinterpret_cast<float*>(bcp_smem + PathT::col_bytes(P));
  const int N = P.num_envs;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int width = 14 * P.R;
  float* my = nullptr;
  if (obs) my = stage_obs ? sobs + threadIdx.x * width : obs + (size_t)e * width;
  unsigned n_live = 0;
  // every lane of the warp goes through the env program (fast path: warp-wide votes); lanes
  // past the end of the shard or with nothing to do carry `on` == false
  const bool valid = e < N;
  Env<T> st;
  st.done = 1;
  st.reset_left = 0;
  if (valid) load_env<T>(f, id, N, e, st);
  // an env that does not step keeps its observation (:179-181; pipelined: while it resets)
  if (valid && st.done && obs && stage_obs) for (int j = 0; j < width; ++j) my[j] = obs[(size_t)e * width + j];
  int a = 0;
  float ax = 0.f, ay = 0.f;
  if (valid) {
    if (P.discrete) a = reinterpret_cast<const int32_t*>(actions)[e];
    else {
      const float2 av = reinterpret_cast<const float2*>(actions)[e];
      ax = av.x; ay = av.y;
    }
  }
  T fx, fy;
  action_force(P, a, ax, ay, fx, fy);
  bool live;
  const int rs = env_call(P, st, P.env_id_offset + e, fx, fy, my, sub, valid, live);
  if (valid) {
    store_env<T>(f, id, N, e, st);
    if (reward) reward[e] = live ? 1.0f : 0.0f;
    if (done) done[e] = (uint8_t)st.done;
    if (reason) reason[e] = (uint8_t)rs;
    if (live) {
      n_live = 1;
      if (st.done && stats.hist) {
        atomicAdd(&stats.hist[st.steps <= P.max_len ? st.steps : P.max_len], 1ull);
        atomicAdd(&stats.header[0], 1ull);
      }
    }
  }
  if (stats.header) {  // env.steps actually taken (reward 1) - what throughput counts
    const unsigned w = __popc(__ballot_sync(0xffffffffu, n_live != 0));
    if ((threadIdx.x & 31) == 0 && w) atomicAdd(&stats.header[1], (unsigned long long)w);
  }
  if (obs && stage_obs) {
    __syncthreads();
    const int base = blockIdx.x * blockDim.x;
    const int n_valid = min((int)blockDim.x, N - base);
    flush_obs(sobs, obs + (size_t)base * width, n_valid * width);
  }
}

// BulletCartpole.reset (bullet_cartpole.py:313-346) for the envs named by `list`
// (compacted: warps stay dense even when few envs reset) or, list == nullptr, all.
template <typename T, bool FRICTION, bool FAST>
__global__ void __launch_bounds__(kBlock) k_reset(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id,
                                                  const int* __restrict__ list, const int* __restrict__ count,
                                                  float* __restrict__ obs) {
  typedef Path<T, FRICTION, FAST> PathT;
  const typename PathT::Step sub = PathT::make(P);
  const int N = P.num_envs;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = list ? *count : N;
  const bool on = i < n;
  const int e = on ? (list ? list[i] : i) : 0;
  Env<T> st;
  if (on) load_env<T>(f, id, N, e, st);
  env_reset(P, st, P.env_id_offset + e, sub, on);
  if (!on) return;
  store_env<T>(f, id, N, e, st);
  if (obs)  // bootstrap: every repeat slot gets the same pose (:338-339)
    for (int r = 0; r < P.R; ++r) write_obs(st, obs + ((size_t)e * P.R + r) * 14);
}

// list <- indices e with flag[e] != 0 (warp-aggregated append; order irrelevant)
static __global__ void k_compact(const uint8_t* __restrict__ flag, int n, int* __restrict__ list, int* __restrict__ count) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const bool on = (e < n) && flag[e];
  const unsigned m = __ballot_sync(0xffffffffu, on);
  if (m == 0) return;
  const int lane = threadIdx.x & 31;
  int base = 0;
  if (lane == __ffs(m) - 1) base = atomicAdd(count, __popc(m));
  base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
  if (on) list[base + __popc(m & ((1u << lane) - 1))] = e;
}

// bare p.stepSimulation() for every env (single-substep parity tests)
template <typename T, bool FRICTION, bool FAST>
__global__ void __launch_bounds__(kBlock) k_substep(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id) {
  typedef Path<T, FRICTION, FAST> PathT;
  const typename PathT::Step sub = PathT::make(P);
  const int N = P.num_envs;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const bool on = e < N;
  Env<T> st;
  if (on) load_env<T>(f, id, N, e, st);
  sub(P, st, on);
  if (on) store_env<T>(f, id, N, e, st);
}

// README.md:76-93 random-agent protocol, n_calls rounds of the pipelined machine in one
// launch: in-kernel Philox actions, finished episodes recorded, resets run in place R*S
// substeps per round (every lane runs one substep per inner iteration whatever its phase).
template <typename T, bool FRICTION, bool FAST>
__global__ void __launch_bounds__(kBlock) k_rollout(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id,
                                                    int n_calls, uint32_t action_mask, int max_episodes,
                                                    float* __restrict__ obs, Stats stats) {
  typedef Path<T, FRICTION, FAST> PathT;
  const typename PathT::Step sub = PathT::make(P);
  const int N = P.num_envs;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long n_eps = 0, n_live = 0, n_sub = 0, n_resets = 0;
  const bool valid = e < N;
  const int64_t genv = P.env_id_offset + e;
  Env<T> st;
  st.done = 1;
  st.reset_left = 0;
  st.episode = 0;
  st.steps = 0;
  if (valid) load_env<T>(f, id, N, e, st);
  float* o = (obs && valid) ? obs + (size_t)e * 14 * P.R : nullptr;
  for (int t = 0; t < n_calls; ++t) {
    // st.episode counts finished resets: an env that is done with episode == max_episodes has
    // delivered its share of the protocol's episodes and idles
    const bool on = valid && !(max_episodes > 0 && st.done && st.reset_left == 0 && st.episode >= max_episodes);
    if (!__any_sync(0xffffffffu, on)) break;
    int a = 0; float ax = 0.f, ay = 0.f;
    if (on) draw_action(P, genv, st, action_mask, a, ax, ay);
    T fx, fy;
    action_force(P, a, ax, ay, fx, fy);
    bool live;
    const int rs = env_call(P, st, genv, fx, fy, o, sub, on, live);
    if (on) {
      n_sub += (unsigned long long)(P.R * P.S);
      if (live) {
        n_live++;
        if (st.done) {
          atomicAdd(&stats.hist[st.steps <= P.max_len ? st.steps : P.max_len], 1ull);
          n_eps++;
        }
      } else if (rs == BCP_REASON_FRESH) {
        n_resets++;
      }
    }
  }
  if (valid) store_env<T>(f, id, N, e, st);
  // warp-reduced counters: one atomic per warp per counter
  for (int off = 16; off > 0; off >>= 1) {
    n_eps += __shfl_down_sync(0xffffffffu, n_eps, off);
    n_live += __shfl_down_sync(0xffffffffu, n_live, off);
    n_sub += __shfl_down_sync(0xffffffffu, n_sub, off);
    n_resets += __shfl_down_sync(0xffffffffu, n_resets, off);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&stats.header[0], n_eps);
    atomicAdd(&stats.header[1], n_live);
    atomicAdd(&stats.header[2], n_sub);
    atomicAdd(&stats.header[3], n_resets);
    atomicAdd(stats.reward_sum, (double)n_live);
  }
}

// canonical record export / import (device staging buffer of doubles)
template <typename T>
__global__ void k_export(int N, const T* __restrict__ f, const int* __restrict__ id, double* __restrict__ out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= N) return;
  Env<T> st;
  load_env<T>(f, id, N, e, st);
  env_to_canonical<T>(st, out + (size_t)e * BCP_STATE_DOUBLES);
}
template <typename T>
__global__ void k_import(int N, T* __restrict__ f, int* __restrict__ id, const double* __restrict__ in) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= N) return;
  Env<T> st;
  env_from_canonical<T>(in + (size_t)e * BCP_STATE_DOUBLES, st);
  store_env<T>(f, id, N, e, st);
}

}  // namespace bcp
