// bcp_kernels.cuh — the __global__ kernels, templated on scalar and friction.
//
// Mapping: one env per lane, 32 consecutive envs per warp ("env group"), SoA state
// so every state access is a full-width coalesced transaction.  An env's whole
// env.step (R x S substeps) runs inside one launch with its state in the thread's
// registers / local frame; HBM sees one state read and one state write per call.
// Observations are staged per block in shared memory and streamed out as float4.
#pragma once
#include <cuda_runtime.h>
#include "bcp_env.cuh"
#include "bcp_state.cuh"

namespace bcp {

constexpr int kBlock = 128;

struct Stats {
  unsigned long long* hist;    // [max_len + 1]
  unsigned long long* header;  // episodes, env_steps, substeps, resets
  double* reward_sum;
};

// coalesced float4 copy-out of a block's staged observations.
// sobs: [nthreads][width+1] (padded rows), out: dense [n_valid][width]
__device__ __forceinline__ void flush_obs(const float* sobs, float* __restrict__ out, int n_valid, int width) {
  const int total = n_valid * width;  // floats; block base is 16-byte aligned (kBlock*width*4 bytes per block)
  const int nvec = total >> 2;
  for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
    float4 o;
    float* op = reinterpret_cast<float*>(&o);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = v * 4 + k;
      const int t = i / width, j = i - t * width;
      op[k] = sobs[t * (width + 1) + j];
    }
    reinterpret_cast<float4*>(out)[v] = o;
  }
  for (int i = (nvec << 2) + threadIdx.x; i < total; i += blockDim.x) {
    const int t = i / width, j = i - t * width;
    out[i] = sobs[t * (width + 1) + j];
  }
}

// BulletCartpole.step for every env of the shard (bullet_cartpole.py:178-275)
template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock) k_step(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id,
                                                 const void* __restrict__ actions, float* __restrict__ obs,
                                                 float* __restrict__ reward, uint8_t* __restrict__ done,
                                                 uint8_t* __restrict__ reason, Stats stats) {
  extern __shared__ float sobs[];
  const int N = P.num_envs;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int width = 14 * P.R;
  float* my = sobs + threadIdx.x * (width + 1);
  if (e < N) {
    Env<T> st;
    load_env<T>(f, id, N, e, st);
    if (st.done) {  // :179-181 - old observation, reward 0, done
      if (obs) for (int j = 0; j < width; ++j) my[j] = obs[(size_t)e * width + j];
      if (reward) reward[e] = 0.0f;
      if (done) done[e] = 1;
      if (reason) reason[e] = BCP_REASON_NONE;
    } else {
      int a = 0;
      float ax = 0.f, ay = 0.f;
      if (P.discrete) a = reinterpret_cast<const int32_t*>(actions)[e];
      else {
        const float2 av = reinterpret_cast<const float2*>(actions)[e];
        ax = av.x; ay = av.y;
      }
      T fx, fy;
      action_force(P, a, ax, ay, fx, fy);
      const int rs = env_step<T, FRICTION>(P, st, fx, fy, my);
      store_env<T>(f, id, N, e, st);
      if (reward) reward[e] = 1.0f;
      if (done) done[e] = (uint8_t)st.done;
      if (reason) reason[e] = (uint8_t)rs;
      if (st.done && stats.hist) {
        atomicAdd(&stats.hist[st.steps <= P.max_len ? st.steps : P.max_len], 1ull);
        atomicAdd(&stats.header[0], 1ull);
      }
    }
  }
  if (obs) {
    __syncthreads();
    const int base = blockIdx.x * blockDim.x;
    const int n_valid = min((int)blockDim.x, N - base);
    flush_obs(sobs, obs + (size_t)base * width, n_valid, width);
  }
}

// BulletCartpole.reset (bullet_cartpole.py:313-346) for the envs named by `list`
// (compacted: warps stay dense even when few envs reset) or, list == nullptr, all.
template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock) k_reset(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id,
                                                  const int* __restrict__ list, const int* __restrict__ count,
                                                  float* __restrict__ obs) {
  const int N = P.num_envs;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = list ? *count : N;
  if (i >= n) return;
  const int e = list ? list[i] : i;
  Env<T> st;
  load_env<T>(f, id, N, e, st);
  env_reset<T, FRICTION>(P, st, P.env_id_offset + e);
  store_env<T>(f, id, N, e, st);
  if (obs) {  // bootstrap: every repeat slot gets the same pose (:338-339)
    float o[14];
    write_obs(st, o);
    for (int r = 0; r < P.R; ++r) {
      float2* dst = reinterpret_cast<float2*>(obs + ((size_t)e * P.R + r) * 14);
#pragma unroll
      for (int j = 0; j < 7; ++j) dst[j] = make_float2(o[2 * j], o[2 * j + 1]);
    }
  }
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
template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock) k_substep(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id) {
  const int N = P.num_envs;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= N) return;
  Env<T> st;
  load_env<T>(f, id, N, e, st);
  substep<T, FRICTION>(P, st, nullptr);
  store_env<T>(f, id, N, e, st);
}

// README.md:76-93 random-agent protocol, n_steps env.steps per env in one launch:
// in-kernel Philox actions, finished episodes recorded and reset in place.
template <typename T, bool FRICTION>
__global__ void __launch_bounds__(kBlock) k_rollout(const __grid_constant__ Params<T> P, T* __restrict__ f, int* __restrict__ id,
                                                    int n_steps, uint32_t action_mask, float* __restrict__ obs, Stats stats) {
  const int N = P.num_envs;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= N) return;
  const int64_t genv = P.env_id_offset + e;
  Env<T> st;
  load_env<T>(f, id, N, e, st);
  unsigned long long n_eps = 0, n_resets = 0;
  for (int t = 0; t < n_steps; ++t) {
    if (st.done) {
      env_reset<T, FRICTION>(P, st, genv);
      n_resets++;
    }
    int a; float ax, ay;
    draw_action(P, genv, st, action_mask, a, ax, ay);
    T fx, fy;
    action_force(P, a, ax, ay, fx, fy);
    // only the last env.step's (R,2,7) observation is observable after the launch
    float* o = (obs && t == n_steps - 1) ? obs + (size_t)e * 14 * P.R : nullptr;
    env_step<T, FRICTION>(P, st, fx, fy, o);
    if (st.done) {
      atomicAdd(&stats.hist[st.steps <= P.max_len ? st.steps : P.max_len], 1ull);
      n_eps++;
    }
  }
  store_env<T>(f, id, N, e, st);
  // block-level reduction of the counters would be cheaper; these are 3 atomics per env per launch
  atomicAdd(&stats.header[0], n_eps);
  atomicAdd(&stats.header[1], (unsigned long long)n_steps);
  atomicAdd(&stats.header[2], (unsigned long long)n_steps * (unsigned long long)(P.R * P.S) +
                                  n_resets * (unsigned long long)(P.settle_steps + P.push_steps));
  atomicAdd(&stats.header[3], n_resets);
  atomicAdd(stats.reward_sum, (double)n_steps);
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
