// bcp_launch_impl.cuh — definitions of Launch<T>; included once per scalar type.
#pragma once
#include "bcp_kernels.cuh"
#include "bcp_launch.cuh"

namespace bcp {

static inline int grid_for(int n) { return (n + kBlock - 1) / kBlock; }

template <typename T> size_t Launch<T>::state_bytes(int n_envs) { return (size_t)NF * (size_t)n_envs * sizeof(T); }

template <typename T>
cudaError_t Launch<T>::step(const Params<T>& P, void* f, int* id, const void* actions, float* obs, float* reward,
                            uint8_t* done, uint8_t* reason, StatsPtrs sp, cudaStream_t s) {
  const Stats st{sp.hist, sp.header, sp.reward_sum};
  const size_t smem = (size_t)kBlock * (14 * P.R + 1) * sizeof(float);
  if (smem > 48 * 1024) {
    cudaFuncSetAttribute(k_step<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_step<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  if (P.any_friction)
    k_step<T, true><<<grid_for(P.num_envs), kBlock, smem, s>>>(P, (T*)f, id, actions, obs, reward, done, reason, st);
  else
    k_step<T, false><<<grid_for(P.num_envs), kBlock, smem, s>>>(P, (T*)f, id, actions, obs, reward, done, reason, st);
  return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::reset(const Params<T>& P, void* f, int* id, const int* list, const int* count, int max_count,
                             float* obs, cudaStream_t s) {
  const int n = list ? max_count : P.num_envs;
  if (n <= 0) return cudaSuccess;
  if (P.any_friction) k_reset<T, true><<<grid_for(n), kBlock, 0, s>>>(P, (T*)f, id, list, count, obs);
  else k_reset<T, false><<<grid_for(n), kBlock, 0, s>>>(P, (T*)f, id, list, count, obs);
  return cudaGetLastError();
}

template <typename T> cudaError_t Launch<T>::substep(const Params<T>& P, void* f, int* id, cudaStream_t s) {
  if (P.any_friction) k_substep<T, true><<<grid_for(P.num_envs), kBlock, 0, s>>>(P, (T*)f, id);
  else k_substep<T, false><<<grid_for(P.num_envs), kBlock, 0, s>>>(P, (T*)f, id);
  return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::rollout(const Params<T>& P, void* f, int* id, int n_steps, uint32_t action_mask, float* obs,
                               StatsPtrs sp, cudaStream_t s) {
  const Stats st{sp.hist, sp.header, sp.reward_sum};
  if (P.any_friction) k_rollout<T, true><<<grid_for(P.num_envs), kBlock, 0, s>>>(P, (T*)f, id, n_steps, action_mask, obs, st);
  else k_rollout<T, false><<<grid_for(P.num_envs), kBlock, 0, s>>>(P, (T*)f, id, n_steps, action_mask, obs, st);
  return cudaGetLastError();
}

template <typename T> cudaError_t Launch<T>::export_state(int n, const void* f, const int* id, double* out, cudaStream_t s) {
  k_export<T><<<grid_for(n), kBlock, 0, s>>>(n, (const T*)f, id, out);
  return cudaGetLastError();
}
template <typename T> cudaError_t Launch<T>::import_state(int n, void* f, int* id, const double* in, cudaStream_t s) {
  k_import<T><<<grid_for(n), kBlock, 0, s>>>(n, (T*)f, id, in);
  return cudaGetLastError();
}

}  // namespace bcp
