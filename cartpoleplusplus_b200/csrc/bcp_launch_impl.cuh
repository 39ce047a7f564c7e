// bcp_launch_impl.cuh — definitions of Launch<T>; included once per scalar type.
// float  -> fast path (bcp_fast.cuh, constraint rows in shared memory)
// double -> parity path (bit-exact with the oracle)
#pragma once
#include <cstdlib>
#include <type_traits>
#include "bcp_kernels.cuh"
#include "bcp_launch.cuh"

namespace bcp {

template <typename T> struct IsFast { static constexpr bool value = std::is_same<T, float>::value; };

struct LaunchShape {
  int block, grid, stage_obs;
  size_t smem;
};

// Block size and dynamic shared memory of a kernel that runs substeps.  Fast path: one
// column of P.col_words words per thread; the observation staging area is added only when
// it does not cost a resident block.
template <typename T> static LaunchShape shape_for(const Params<T>& P, int n, bool wants_obs) {
  static int forced = -1;
  if (forced < 0) {
    const char* s = getenv("BCP_BLOCK");
    forced = s ? atoi(s) : 0;
  }
  LaunchShape L;
  L.block = (forced >= 32 && forced <= kBlock && forced % 32 == 0) ? forced : kBlock;
  const size_t cols = IsFast<T>::value ? (size_t)P.col_words * L.block * sizeof(T) : 0;
  const size_t stage = (size_t)L.block * 14 * P.R * sizeof(float);
  const size_t sm_bytes = 227 * 1024, reserve = 1024;
  auto resident = [&](size_t b) { return b + reserve > sm_bytes ? (size_t)0 : sm_bytes / (b + reserve); };
  L.stage_obs = (wants_obs && resident(cols + stage) > 0 && resident(cols + stage) == resident(cols ? cols : 1)) ? 1 : 0;
  if (!IsFast<T>::value && wants_obs && stage <= 64 * 1024) L.stage_obs = 1;
  L.smem = cols + (L.stage_obs ? stage : 0);
  L.grid = (n + L.block - 1) / L.block;
  return L;
}
template <typename K> static cudaError_t allow_smem(K kernel, size_t smem) {
  if (smem <= 48 * 1024) return cudaSuccess;
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}
#define BCP_LAUNCH(KERNEL, L, s, ...)                                                   \
  do {                                                                                  \
    cudaError_t _e;                                                                     \
    if (P.any_friction) {                                                               \
      _e = allow_smem(KERNEL<T, true, IsFast<T>::value>, (L).smem);                     \
      if (_e != cudaSuccess) return _e;                                                 \
      KERNEL<T, true, IsFast<T>::value><<<(L).grid, (L).block, (L).smem, s>>>(__VA_ARGS__);  \
    } else {                                                                            \
      _e = allow_smem(KERNEL<T, false, IsFast<T>::value>, (L).smem);                    \
      if (_e != cudaSuccess) return _e;                                                 \
      KERNEL<T, false, IsFast<T>::value><<<(L).grid, (L).block, (L).smem, s>>>(__VA_ARGS__); \
    }                                                                                   \
    return cudaGetLastError();                                                          \
  } while (0)

static inline int grid_for(int n) { return (n + kBlock - 1) / kBlock; }

template <typename T> size_t Launch<T>::state_bytes(int n_envs) { return (size_t)NF * (size_t)n_envs * sizeof(T); }

template <typename T>
cudaError_t Launch<T>::step(const Params<T>& P, void* f, int* id, const void* actions, float* obs, float* reward,
                            uint8_t* done, uint8_t* reason, StatsPtrs sp, cudaStream_t s) {
  const Stats st{sp.hist, sp.header, sp.reward_sum};
  const LaunchShape L = shape_for(P, P.num_envs, obs != nullptr);
  BCP_LAUNCH(k_step, L, s, P, (T*)f, id, actions, obs, reward, done, reason, st, L.stage_obs);
}

template <typename T>
cudaError_t Launch<T>::reset(const Params<T>& P, void* f, int* id, const int* list, const int* count, int max_count,
                             float* obs, cudaStream_t s) {
  const int n = list ? max_count : P.num_envs;
  if (n <= 0) return cudaSuccess;
  const LaunchShape L = shape_for(P, n, false);
  BCP_LAUNCH(k_reset, L, s, P, (T*)f, id, list, count, obs);
}

template <typename T> cudaError_t Launch<T>::substep(const Params<T>& P, void* f, int* id, cudaStream_t s) {
  const LaunchShape L = shape_for(P, P.num_envs, false);
  BCP_LAUNCH(k_substep, L, s, P, (T*)f, id);
}

template <typename T>
cudaError_t Launch<T>::rollout(const Params<T>& P, void* f, int* id, int n_calls, uint32_t action_mask,
                               int max_episodes, float* obs, StatsPtrs sp, cudaStream_t s) {
  const Stats st{sp.hist, sp.header, sp.reward_sum};
  const LaunchShape L = shape_for(P, P.num_envs, false);
  BCP_LAUNCH(k_rollout, L, s, P, (T*)f, id, n_calls, action_mask, max_episodes, obs, st);
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
