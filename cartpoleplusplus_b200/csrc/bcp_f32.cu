// FP32 fast mode: FMA contraction allowed (compiled with default -fmad=true).
#include "bcp_launch_impl.cuh"
namespace bcp {
template struct Launch<float>;
cudaError_t launch_compact(const uint8_t* flag, int n, int* list, int* count, cudaStream_t s) {
  cudaError_t err = cudaMemsetAsync(count, 0, sizeof(int), s);
  if (err != cudaSuccess) return err;
  k_compact<<<(n + 255) / 256, 256, 0, s>>>(flag, n, list, count);
  return cudaGetLastError();
}
}  // namespace bcp
