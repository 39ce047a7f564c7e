// Issue cost of FFMA2 (fma.rn.f32x2) against FFMA on sm_100a: N independent accumulators per thread updated with one
// broadcast scalar (the  res[k] -= A[k] * d  shape of the impulse-space solver), one warp per scheduler and full occupancy.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_bench ffma2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b),
                     rc = *reinterpret_cast<unsigned long long*>(&c), rd;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  return *reinterpret_cast<float2*>(&rd);
}
template <bool PACKED, int CHAIN> __global__ void k(float* x, int iters, long long* cyc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float a[12], r[12];
  for (int j = 0; j < 12; ++j) { a[j] = x[i * 24 + j]; r[j] = x[i * 24 + 12 + j]; }
  float d = x[i * 24];
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (PACKED) {
#pragma unroll
      for (int j = 0; j < 12; j += 2) {
        const float2 v = ffma2(make_float2(a[j], a[j + 1]), make_float2(d, d), make_float2(r[j], r[j + 1]));
        r[j] = v.x; r[j + 1] = v.y;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 12; ++j) r[j] = fmaf(a[j], d, r[j]);
    }
    if (CHAIN) {  // a dependent tail like the row's clamp: d for the next round comes out of r[it % 12]
      d = fmaxf(r[0] * 0.999f, -1.0f) + 1e-3f;
    }
  }
  const long long t1 = clock64();
  float s = d;
  for (int j = 0; j < 12; ++j) s += r[j];
  x[i * 24] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <bool PACKED, int CHAIN> void run(const char* name, int blocks, int threads, float* x, long long* cyc) {
  const int iters = 4096;
  k<PACKED, CHAIN><<<blocks, threads>>>(x, iters, cyc);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<PACKED, CHAIN><<<blocks, threads>>>(x, iters, cyc);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-28s blocks %4d x %4d: %.3f ms, %.2f cycles per round of 12 FMAs (block 0 warp 0)\n", name, blocks, threads, ms, (double)c / iters);
}
int main() {
  float* x; long long* cyc;
  cudaMalloc(&x, 148 * 8 * 1024 * 24 * 4); cudaMemset(x, 0, 148 * 8 * 1024 * 24 * 4); cudaMalloc(&cyc, 8);
  run<false, 0>("FFMA  independent", 148, 128, x, cyc);
  run<true, 0>("FFMA2 independent", 148, 128, x, cyc);
  run<false, 1>("FFMA  + dependent tail", 148, 128, x, cyc);
  run<true, 1>("FFMA2 + dependent tail", 148, 128, x, cyc);
  run<false, 0>("FFMA  independent, full occ", 148 * 2, 1024, x, cyc);
  run<true, 0>("FFMA2 independent, full occ", 148 * 2, 1024, x, cyc);
  return 0;
}
