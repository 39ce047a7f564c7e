// FP32 fast mode: FMA contraction allowed (compiled with default -fmad=true).
#include <cstdlib>
#include "bcp_launch_impl.cuh"
#include "bcp_split.cuh"
namespace bcp {
template struct Launch<float>;

size_t split_record_words(const Params<float>& P) { return (size_t)record_words(P); }
size_t split_counter_words(const Params<float>& P) { return (size_t)(P.R * P.S) * Q_WORDS; }
size_t split_bins_words(const Params<float>& P) { return (size_t)N_BINS * (size_t)P.num_envs; }

template <bool FRICTION>
static cudaError_t step_split_impl(const Params<float>& P, float* f, int* id, SplitBuffers& b, const SplitIO& io,
                                   const Stats& st, cudaStream_t s, int* launches, std::vector<cudaEvent_t>* prof) {
  auto mark = [&]() {
    if (!prof) return;
    cudaEvent_t ev;
    cudaEventCreate(&ev);
    cudaEventRecord(ev, s);
    prof->push_back(ev);
  };
  const int RS = P.R * P.S, N = P.num_envs;
  const int rec_words = (int)split_record_words(P);
  cudaError_t err = cudaMemsetAsync(b.qc, 0, split_counter_words(P) * sizeof(int), s);
  if (err != cudaSuccess) return err;
  const int nxt = b.cur < 0 ? 0 : 1 - b.cur;
  err = cudaMemsetAsync(b.cnts[nxt], 0, N_BINS * sizeof(int), s);
  if (err != cudaSuccess) return err;
  VisitOrder vo;
  vo.bins = b.cur < 0 ? nullptr : b.bins[b.cur];
  vo.cnt = b.cur < 0 ? nullptr : b.cnts[b.cur];
  vo.next_bins = b.bins[nxt];
  vo.next_cnt = b.cnts[nxt];
  // k_solve: persistent, as many blocks as stay resident (shared-memory columns bound it)
  LaunchShape L = shape_for(P, N, false);
  // lane-major columns of float4: stride / 4 odd => conflict-free 16-byte accesses
  const int col4 = (rec_words >> 2) | 1;  // a column holds the whole record (header + rows)
  const int col_stride = col4 * 4;
  L.smem = (size_t)col_stride * L.block * sizeof(float);
  err = allow_smem(k_solve<float, FRICTION>, L.smem);
  if (err != cudaSuccess) return err;
  static int sms = 0, per_sm = 0;
  static size_t per_sm_smem = (size_t)-1;
  if (!sms || per_sm_smem != L.smem) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_solve<float, FRICTION>, L.block, L.smem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) per_sm = 1;
    per_sm_smem = L.smem;
  }
  static int fetch_min = 0;
  if (!fetch_min) {
    const char* fm = getenv("BCP_FETCH_MIN");
    fetch_min = fm ? atoi(fm) : kFetchMin;
    if (fetch_min < 1) fetch_min = 1;
    if (fetch_min > 32) fetch_min = 32;
  }
  const int solve_grid = min(sms * per_sm, (N + L.block - 1) / L.block);
  static int prep_block = 0;
  if (!prep_block) {
    const char* pb = getenv("BCP_PREP_BLOCK");
    prep_block = pb ? atoi(pb) : kBlock;
    if (prep_block < 32 || prep_block > kBlock || prep_block % 32) prep_block = kBlock;
  }
  const int prep_grid = (N + prep_block - 1) / prep_block;
  const size_t prep_smem = prep_smem_bytes(prep_block);
  err = allow_smem(k_prep<float, FRICTION>, prep_smem);
  if (err != cudaSuccess) return err;
  mark();
  for (int k = 0; k <= RS; ++k) {
    const int s_fin = k - 1, s_set = (k < RS) ? k : -1;
    int* qc = b.qc + (size_t)(k < RS ? k : 0) * Q_WORDS;
    k_prep<float, FRICTION><<<prep_grid, prep_block, prep_smem, s>>>(P, f, id, b.recs, rec_words, reinterpret_cast<int2*>(b.queue), qc, io, st, vo, s_fin, s_set);
    if (launches) ++*launches;
    mark();
    if (k < RS) {
      k_solve<float, FRICTION><<<solve_grid, L.block, L.smem, s>>>(P, b.recs, rec_words, reinterpret_cast<const int2*>(b.queue), qc, col_stride, fetch_min);
      if (launches) ++*launches;
      mark();
    }
  }
  b.cur = nxt;
  return cudaGetLastError();
}
cudaError_t launch_step_split(const Params<float>& P, float* f, int* id, SplitBuffers& b, const void* actions, float* obs,
                              float* reward, uint8_t* done, uint8_t* reason, StatsPtrs sp, cudaStream_t s, int* launches,
                              std::vector<cudaEvent_t>* prof) {
  const Stats st{sp.hist, sp.header, sp.reward_sum};
  const SplitIO io{actions, obs, reward, done, reason};
  if (P.any_friction) return step_split_impl<true>(P, f, id, b, io, st, s, launches, prof);
  return step_split_impl<false>(P, f, id, b, io, st, s, launches, prof);
}

cudaError_t launch_compact(const uint8_t* flag, int n, int* list, int* count, cudaStream_t s) {
  cudaError_t err = cudaMemsetAsync(count, 0, sizeof(int), s);
  if (err != cudaSuccess) return err;
  k_compact<<<(n + 255) / 256, 256, 0, s>>>(flag, n, list, count);
  return cudaGetLastError();
}

// FMA-chain microbenchmark: 8 independent accumulators per thread, no memory traffic.
template <typename T> __global__ void k_fma_peak(T* out, int iters, T a, T b) {
  T x0 = a, x1 = a + T(1), x2 = a + T(2), x3 = a + T(3), x4 = a + T(4), x5 = a + T(5), x6 = a + T(6), x7 = a + T(7);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      x0 = fma(x0, b, a); x1 = fma(x1, b, a); x2 = fma(x2, b, a); x3 = fma(x3, b, a);
      x4 = fma(x4, b, a); x5 = fma(x5, b, a); x6 = fma(x6, b, a); x7 = fma(x7, b, a);
    }
  }
  const T s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (s == T(12345.678)) out[0] = s;  // keeps the chains alive
}
cudaError_t launch_fma_peak(int use_f64, int iters, double* tflops) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  void* out = nullptr;
  cudaError_t err = cudaMalloc(&out, 64);
  if (err != cudaSuccess) return err;
  const int blocks = sms * 8, threads = 256;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    if (use_f64) k_fma_peak<double><<<blocks, threads>>>((double*)out, iters, 0.5, 0.999);
    else k_fma_peak<float><<<blocks, threads>>>((float*)out, iters, 0.5f, 0.999f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  err = cudaGetLastError();
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  const double flops = 2.0 * 64.0 * (double)iters * (double)blocks * (double)threads;
  *tflops = flops / (best * 1e-3) / 1e12;
  return err;
}
}  // namespace bcp
