// FP32 fast mode: FMA contraction allowed (compiled with default -fmad=true).
#include <cstdlib>
#include "bcp_launch_impl.cuh"
#include "bcp_split.cuh"
namespace bcp {
template struct Launch<float>;

size_t split_record_words(const Params<float>& P) { return (size_t)record_words(P); }
size_t split_counter_words(const Params<float>& P) { return (size_t)(P.R * P.S) * 3 * Q_WORDS; }  // main queue, generic queue, continuation counts
size_t split_cont_words(const Params<float>& P) { return (size_t)LAM_WORDS * (size_t)P.num_envs; }
size_t split_bins_words(const Params<float>& P) { return (size_t)N_BINS * (size_t)P.num_envs; }

bool split_uses_lambda(const Params<float>& P) {
  // Impulse-space solve phase wherever its two regimes cover the configuration.  Round 1 switched to row space from
  // 131 072 envs per handle (one pipeline: beyond a wave the matrix builds cost more than the shorter chain saves);
  // with the handle run as cohorts of <= 65 536 envs the impulse-space phase wins at every size measured
  // (131 072: 1.79e8 vs 1.56e8 env-steps/s, 262 144: 1.86e8 vs 1.63e8, 1 M: 1.91e8 vs 1.79e8).  BCP_LAMBDA=0 / 1 forces it.
  const char* le = getenv("BCP_LAMBDA");
  const bool wanted = le ? atoi(le) != 0 : true;
  return wanted && lam_config_ok(P);
}

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
  static int visit_rev = -1;
  if (visit_rev < 0) {
    const char* vr = getenv("BCP_VISIT_REV");
    visit_rev = vr ? atoi(vr) : 1;
  }
  vo.rev = visit_rev;
  // k_solve: persistent, as many blocks as stay resident (shared-memory columns bound it)
  LaunchShape L = shape_for(P, N, false);
  // lane-major columns of float4: stride / 4 odd => conflict-free 16-byte accesses
  const int col4 = (rec_words >> 2) | 1;  // a column holds the whole record (header + rows)
  const int col_stride = col4 * 4;
  L.smem = (size_t)col_stride * L.block * sizeof(float);
  err = allow_smem(k_solve<float, FRICTION>, L.smem);
  if (err != cudaSuccess) return err;
  err = allow_smem(k_solve_lam<float, FRICTION>, L.smem);
  if (err != cudaSuccess) return err;
  err = allow_smem(k_solve_lam_once<float, FRICTION>, L.smem);
  if (err != cudaSuccess) return err;
  err = allow_smem(k_solve_gen_once<float, FRICTION>, L.smem);
  if (err != cudaSuccess) return err;
  // Impulse-space solver (bcp_lambda.cuh) on the main queue.  It shortens the dependent chain of an
  // iteration, which is what bounds the solve phase while a launch is about one wave of envs
  // (measured on B200: +6..7 % env-steps/s at 32 768 and 65 536 envs, equal at 131 072); beyond
  // that the phase is issue-bound and building the matrices costs more than it saves (-3 % at
  // 262 144, -8 % at 1 M).  BCP_LAMBDA=1 / 0 forces it on / off.
  const int use_lam = ((b.lam < 0 ? split_uses_lambda(P) : b.lam != 0) && b.cont) ? 1 : 0;
  static int handoff = -1;
  if (handoff < 0) {
    const char* he = getenv("BCP_HANDOFF");
    handoff = he ? atoi(he) : 8;  // iterations after which a solve moves to k_solve_tail (0: never)
  }
  static int lam_gen = -1;
  if (lam_gen < 0) {
    const char* lg = getenv("BCP_LAM_GEN");
    lam_gen = lg ? atoi(lg) : 1;
  }
  static int lam_once = -1;
  if (lam_once < 0) {
    const char* lo = getenv("BCP_LAM_ONCE");
    lam_once = lo ? atoi(lo) : 1;
  }
  int tail_grid = (N + kBlock - 1) / kBlock;
  if (tail_grid > 2 * 148) tail_grid = 2 * 148;
  static int side_env = -1;
  if (side_env < 0) {
    const char* se = getenv("BCP_SIDE");
    side_env = se ? atoi(se) : 1;
  }
  // The few envs of the generic queue (pole touching cart AND ground) still solve row by row, and one
  // that needs all 50 iterations takes as long as the whole impulse-space phase: their k_solve gets
  // a handful of blocks on a high-priority side stream and runs beside k_solve_lam / k_solve_tail.
  const bool side = use_lam && side_env;
  if (side && !b.side) {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    err = cudaStreamCreateWithPriority(&b.side, cudaStreamNonBlocking, hi);
    if (err != cudaSuccess) return err;
    cudaEventCreateWithFlags(&b.ev_prep, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&b.ev_gen, cudaEventDisableTiming);
  }
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
  const int gen_grid = side ? min(solve_grid, 16) : solve_grid;
  const int gen_once_grid = min((N + L.block - 1) / L.block, 4 * sms);
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
    int* qc = b.qc + (size_t)(k < RS ? k : 0) * 3 * Q_WORDS;
    if (s_set >= 0)
      k_prep<float, FRICTION><<<prep_grid, prep_block, prep_smem, s>>>(P, f, id, b.recs, rec_words, reinterpret_cast<int2*>(b.queue), qc, io, st, vo, s_fin, s_set, use_lam);
    else  // the closing kernel: the lean instance (no frame, four blocks per SM)
      k_prep<float, FRICTION, false><<<(N + kBlock - 1) / kBlock, kBlock, 0, s>>>(P, f, id, b.recs, rec_words, reinterpret_cast<int2*>(b.queue), qc, io, st, vo, s_fin, -1, use_lam);
    if (launches) ++*launches;
    if (b.after_prep) b.after_prep(b.hook_ctx, k, s);
    mark();
    if (k < RS) {
      if (use_lam) {
        if (side) {
          cudaEventRecord(b.ev_prep, s);
          cudaStreamWaitEvent(b.side, b.ev_prep, 0);
          if (lam_gen)  // impulse space for these too: half the latency per iteration of solves that mostly run all 50
            // one warp per 32 queued envs; blocks past the queue's end leave at once.  Normally a block or two have work; a batch
            // whose envs were all reset together sends a wave of them through this regime at the same time (poles falling off
            // the carts: tens of thousands of 50-iteration solves in the same calls), which 16 blocks took 20 ms to work off.
            k_solve_gen_once<float, FRICTION><<<gen_once_grid, L.block, L.smem, b.side>>>(P, b.recs, rec_words, reinterpret_cast<const int2*>(b.queue) + N, qc + Q_WORDS, col_stride);
          else
            k_solve<float, FRICTION><<<gen_grid, L.block, L.smem, b.side>>>(P, b.recs, rec_words, reinterpret_cast<const int2*>(b.queue) + N, qc + Q_WORDS, col_stride, fetch_min);
          cudaEventRecord(b.ev_gen, b.side);
        }
        if (handoff > 0 && lam_once)  // at most `handoff` iterations per solve here: one chunk of the queue per warp, no pool
          k_solve_lam_once<float, FRICTION><<<(N + L.block - 1) / L.block, L.block, L.smem, s>>>(P, b.recs, rec_words, reinterpret_cast<const int2*>(b.queue), qc,
                                                                                                 col_stride, b.cont, reinterpret_cast<int4*>(b.contm), handoff);
        else
          k_solve_lam<float, FRICTION><<<solve_grid, L.block, L.smem, s>>>(P, b.recs, rec_words, reinterpret_cast<const int2*>(b.queue), qc, col_stride, fetch_min,
                                                                            b.cont, reinterpret_cast<int4*>(b.contm), handoff);
        if (handoff > 0) {
          k_solve_tail<float, FRICTION><<<tail_grid, kBlock, 0, s>>>(P, b.recs, rec_words, qc, b.cont, reinterpret_cast<const int4*>(b.contm));
          if (launches) ++*launches;
        }
        if (side) cudaStreamWaitEvent(s, b.ev_gen, 0);  // the next k_prep finishes these envs too
        else k_solve<float, FRICTION><<<gen_grid, L.block, L.smem, s>>>(P, b.recs, rec_words, reinterpret_cast<const int2*>(b.queue) + N, qc + Q_WORDS, col_stride, fetch_min);
        if (launches) ++*launches;
      } else {
        k_solve<float, FRICTION><<<solve_grid, L.block, L.smem, s>>>(P, b.recs, rec_words, reinterpret_cast<const int2*>(b.queue), qc, col_stride, fetch_min);
      }
      if (launches) ++*launches;
      if (b.trace)  // instrumentation of the parity tests: what every env's solve of substep k did
        cudaMemcpy2DAsync(b.trace + (size_t)k * (b.trace_stride > 0 ? b.trace_stride : N) * 4, 4 * sizeof(int), b.recs, (size_t)rec_words * sizeof(float), 4 * sizeof(int), (size_t)N,
                          cudaMemcpyDeviceToDevice, s);
      mark();
    }
  }
  b.cur = nxt;
  b.side_ready = 1;
  return cudaGetLastError();
}
cudaError_t launch_step_split(const Params<float>& P, float* f, int* id, SplitBuffers& b, const void* actions, float* obs,
                              float* reward, uint8_t* done, uint8_t* reason, StatsPtrs sp, cudaStream_t s, int* launches,
                              std::vector<cudaEvent_t>* prof, size_t obs_env_stride, size_t obs_rep_stride) {
  const Stats st{sp.hist, sp.header, sp.reward_sum};
  const SplitIO io{actions, obs, reward, done, reason, obs_env_stride ? obs_env_stride : (size_t)14 * P.R, obs_rep_stride ? obs_rep_stride : (size_t)14,
                   b.roll_max_episodes, b.rollout};
  if (P.any_friction) return step_split_impl<true>(P, f, id, b, io, st, s, launches, prof);
  return step_split_impl<false>(P, f, id, b, io, st, s, launches, prof);
}

cudaError_t launch_draw_actions(const Params<float>& P, const int* id, void* actions, uint32_t mask, cudaStream_t s) {
  k_draw_actions<float><<<(P.num_envs + 255) / 256, 256, 0, s>>>(P, id, actions, mask);
  return cudaGetLastError();
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
