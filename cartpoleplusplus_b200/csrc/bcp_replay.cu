// bcp_replay.cu — device-side replay memory (include/bcp_replay.h): the reference's
// ReplayMemory (replay_memory.py:11-165) as an HBM-resident ring + float16 state pool + FIFO of
// free pool slots, fed straight from the batched env step.
//
// The reference's add_episode is sequential (one pop of the free list per state, one or two
// pushes per evicted row).  Everything it decides is a function of POSITION in that sequence,
// so a commit of K finished episodes / T transitions is planned with prefix sums instead:
//   transition g (0..T-1, episodes in env-id order) lands in row (insert + g) mod N;
//   it is preceded by  k + 1 + g  pops (k = its episode's rank): its state_2 slot is entry
//   head + k + 1 + g of the free-slot stream, its state_1 slot the entry before that (the
//   previous transition's state_2, or the episode's initial-state pop);
//   if the ring is full at g it first pushes the evicted row's state_1 slot and, when that row
//   was terminal, its state_2 slot (replay_memory.py:83-93): the push offsets are a prefix sum
//   of 1 + (terminal_mask_old == 0) over g.
// The free list is a circular array indexed by absolute stream position.  A pass handles at
// most N transitions (so that every evicted row was written before the pass); commits larger
// than the ring run several passes.  All index work is integer and exact; states are rounded
// float32 -> float16 (round to nearest even) exactly as numpy's assignment does.
//
// HBM-bound byte work: per transition the commit reads 2 W + 4 (A + 1) staged bytes and writes
// 2 W + 4 (A + 4) ring/pool bytes; observe reads 4 W + 4 A + 6 and writes 2 W + 4 A + 4 per env.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>
#include <new>
#include <type_traits>

#include "../../include/bcp_replay.h"
#include "bcp_math.cuh"

namespace bcp {
int abi_fail(const char* what, cudaError_t e);  // bcp_abi.cu: sets bcp_last_error()
}
using bcp::abi_fail;

namespace {

#define RCK(call)                                      \
  do {                                                 \
    cudaError_t _e = (call);                           \
    if (_e != cudaSuccess) return abi_fail(#call, _e); \
  } while (0)

typedef long long i64;
typedef unsigned long long u64;

// device control block
enum { C_INSERT = 0, C_FULL, C_HEAD, C_TAIL, C_N_ADD_EPISODE, C_N_ADD, C_N_EVICTED_S2, C_ERROR, C_K, C_T, C_PB0, C_PB1, C_WORDS = 16 };

struct View {
  // ring (replay_memory.py:20-24) and pool (:32-33)
  int* s1_idx; int* s2_idx; float* action; float* reward; float* mask; __half* state;
  int* fq;  // free pool slots, circular by absolute stream position (:36)
  i64* ctl;
  // per-env staging of the open episode (float16 states: what the pool will hold)
  __half* st_state; float* st_action; float* st_reward; int* st_len; int* st_fin;
  // scratch slot (index B): the episode of bcp_replay_add_episode
  __half* sc_state; float* sc_action; float* sc_reward;
  // commit planning
  u64* slot_pref; u64* blk_tot; int* ep_slot; i64* ep_off; int* ep_len; int* blk_push; int* pushcum;
  i64 N, M, C;
  int B, L, W, A;
};

__device__ __forceinline__ const __half* slot_state(const View& v, int s, i64 t) {
  return s < v.B ? v.st_state + ((size_t)s * (v.L + 1) + t) * v.W : v.sc_state + (size_t)t * v.W;
}
__device__ __forceinline__ const float* slot_action(const View& v, int s) {
  return s < v.B ? v.st_action + (size_t)s * v.L * v.A : v.sc_action;
}
__device__ __forceinline__ const float* slot_reward(const View& v, int s) {
  return s < v.B ? v.st_reward + (size_t)s * v.L : v.sc_reward;
}

// exclusive block scan (NT threads, NT / 32 <= 32 warps)
template <typename T, int NT> __device__ __forceinline__ T block_excl_scan(T x, T& total, T* sh /* [33] */) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  T inc = x;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const T y = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += y;
  }
  __syncthreads();  // sh may still be read from a previous call
  if (lane == 31) sh[w] = inc;
  __syncthreads();
  if (w == 0) {
    T s = lane < NT / 32 ? sh[lane] : T(0);
    T si = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const T y = __shfl_up_sync(0xffffffffu, si, d);
      if (lane >= d) si += y;
    }
    if (lane < NT / 32) sh[lane] = si - s;
    if (lane == 31) sh[32] = si;
  }
  __syncthreads();
  total = sh[32];
  return sh[w] + inc - x;
}
template <typename T, int NT> __device__ __forceinline__ T block_sum(T x, T* sh /* [33] */) {
  T tot;
  block_excl_scan<T, NT>(x, tot, sh);
  return tot;
}

// ---- begin / observe ----
// What the caller hands over after env.reset() (mode 1) or env.step() (mode 0); mode -1: nothing
// (bcp_replay_add_episode feeds the scratch slot instead).
struct Feed {
  const void* actions; const float* obs; const float* reward; const uint8_t* done; const uint8_t* reason; const uint8_t* mask;
  int act_i32, mode;
};
// 1: obs[e] opens an episode, 2: obs[e] is the state_2 of a transition, 0: nothing to record
__device__ __forceinline__ int feed_kind(const View& v, const Feed& f, int e, int len, int fin) {
  if (f.mode == 1) return (!f.mask || f.mask[e]) ? 1 : 0;
  const int code = f.reason ? f.reason[e] : 0;
  if (code == 5) return 1;                           // BCP_REASON_FRESH: a reset observation
  if (code == 4 || len < 0 || fin) return 0;         // BCP_REASON_RESETTING / no open episode (step after done)
  return len < v.L ? 2 : 3;                          // 3: the staging row is full (reported by the bookkeeping kernel)
}
// observations -> float16 staging rows, one thread per 4-byte granule of the output (two halves;
// one when the state has an odd number of words): fully parallel, coalesced on both sides
template <bool EVEN>
__global__ void __launch_bounds__(256) k_rp_states(View v, Feed f) {
  const int Wc = EVEN ? v.W / 2 : v.W;
  const size_t total = (size_t)v.B * Wc;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int e = (int)(i / Wc), w = (int)(i - (size_t)e * Wc);
    const int len = v.st_len[e];
    const int kind = feed_kind(v, f, e, len, v.st_fin[e]);
    if (kind != 1 && kind != 2) continue;
    const size_t row = ((size_t)e * (v.L + 1) + (kind == 1 ? 0 : len + 1)) * v.W;
    if (EVEN) {
      const float2 x = reinterpret_cast<const float2*>(f.obs)[i];
      reinterpret_cast<__half2*>(v.st_state + row)[w] = __floats2half2_rn(x.x, x.y);
    } else {
      v.st_state[row + w] = __float2half_rn(f.obs[i]);
    }
  }
}
// per-env bookkeeping of the same call (runs after k_rp_states, inside the slot scan below)
__device__ __forceinline__ void feed_book(const View& v, const Feed& f, int e) {
  const int len = v.st_len[e];
  const int kind = feed_kind(v, f, e, len, v.st_fin[e]);
  if (kind == 1) {
    v.st_len[e] = 0;
    v.st_fin[e] = 0;
  } else if (kind == 2) {
    float* a = v.st_action + ((size_t)e * v.L + len) * v.A;
    if (f.act_i32) a[0] = (float)reinterpret_cast<const int32_t*>(f.actions)[e];
    else for (int i = 0; i < v.A; ++i) a[i] = reinterpret_cast<const float*>(f.actions)[(size_t)e * v.A + i];
    v.st_reward[(size_t)e * v.L + len] = f.reward[e];
    v.st_len[e] = len + 1;
    if (f.done[e]) v.st_fin[e] = 1;
  } else if (kind == 3) {
    atomicOr((u64*)&v.ctl[C_ERROR], (u64)BCP_REPLAY_ERR_STAGING);
  }
}

// float32 episode of bcp_replay_add_episode -> the scratch slot (float16 states)
__global__ void k_rp_scratch(View v, const float* __restrict__ states /* [(n + 1)][W] */, int n) {
  const size_t total = (size_t)(n + 1) * v.W;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
    v.sc_state[i] = __float2half_rn(states[i]);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    v.st_len[v.B] = n;
    v.st_fin[v.B] = 1;
  }
}

// ---- commit planning: finished episodes in slot order, their transition offsets ----
constexpr int kScanT = 1024;
constexpr u64 kLenMask = (1ull << 40) - 1;
__global__ void __launch_bounds__(kScanT) k_rp_scan1(View v, Feed f) {
  __shared__ u64 sh[33];
  const int s = blockIdx.x * kScanT + threadIdx.x;
  if (f.mode >= 0 && s < v.B) feed_book(v, f, s);
  const u64 x = (s <= v.B && v.st_fin[s]) ? ((1ull << 40) | (u64)v.st_len[s]) : 0ull;  // (count, transitions)
  u64 tot;
  const u64 ex = block_excl_scan<u64, kScanT>(x, tot, sh);
  if (s <= v.B) v.slot_pref[s] = ex;
  if (threadIdx.x == 0) v.blk_tot[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(kScanT) k_rp_scan2(View v) {
  __shared__ u64 sh[33];
  u64 part = 0;
  for (int j = threadIdx.x; j < (int)blockIdx.x; j += kScanT) part += v.blk_tot[j];
  const u64 base = block_sum<u64, kScanT>(part, sh);
  const int s = blockIdx.x * kScanT + threadIdx.x;
  if (s <= v.B && v.st_fin[s]) {
    const u64 p = base + v.slot_pref[s];
    const i64 k = (i64)(p >> 40);
    v.ep_slot[k] = s;
    v.ep_off[k] = (i64)(p & kLenMask);
    v.ep_len[k] = v.st_len[s];
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
    const u64 p = base + v.blk_tot[blockIdx.x];
    v.ctl[C_K] = (i64)(p >> 40);
    v.ctl[C_T] = (i64)(p & kLenMask);
    v.ctl[C_PB0] = 0;
  }
}

// ---- a pass over transitions [r0, r0 + Tr): evictions (pushes), then rows + states (pops) ----
constexpr int kPushT = 256;
struct PassRange { i64 r0, tr, lo, hi; };
__device__ __forceinline__ PassRange pass_range(const View& v, int round) {
  PassRange p;
  const i64 T = v.ctl[C_T];
  p.r0 = (i64)round * v.N;
  p.tr = T - p.r0 < v.N ? T - p.r0 : v.N;
  if (p.tr < 0) p.tr = 0;
  const i64 chunk = (p.tr + gridDim.x - 1) / gridDim.x;
  p.lo = chunk * blockIdx.x < p.tr ? chunk * blockIdx.x : p.tr;
  p.hi = p.lo + chunk < p.tr ? p.lo + chunk : p.tr;
  return p;
}
// slots the row at transition g frees before it is overwritten (replay_memory.py:83-93)
__device__ __forceinline__ int evictions(const View& v, i64 insert0, bool full0, i64 g, i64& pos) {
  const i64 lin = insert0 + g;
  pos = lin % v.N;
  if (!(full0 || lin >= v.N)) return 0;
  return v.mask[pos] == 0.0f ? 2 : 1;
}
__global__ void __launch_bounds__(kPushT) k_rp_push1(View v, int round) {
  __shared__ int sh[33];
  const PassRange p = pass_range(v, round);
  const i64 insert0 = v.ctl[C_INSERT];
  const bool full0 = v.ctl[C_FULL] != 0;
  int c = 0;
  for (i64 i = p.lo + threadIdx.x; i < p.hi; i += kPushT) {
    i64 pos;
    c += evictions(v, insert0, full0, p.r0 + i, pos);
  }
  const int tot = block_sum<int, kPushT>(c, sh);
  if (threadIdx.x == 0) v.blk_push[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(kPushT) k_rp_push2(View v, int round) {
  __shared__ i64 shl[33];
  __shared__ int sh[33];
  const PassRange p = pass_range(v, round);
  if (p.tr == 0 && !(blockIdx.x == gridDim.x - 1)) return;
  const i64 insert0 = v.ctl[C_INSERT], tail = v.ctl[C_TAIL];
  const bool full0 = v.ctl[C_FULL] != 0;
  i64 part = 0;
  for (int j = threadIdx.x; j < (int)blockIdx.x; j += kPushT) part += v.blk_push[j];
  i64 run = v.ctl[C_PB0 + (round & 1)] + block_sum<i64, kPushT>(part, shl);  // pushes before this block's range
  int n_s2 = 0;
  for (i64 i0 = p.lo; i0 < p.hi; i0 += kPushT) {
    const i64 i = i0 + threadIdx.x;
    i64 pos = 0;
    const int c = i < p.hi ? evictions(v, insert0, full0, p.r0 + i, pos) : 0;
    int tot;
    const int ex = block_excl_scan<int, kPushT>(c, tot, sh);
    if (i < p.hi) {
      const i64 at = tail + run + ex;
      if (c >= 1) v.fq[at % v.C] = v.s1_idx[pos];
      if (c == 2) { v.fq[(at + 1) % v.C] = v.s2_idx[pos]; ++n_s2; }
      v.pushcum[i] = (int)(run + ex + c);  // pushes up to and including g, since the commit began
    }
    run += tot;
  }
  const int s2_tot = block_sum<int, kPushT>(n_s2, sh);
  if (threadIdx.x == 0) {
    if (s2_tot) atomicAdd((u64*)&v.ctl[C_N_EVICTED_S2], (u64)s2_tot);  // stats['cache_evicted_s2'] :92
    if (blockIdx.x == gridDim.x - 1) v.ctl[C_PB0 + ((round + 1) & 1)] = run;
  }
}
// one warp per 32 consecutive transitions of the pass: rows of the ring, states into the pool.
// A lane finds its episode by bisection over the episodes' transition offsets.
template <bool EVEN>
__global__ void __launch_bounds__(256) k_rp_pop(View v, int round) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const i64 K = v.ctl[C_K], T = v.ctl[C_T];
  const i64 r0 = (i64)round * v.N;
  if (r0 >= T) return;
  const i64 r1 = T - r0 < v.N ? T : r0 + v.N;
  const i64 insert0 = v.ctl[C_INSERT], head = v.ctl[C_HEAD], tail = v.ctl[C_TAIL], pb = v.ctl[C_PB0 + (round & 1)];
  const int Wc = EVEN ? v.W / 2 : v.W;  // copy granule: 4 bytes when the state has an even number of halves
  typedef typename std::conditional<EVEN, uint32_t, unsigned short>::type G;
  G* const pool = reinterpret_cast<G*>(v.state);
  const i64 warps = (i64)gridDim.x * (blockDim.x >> 5);
  for (i64 g0 = r0 + 32 * ((i64)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)); g0 < r1; g0 += 32 * warps) {
    const i64 g = g0 + lane;
    const bool act = g < r1;
    int s1 = 0, s2 = 0, n = 1;
    const G* src = nullptr;   // the transition's state_2 row in the staging area
    const G* src0 = nullptr;  // the episode's initial state (first transition only)
    if (act) {
      i64 lo = 0, hi = K;  // last episode with ep_off <= g
      while (hi - lo > 1) {
        const i64 mid = (lo + hi) >> 1;
        if (v.ep_off[mid] <= g) lo = mid; else hi = mid;
      }
      const i64 k = lo;
      const int slot = v.ep_slot[k], len = v.ep_len[k];
      n = (int)(g - v.ep_off[k]);
      const i64 a2 = head + k + 1 + g;  // pops before this one: k + 1 initial states, g state_2s
      const i64 i = g - r0;
      const i64 pushed = v.pushcum[i];
      const i64 pushed_before = n == 0 ? (i > 0 ? (i64)v.pushcum[i - 1] : pb) : pushed;  // the initial-state pop precedes g's pushes
      if (a2 >= tail + pushed || a2 - 1 >= tail + pushed_before) atomicOr((u64*)&v.ctl[C_ERROR], (u64)BCP_REPLAY_ERR_UNDERFLOW);
      s2 = v.fq[a2 % v.C];
      s1 = v.fq[(a2 - 1) % v.C];
      const i64 pos = (insert0 + g) % v.N;
      v.s1_idx[pos] = s1;  // :96-107
      v.s2_idx[pos] = s2;
      const float* sa = slot_action(v, slot) + (size_t)n * v.A;
      for (int a = 0; a < v.A; ++a) v.action[pos * v.A + a] = sa[a];
      v.reward[pos] = slot_reward(v, slot)[n];
      v.mask[pos] = (n == len - 1) ? 0.0f : 1.0f;  // :102
      src = reinterpret_cast<const G*>(slot_state(v, slot, (i64)n + 1));
      if (n == 0) src0 = reinterpret_cast<const G*>(slot_state(v, slot, 0));
    }
    // states: granule idx of the warp's 32 rows is copied by lane idx % 32
    const u64 srcb = (u64)src;
#pragma unroll 2
    for (int base = 0; base < 32 * Wc; base += 32) {
      const int idx = base + lane;
      const int j = idx / Wc, w = idx - j * Wc;
      const int s2j = __shfl_sync(FULL, s2, j & 31);
      const G* sj = (const G*)__shfl_sync(FULL, srcb, j & 31);
      if (sj) pool[(size_t)s2j * Wc + w] = sj[w];
    }
    unsigned firsts = __ballot_sync(FULL, src0 != nullptr);  // episodes that start in this chunk (:66-67)
    while (firsts) {
      const int j = __ffs(firsts) - 1;
      firsts &= firsts - 1;
      const int s1j = __shfl_sync(FULL, s1, j);
      const G* s0 = (const G*)__shfl_sync(FULL, (u64)src0, j);
      for (int w = lane; w < Wc; w += 32) pool[(size_t)s1j * Wc + w] = s0[w];
    }
  }
}
__global__ void __launch_bounds__(256) k_rp_finalize(View v, int rounds_launched) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s <= v.B && v.st_fin[s]) {
    v.st_fin[s] = 0;
    v.st_len[s] = -1;
  }
  if (s == 0) {
    const i64 K = v.ctl[C_K], T = v.ctl[C_T];
    const i64 rounds = (T + v.N - 1) / v.N;
    if (rounds > rounds_launched) atomicOr((u64*)&v.ctl[C_ERROR], (u64)BCP_REPLAY_ERR_ROUNDS);
    const i64 lin = v.ctl[C_INSERT] + T;
    if (lin >= v.N) v.ctl[C_FULL] = 1;  // :110-113
    v.ctl[C_INSERT] = lin % v.N;
    v.ctl[C_HEAD] += K + T;
    v.ctl[C_TAIL] += v.ctl[C_PB0 + (rounds & 1)];
    v.ctl[C_N_ADD_EPISODE] += K;  // stats['>add_episode'] :65
    v.ctl[C_N_ADD] += T;          // stats['>add'] :77
  }
}

// ---- batch gather (replay_memory.py:131-138): a half-warp per row when the state fits 16 granules ----
template <bool EVEN>
__global__ void __launch_bounds__(256) k_rp_batch(View v, const int64_t* __restrict__ idxs, int n, u64 seed, u64 counter,
                                                  __half* __restrict__ o_s1, float* __restrict__ o_action, float* __restrict__ o_reward,
                                                  float* __restrict__ o_mask, __half* __restrict__ o_s2, int64_t* __restrict__ o_idx) {
  const i64 size = v.ctl[C_FULL] ? v.N : v.ctl[C_INSERT];  // :120-121
  if (size == 0) return;
  const int Wc = EVEN ? v.W / 2 : v.W;
  const int gs = Wc <= 16 ? 16 : 32;  // lanes per row
  const int sub = threadIdx.x & (gs - 1);
  typedef typename std::conditional<EVEN, uint32_t, unsigned short>::type G;
  const G* pool = reinterpret_cast<const G*>(v.state);
  G* d1 = reinterpret_cast<G*>(o_s1);
  G* d2 = reinterpret_cast<G*>(o_s2);
  const int groups = gridDim.x * (blockDim.x / gs);
  for (int i = blockIdx.x * (blockDim.x / gs) + threadIdx.x / gs; i < n; i += groups) {
    i64 idx;
    if (idxs) {
      idx = idxs[i];
      if (idx < 0 || idx >= size) {
        if (sub == 0) atomicOr((u64*)&v.ctl[C_ERROR], (u64)BCP_REPLAY_ERR_INDEX);
        idx = 0;
      }
    } else {  // np.random.randint(0, size) :123-129, here a counter-based draw
      bcp::U4 c;
      c.v[0] = (uint32_t)i; c.v[1] = (uint32_t)counter; c.v[2] = (uint32_t)(counter >> 32); c.v[3] = 0x5eb1a7u;
      const bcp::U4 r = bcp::philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
      const u64 r64 = ((u64)r.v[0] << 32) | r.v[1];
      idx = (i64)__umul64hi(r64, (u64)size);
    }
    const int s1 = v.s1_idx[idx], s2 = v.s2_idx[idx];
    for (int w = sub; w < Wc; w += gs) {
      d1[(size_t)i * Wc + w] = pool[(size_t)s1 * Wc + w];
      d2[(size_t)i * Wc + w] = pool[(size_t)s2 * Wc + w];
    }
    for (int a = sub; a < v.A; a += gs) o_action[(size_t)i * v.A + a] = v.action[idx * v.A + a];
    if (sub == 0) {
      o_reward[i] = v.reward[idx];
      o_mask[i] = v.mask[idx];
      if (o_idx) o_idx[i] = idx;
    }
  }
}

__global__ void k_rp_init(View v) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x; i < v.C; i += stride) v.fq[i] = (int)(i % v.M);  // range(M) :36
  for (i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x; i <= v.B; i += stride) { v.st_len[i] = -1; v.st_fin[i] = 0; }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    for (int i = 0; i < C_WORDS; ++i) v.ctl[i] = 0;
    v.ctl[C_TAIL] = v.M;
  }
}

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

}  // namespace

struct bcp_replay {
  bcp_replay_config cfg;
  View v;
  void* arena = nullptr;  // one allocation behind every pointer of the view
  float* in_states = nullptr;  // add_episode: float32 episode as given, [(cap + 1)][W]
  int scratch_cap = 0;
  int sms = 0;
  int64_t launches = 0;
  // bcp_replay_observe as one CUDA graph launch: the call's kernels with their arguments, captured
  // the first time a set of caller buffers is seen (an env loop passes the same buffers every call)
  cudaGraphExec_t graph = nullptr;
  cudaStream_t cap_stream = nullptr;
  const void* graph_key[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  int graph_act_i32 = 0, graph_nodes = 0, use_graph = 1;
};

namespace {

int grow_scratch(bcp_replay* rp, int n) {
  if (n <= rp->scratch_cap) return 0;
  int cap = rp->scratch_cap ? rp->scratch_cap : 256;
  while (cap < n) cap *= 2;
  RCK(cudaDeviceSynchronize());  // earlier commits may still read the old scratch
  if (rp->graph) { cudaGraphExecDestroy(rp->graph); rp->graph = nullptr; }  // the graph's kernels carry the old pointers
  cudaFree(rp->v.sc_state); cudaFree(rp->v.sc_action); cudaFree(rp->v.sc_reward); cudaFree(rp->in_states);
  rp->v.sc_state = nullptr; rp->v.sc_action = nullptr; rp->v.sc_reward = nullptr; rp->in_states = nullptr;
  rp->scratch_cap = 0;
  const size_t W = rp->cfg.state_words, A = rp->cfg.action_dim;
  RCK(cudaMalloc((void**)&rp->v.sc_state, (size_t)(cap + 1) * W * sizeof(__half)));
  RCK(cudaMalloc((void**)&rp->v.sc_action, (size_t)cap * A * sizeof(float)));
  RCK(cudaMalloc((void**)&rp->v.sc_reward, (size_t)cap * sizeof(float)));
  RCK(cudaMalloc((void**)&rp->in_states, (size_t)(cap + 1) * W * sizeof(float)));
  rp->scratch_cap = cap;
  return 0;
}

// add_episode of every finished slot, in slot order.  max_T bounds the transitions the commit can hold.
int launch_states(bcp_replay* rp, const Feed& f, cudaStream_t s) {
  const View& v = rp->v;
  const bool even = v.W % 2 == 0 && ((uintptr_t)f.obs & 7) == 0;  // float2 loads need an 8-byte aligned source
  const size_t total = (size_t)v.B * (even ? v.W / 2 : v.W);
  int grid = (int)((total + 255) / 256);
  if (grid > rp->sms * 16) grid = rp->sms * 16;
  if (even) k_rp_states<true><<<grid, 256, 0, s>>>(v, f);
  else k_rp_states<false><<<grid, 256, 0, s>>>(v, f);
  rp->launches += 1;
  RCK(cudaGetLastError());
  return 0;
}

int commit(bcp_replay* rp, const Feed& f, i64 max_T, cudaStream_t s) {
  const View& v = rp->v;
  const int slots = v.B + 1;
  const int nb1 = (slots + kScanT - 1) / kScanT;
  k_rp_scan1<<<nb1, kScanT, 0, s>>>(v, f);
  k_rp_scan2<<<nb1, kScanT, 0, s>>>(v);
  int rounds = (int)((max_T + v.N - 1) / v.N);
  if (rounds < 1) rounds = 1;
  const i64 per_round = max_T < v.N ? max_T : v.N;
  int g_push = (int)((per_round + 4 * kPushT - 1) / (4 * kPushT));  // >= 4 tiles per block, at most 4 blocks per SM
  if (g_push > rp->sms * 4) g_push = rp->sms * 4;
  if (g_push < 1) g_push = 1;
  int g_pop = (int)((per_round + 255) / 256);  // a warp per 32 transitions
  if (g_pop > rp->sms * 8) g_pop = rp->sms * 8;
  if (g_pop < 1) g_pop = 1;
  for (int r = 0; r < rounds; ++r) {
    k_rp_push1<<<g_push, kPushT, 0, s>>>(v, r);
    k_rp_push2<<<g_push, kPushT, 0, s>>>(v, r);
    if (v.W % 2 == 0) k_rp_pop<true><<<g_pop, 256, 0, s>>>(v, r);
    else k_rp_pop<false><<<g_pop, 256, 0, s>>>(v, r);
  }
  k_rp_finalize<<<(slots + 255) / 256, 256, 0, s>>>(v, rounds);
  rp->launches += 3 + 3 * rounds;
  RCK(cudaGetLastError());
  return 0;
}

}  // namespace

extern "C" {

size_t bcp_replay_config_size(void) { return sizeof(bcp_replay_config); }

int bcp_replay_create(const bcp_replay_config* c, bcp_replay** out) {
  if (!out) return abi_fail("bcp_replay_create: null out", cudaSuccess);
  *out = nullptr;
  if (!c) return abi_fail("bcp_replay_create: null config", cudaSuccess);
  if (c->abi_version != BCP_REPLAY_ABI_VERSION) return abi_fail("bcp_replay_create: abi_version mismatch", cudaSuccess);
  if (c->buffer_size < 1 || c->buffer_size > (1ll << 31) - 1) return abi_fail("bcp_replay_create: buffer_size out of range", cudaSuccess);
  if (c->state_buffer_size < c->buffer_size + c->buffer_size / 2 || c->state_buffer_size > (1ll << 31) - 1)
    return abi_fail("bcp_replay_create: load_factor has to be at least 1.5", cudaSuccess);  // replay_memory.py:13
  if (c->state_words < 1 || c->action_dim < 1) return abi_fail("bcp_replay_create: state_words and action_dim must be >= 1", cudaSuccess);
  if (c->num_envs < 0 || (c->num_envs > 0 && c->max_episode_len < 1)) return abi_fail("bcp_replay_create: bad num_envs / max_episode_len", cudaSuccess);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) return abi_fail("bcp_replay_create: no CUDA device (this library has no CPU path)", e);
  if (c->device_ordinal < 0 || c->device_ordinal >= ndev) return abi_fail("bcp_replay_create: device_ordinal out of range", cudaSuccess);
  bcp_replay* rp = new (std::nothrow) bcp_replay();
  if (!rp) return abi_fail("bcp_replay_create: out of host memory", cudaSuccess);
  rp->cfg = *c;
  DeviceGuard g(c->device_ordinal);
  cudaDeviceGetAttribute(&rp->sms, cudaDevAttrMultiProcessorCount, c->device_ordinal);
  if (const char* e2 = getenv("BCP_REPLAY_GRAPH")) rp->use_graph = atoi(e2) != 0;  // 0: plain launches (per-kernel profiling)
  View& v = rp->v;
  v = View{};
  v.N = c->buffer_size; v.M = c->state_buffer_size; v.C = v.M + 2 * v.N + 8;
  v.B = c->num_envs; v.L = c->num_envs ? c->max_episode_len : 0; v.W = c->state_words; v.A = c->action_dim;
  // one arena, 256-byte aligned pieces
  size_t at = 0;
  auto piece = [&](size_t bytes) { const size_t o = at; at += (bytes + 255) & ~(size_t)255; return o; };
  const size_t N = (size_t)v.N, M = (size_t)v.M, B = (size_t)v.B, L = (size_t)v.L, W = (size_t)v.W, A = (size_t)v.A;
  const size_t nb1 = (B + 1 + kScanT - 1) / kScanT;
  const size_t o_s1 = piece(N * 4), o_s2 = piece(N * 4), o_act = piece(N * A * 4), o_rew = piece(N * 4), o_msk = piece(N * 4);
  const size_t o_state = piece(M * W * 2), o_fq = piece((size_t)v.C * 4), o_ctl = piece(C_WORDS * 8);
  const size_t o_sts = piece(B * (L + 1) * W * 2), o_sta = piece(B * L * A * 4), o_str = piece(B * L * 4);
  const size_t o_len = piece((B + 1) * 4), o_fin = piece((B + 1) * 4);
  const size_t o_pref = piece((B + 1) * 8), o_tot = piece(nb1 * 8), o_eps = piece((B + 1) * 4), o_epo = piece((B + 1) * 8), o_epl = piece((B + 1) * 4);
  const size_t o_bp = piece((size_t)rp->sms * 4 * 4 + 64), o_pc = piece(N * 4);
  e = cudaMalloc(&rp->arena, at);
  if (e != cudaSuccess) { delete rp; return abi_fail("bcp_replay_create: cudaMalloc", e); }
  cudaMemset(rp->arena, 0, at);
  char* b = (char*)rp->arena;
  v.s1_idx = (int*)(b + o_s1); v.s2_idx = (int*)(b + o_s2); v.action = (float*)(b + o_act); v.reward = (float*)(b + o_rew);
  v.mask = (float*)(b + o_msk); v.state = (__half*)(b + o_state); v.fq = (int*)(b + o_fq); v.ctl = (i64*)(b + o_ctl);
  v.st_state = (__half*)(b + o_sts); v.st_action = (float*)(b + o_sta); v.st_reward = (float*)(b + o_str);
  v.st_len = (int*)(b + o_len); v.st_fin = (int*)(b + o_fin);
  v.slot_pref = (u64*)(b + o_pref); v.blk_tot = (u64*)(b + o_tot); v.ep_slot = (int*)(b + o_eps); v.ep_off = (i64*)(b + o_epo);
  v.ep_len = (int*)(b + o_epl); v.blk_push = (int*)(b + o_bp); v.pushcum = (int*)(b + o_pc);
  k_rp_init<<<rp->sms * 4, 256>>>(v);
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { cudaFree(rp->arena); delete rp; return abi_fail("bcp_replay_create: init", e); }
  rp->launches = 1;
  *out = rp;
  return 0;
}

int bcp_replay_destroy(bcp_replay* rp) {
  if (!rp) return 0;
  DeviceGuard g(rp->cfg.device_ordinal);
  cudaDeviceSynchronize();
  if (rp->graph) cudaGraphExecDestroy(rp->graph);
  if (rp->cap_stream) cudaStreamDestroy(rp->cap_stream);
  cudaFree(rp->arena);
  cudaFree(rp->v.sc_state); cudaFree(rp->v.sc_action); cudaFree(rp->v.sc_reward); cudaFree(rp->in_states);
  delete rp;
  return 0;
}

int bcp_replay_begin(bcp_replay* rp, const float* obs, const uint8_t* mask, void* stream) {
  if (!rp) return abi_fail("bcp_replay_begin: null handle", cudaSuccess);
  if (!obs) return abi_fail("bcp_replay_begin: null obs", cudaSuccess);
  if (rp->v.B < 1) return abi_fail("bcp_replay_begin: the handle was created with num_envs = 0", cudaSuccess);
  DeviceGuard g(rp->cfg.device_ordinal);
  const Feed f{nullptr, obs, nullptr, nullptr, nullptr, mask, 0, 1};
  if (launch_states(rp, f, (cudaStream_t)stream)) return -1;
  k_rp_scan1<<<(rp->v.B + kScanT) / kScanT, kScanT, 0, (cudaStream_t)stream>>>(rp->v, f);  // the per-env bookkeeping alone
  rp->launches += 1;
  RCK(cudaGetLastError());
  return 0;
}

int bcp_replay_observe(bcp_replay* rp, const void* actions, int32_t actions_are_int32, const float* obs, const float* reward,
                       const uint8_t* done, const uint8_t* reason, void* stream) {
  if (!rp) return abi_fail("bcp_replay_observe: null handle", cudaSuccess);
  if (!actions || !obs || !reward || !done) return abi_fail("bcp_replay_observe: null argument", cudaSuccess);
  if (rp->v.B < 1) return abi_fail("bcp_replay_observe: the handle was created with num_envs = 0", cudaSuccess);
  if (actions_are_int32 && rp->v.A != 1) return abi_fail("bcp_replay_observe: int32 actions need action_dim = 1", cudaSuccess);
  DeviceGuard g(rp->cfg.device_ordinal);
  const Feed f{actions, obs, reward, done, reason, nullptr, actions_are_int32, 0};
  const i64 max_T = (i64)rp->v.B * rp->v.L;
  if (!rp->use_graph) {
    if (launch_states(rp, f, (cudaStream_t)stream)) return -1;
    return commit(rp, f, max_T, (cudaStream_t)stream);
  }
  const void* key[6] = {actions, obs, reward, done, reason, nullptr};
  bool same = rp->graph != nullptr && rp->graph_act_i32 == actions_are_int32;
  for (int i = 0; i < 6 && same; ++i) same = key[i] == rp->graph_key[i];
  if (!same) {
    if (rp->graph) { cudaGraphExecDestroy(rp->graph); rp->graph = nullptr; }
    if (!rp->cap_stream) RCK(cudaStreamCreateWithFlags(&rp->cap_stream, cudaStreamNonBlocking));
    const int64_t before = rp->launches;
    RCK(cudaStreamBeginCapture(rp->cap_stream, cudaStreamCaptureModeThreadLocal));
    int rc = launch_states(rp, f, rp->cap_stream);
    if (!rc) rc = commit(rp, f, max_T, rp->cap_stream);
    cudaGraph_t g = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(rp->cap_stream, &g);
    rp->graph_nodes = (int)(rp->launches - before);
    rp->launches = before;
    if (rc) { if (g) cudaGraphDestroy(g); return -1; }
    if (ce != cudaSuccess) return abi_fail("bcp_replay_observe: stream capture", ce);
    const cudaError_t ie = cudaGraphInstantiate(&rp->graph, g, 0);
    cudaGraphDestroy(g);
    if (ie != cudaSuccess) { rp->graph = nullptr; return abi_fail("bcp_replay_observe: cudaGraphInstantiate", ie); }
    for (int i = 0; i < 6; ++i) rp->graph_key[i] = key[i];
    rp->graph_act_i32 = actions_are_int32;
  }
  RCK(cudaGraphLaunch(rp->graph, (cudaStream_t)stream));
  rp->launches += rp->graph_nodes;
  return 0;
}

int bcp_replay_add_episode(bcp_replay* rp, const float* initial_state, const float* actions, const float* rewards,
                           const float* states2, int32_t n, void* stream) {
  if (!rp) return abi_fail("bcp_replay_add_episode: null handle", cudaSuccess);
  if (!initial_state || !actions || !rewards || !states2) return abi_fail("bcp_replay_add_episode: null argument", cudaSuccess);
  if (n < 1) return abi_fail("bcp_replay_add_episode: an episode needs at least one transition", cudaSuccess);  // :65
  DeviceGuard g(rp->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  if (grow_scratch(rp, n)) return -1;
  const size_t W = rp->v.W, A = rp->v.A;
  RCK(cudaMemcpyAsync(rp->in_states, initial_state, W * sizeof(float), cudaMemcpyDefault, s));
  RCK(cudaMemcpyAsync(rp->in_states + W, states2, (size_t)n * W * sizeof(float), cudaMemcpyDefault, s));
  RCK(cudaMemcpyAsync(rp->v.sc_action, actions, (size_t)n * A * sizeof(float), cudaMemcpyDefault, s));
  RCK(cudaMemcpyAsync(rp->v.sc_reward, rewards, (size_t)n * sizeof(float), cudaMemcpyDefault, s));
  const size_t total = (size_t)(n + 1) * W;
  int grid = (int)((total + 255) / 256);
  if (grid > rp->sms * 8) grid = rp->sms * 8;
  k_rp_scratch<<<grid, 256, 0, s>>>(rp->v, rp->in_states, n);
  rp->launches += 1;
  RCK(cudaGetLastError());
  const Feed none{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, -1};
  return commit(rp, none, (i64)n, s);
}

int bcp_replay_batch(bcp_replay* rp, const int64_t* idxs, int32_t n, uint64_t seed, uint64_t counter, void* out_state_1,
                     float* out_action, float* out_reward, float* out_mask, void* out_state_2, int64_t* out_idx, void* stream) {
  if (!rp) return abi_fail("bcp_replay_batch: null handle", cudaSuccess);
  if (n < 0) return abi_fail("bcp_replay_batch: negative batch size", cudaSuccess);
  if (n == 0) return 0;
  if (!out_state_1 || !out_action || !out_reward || !out_mask || !out_state_2) return abi_fail("bcp_replay_batch: null output", cudaSuccess);
  DeviceGuard g(rp->cfg.device_ordinal);
  int grid = (n + 15) / 16;
  if (grid > rp->sms * 8) grid = rp->sms * 8;
  if (rp->v.W % 2 == 0)
    k_rp_batch<true><<<grid, 256, 0, (cudaStream_t)stream>>>(rp->v, idxs, n, seed, counter, (__half*)out_state_1, out_action, out_reward, out_mask, (__half*)out_state_2, out_idx);
  else
    k_rp_batch<false><<<grid, 256, 0, (cudaStream_t)stream>>>(rp->v, idxs, n, seed, counter, (__half*)out_state_1, out_action, out_reward, out_mask, (__half*)out_state_2, out_idx);
  rp->launches += 1;
  RCK(cudaGetLastError());
  return 0;
}

int bcp_replay_info(bcp_replay* rp, int64_t* host_info, void* stream) {
  if (!rp || !host_info) return abi_fail("bcp_replay_info: null argument", cudaSuccess);
  DeviceGuard g(rp->cfg.device_ordinal);
  i64 c[C_WORDS];
  RCK(cudaMemcpyAsync(c, rp->v.ctl, sizeof(c), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  RCK(cudaStreamSynchronize((cudaStream_t)stream));
  host_info[0] = c[C_INSERT];
  host_info[1] = c[C_FULL];
  host_info[2] = c[C_FULL] ? rp->v.N : c[C_INSERT];
  host_info[3] = c[C_TAIL] - c[C_HEAD];
  host_info[4] = c[C_N_ADD_EPISODE];
  host_info[5] = c[C_N_ADD];
  host_info[6] = c[C_N_EVICTED_S2];
  host_info[7] = c[C_ERROR];
  host_info[8] = c[C_K];
  host_info[9] = c[C_T];
  return 0;
}

int bcp_replay_export(bcp_replay* rp, int32_t* state_1_idx, float* action, float* reward, float* terminal_mask,
                      int32_t* state_2_idx, void* state_f16, int32_t* free_slots, void* stream) {
  if (!rp) return abi_fail("bcp_replay_export: null handle", cudaSuccess);
  DeviceGuard g(rp->cfg.device_ordinal);
  cudaStream_t s = (cudaStream_t)stream;
  const View& v = rp->v;
  const size_t N = (size_t)v.N;
  RCK(cudaStreamSynchronize(s));
  if (state_1_idx) RCK(cudaMemcpy(state_1_idx, v.s1_idx, N * 4, cudaMemcpyDeviceToHost));
  if (state_2_idx) RCK(cudaMemcpy(state_2_idx, v.s2_idx, N * 4, cudaMemcpyDeviceToHost));
  if (action) RCK(cudaMemcpy(action, v.action, N * v.A * 4, cudaMemcpyDeviceToHost));
  if (reward) RCK(cudaMemcpy(reward, v.reward, N * 4, cudaMemcpyDeviceToHost));
  if (terminal_mask) RCK(cudaMemcpy(terminal_mask, v.mask, N * 4, cudaMemcpyDeviceToHost));
  if (state_f16) RCK(cudaMemcpy(state_f16, v.state, (size_t)v.M * v.W * 2, cudaMemcpyDeviceToHost));
  if (free_slots) {
    i64 c[C_WORDS];
    RCK(cudaMemcpy(c, v.ctl, sizeof(c), cudaMemcpyDeviceToHost));
    i64 n = c[C_TAIL] - c[C_HEAD];
    if (n < 0) n = 0;
    if (n > v.M) n = v.M;
    for (i64 done = 0; done < n;) {  // the list in order: at most two contiguous pieces of the circular array
      const i64 at = (c[C_HEAD] + done) % v.C;
      const i64 len = (v.C - at < n - done) ? v.C - at : n - done;
      RCK(cudaMemcpy(free_slots + done, v.fq + at, (size_t)len * 4, cudaMemcpyDeviceToHost));
      done += len;
    }
  }
  return 0;
}

int64_t bcp_replay_launch_count(const bcp_replay* rp) { return rp ? rp->launches : 0; }

}  // extern "C"
