// bcp_math.cuh — scalar helpers of the batched BulletCartpole kernels.
//
// Everything here is templated on the physics scalar T (float = fast mode,
// double = parity mode) and is __host__ __device__ so the very same code can be
// exercised by the host-side kernel harness in tests/ (no GPU in the build box).
//
// The elementary functions are *defined* here as fixed sequences of IEEE
// operations (no libdevice sin/cos/atan2): with -fmad=false the FP64 kernel is
// then bit-reproducible against the CPU oracle, which restates the same
// definitions independently in C (oracle/bcp_oracle.c).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define BCP_HD __host__ __device__ __forceinline__
#define BCP_HD_NOINLINE __host__ __device__ __noinline__
#else
#define BCP_HD inline
#define BCP_HD_NOINLINE
#include <cmath>
#endif

namespace bcp {

template <typename T> struct Lim;
template <> struct Lim<float> {
  static BCP_HD float eps() { return 1.1920928955078125e-7f; }  // SIMD_EPSILON (FLT_EPSILON)
  static BCP_HD float large() { return 1e18f; }                 // BT_LARGE_FLOAT
};
template <> struct Lim<double> {
  static BCP_HD double eps() { return 2.220446049250313e-16; }
  static BCP_HD double large() { return 1e18; }
};

// Flop census hook: tests/host_harness/flopcount.cpp instantiates the parity substep with a counting scalar and
// specialises this to attribute the operations to a phase of the substep; for float and double it is nothing.
enum { CENSUS_NARROWPHASE = 0, CENSUS_MANIFOLD, CENSUS_DYNAMICS, CENSUS_ROWS, CENSUS_PGS, CENSUS_INTEGRATE, CENSUS_PHASES };
template <typename T> BCP_HD void census_phase(int) {}

BCP_HD float bsqrt(float x) { return sqrtf(x); }
BCP_HD double bsqrt(double x) { return sqrt(x); }
BCP_HD float babs(float x) { return fabsf(x); }
BCP_HD double babs(double x) { return fabs(x); }

#define BCP_PI 3.141592653589793

// sin / cos on |x| <= pi/4 (Taylor, Horner, one rounding per operation)
template <typename T> BCP_HD T sin_core(T x) {
  const T x2 = x * x;
  T p = T(-2.8114572543455206e-15);
  p = p * x2 + T(7.647163731819816e-13);
  p = p * x2 + T(-1.6059043836821613e-10);
  p = p * x2 + T(2.505210838544172e-08);
  p = p * x2 + T(-2.7557319223985893e-06);
  p = p * x2 + T(0.0001984126984126984);
  p = p * x2 + T(-0.008333333333333333);
  p = p * x2 + T(0.16666666666666666);
  return x - (x * x2) * p;
}
template <typename T> BCP_HD T cos_core(T x) {
  const T x2 = x * x;
  T p = T(4.779477332387385e-14);
  p = p * x2 + T(-1.1470745597729725e-11);
  p = p * x2 + T(2.08767569878681e-09);
  p = p * x2 + T(-2.755731922398589e-07);
  p = p * x2 + T(2.48015873015873e-05);
  p = p * x2 + T(-0.001388888888888889);
  p = p * x2 + T(0.041666666666666664);
  p = p * x2 + T(-0.5);
  return T(1.0) + x2 * p;
}
// (sin, cos)(2*pi*u), u = k / 2^24 in [0,1): exact octant reduction
template <typename T> BCP_HD void sincos_2pi(T u, T& s_out, T& c_out) {
  const T t = u * T(8.0);
  const int oct = (int)t;
  const int q = (oct + 1) >> 1;
  const T f = t - T(2.0) * (T)q;
  const T r = f * T(BCP_PI / 4.0);
  const T s = sin_core(r), c = cos_core(r);
  switch (q & 3) {
    case 0: s_out = s; c_out = c; break;
    case 1: s_out = c; c_out = -s; break;
    case 2: s_out = -s; c_out = -c; break;
    default: s_out = -c; c_out = s; break;
  }
}
// atan2 for ordering clipped polygon vertices (cullPoints2 only)
template <typename T> BCP_HD T atan_unit(T t) {
  for (int k = 0; k < 3; ++k) t = t / (T(1.0) + bsqrt(T(1.0) + t * t));
  const T t2 = t * t;
  T p = T(1.0 / 17.0);
  p = T(-1.0 / 15.0) + t2 * p;
  p = T(1.0 / 13.0) + t2 * p;
  p = T(-1.0 / 11.0) + t2 * p;
  p = T(1.0 / 9.0) + t2 * p;
  p = T(-1.0 / 7.0) + t2 * p;
  p = T(1.0 / 5.0) + t2 * p;
  p = T(-1.0 / 3.0) + t2 * p;
  p = T(1.0) + t2 * p;
  return T(8.0) * (t * p);
}
template <typename T> BCP_HD T det_atan2(T y, T x) {
  const T ay = babs(y), ax = babs(x);
  T a;
  if (ax == T(0.0) && ay == T(0.0)) return T(0.0);
  if (ay <= ax) a = atan_unit(ay / ax);
  else a = T(BCP_PI / 2.0) - atan_unit(ax / ay);
  if (x < T(0.0)) a = T(BCP_PI) - a;
  return (y < T(0.0)) ? -a : a;
}

// ---- Philox4x32-10 (counter-based RNG, Salmon et al. SC'11) -----------------
struct U4 { uint32_t v[4]; };
BCP_HD U4 philox4x32_10(U4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c.v[0];
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c.v[2];
    U4 n;
    n.v[0] = (uint32_t)(p1 >> 32) ^ c.v[1] ^ k0;
    n.v[1] = (uint32_t)p1;
    n.v[2] = (uint32_t)(p0 >> 32) ^ c.v[3] ^ k1;
    n.v[3] = (uint32_t)p0;
    c = n;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c;
}
enum : uint32_t { RNG_PUSH = 1u, RNG_ACTION = 2u };
// key = seed; counter = (global env id lo/hi, episode, purpose<<24 | index)
BCP_HD U4 env_rng(uint64_t seed, int64_t global_env, uint32_t episode, uint32_t purpose, uint32_t index) {
  U4 c;
  c.v[0] = (uint32_t)((uint64_t)global_env & 0xffffffffu);
  c.v[1] = (uint32_t)((uint64_t)global_env >> 32);
  c.v[2] = episode;
  c.v[3] = (purpose << 24) | (index & 0xffffffu);
  return philox4x32_10(c, (uint32_t)(seed & 0xffffffffu), (uint32_t)(seed >> 32));
}
template <typename T> BCP_HD T u24(uint32_t x) { return (T)(x >> 8) * T(5.9604644775390625e-08); }

// ---- 3-vectors / 3x3 (row-major body->world basis) ---------------------------
template <typename T> BCP_HD T dot3(const T* a, const T* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename T> BCP_HD void cross3(const T* a, const T* b, T* o) {
  const T x = a[1] * b[2] - a[2] * b[1];
  const T y = a[2] * b[0] - a[0] * b[2];
  const T z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
// btMatrix3x3::setRotation
template <typename T> BCP_HD void quat_to_mat(const T* q, T* m) {
  const T d = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  const T s = T(2.0) / d;
  const T xs = q[0] * s, ys = q[1] * s, zs = q[2] * s;
  const T wx = q[3] * xs, wy = q[3] * ys, wz = q[3] * zs;
  const T xx = q[0] * xs, xy = q[0] * ys, xz = q[0] * zs;
  const T yy = q[1] * ys, yz = q[1] * zs, zz = q[2] * zs;
  m[0] = T(1.0) - (yy + zz); m[1] = xy - wz; m[2] = xz + wy;
  m[3] = xy + wz; m[4] = T(1.0) - (xx + zz); m[5] = yz - wx;
  m[6] = xz - wy; m[7] = yz + wx; m[8] = T(1.0) - (xx + yy);
}
template <typename T> BCP_HD void mat_mul_vec(const T* m, const T* v, T* o) {
  const T x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  const T y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  const T z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
template <typename T> BCP_HD void matT_mul_vec(const T* m, const T* v, T* o) {
  const T x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2];
  const T y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2];
  const T z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
template <typename T> BCP_HD T clampv(T v, T lo, T hi) { return v < lo ? lo : (v > hi ? hi : v); }

}  // namespace bcp
