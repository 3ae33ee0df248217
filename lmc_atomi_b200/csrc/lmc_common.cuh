// Shared device/host helpers of the lmc_b200 library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#include "../../include/lmc_b200.h"

// ------------------------------------------------------------------ error plumbing
void lmc_set_error(const char* fmt, ...);

#define LMC_CUDA_OK(expr)                                                          \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) {                                                       \
      lmc_set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,           \
                    cudaGetErrorString(_e));                                       \
      return -2;                                                                   \
    }                                                                              \
  } while (0)

#define LMC_REQUIRE(cond, ...)                                                     \
  do {                                                                             \
    if (!(cond)) {                                                                 \
      lmc_set_error(__VA_ARGS__);                                                  \
      return -1;                                                                   \
    }                                                                              \
  } while (0)

// ------------------------------------------------------------------ Philox4x32-10
// Counter-based generator (Salmon et al., SC'11).  counter = (c0,c1,c2,c3), key = (k0,k1).
// Also restated in numpy by oracle/philox_ref.py for the production-mode parity tests.
#define LMC_PHILOX_M0 0xD2511F53u
#define LMC_PHILOX_M1 0xCD9E8D57u
#define LMC_PHILOX_W0 0x9E3779B9u
#define LMC_PHILOX_W1 0xBB67AE85u

__host__ __device__ __forceinline__ uint4 lmc_philox4x32_10(uint4 ctr, uint2 key) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
#ifdef __CUDA_ARCH__
    uint32_t hi0 = __umulhi(LMC_PHILOX_M0, ctr.x), lo0 = LMC_PHILOX_M0 * ctr.x;
    uint32_t hi1 = __umulhi(LMC_PHILOX_M1, ctr.z), lo1 = LMC_PHILOX_M1 * ctr.z;
#else
    uint64_t p0 = (uint64_t)LMC_PHILOX_M0 * ctr.x, p1 = (uint64_t)LMC_PHILOX_M1 * ctr.z;
    uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#endif
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += LMC_PHILOX_W0;
    key.y += LMC_PHILOX_W1;
  }
  return ctr;
}

// Same generator with the ten round keys precomputed (host side / once per kernel): per
// round two 32x32->64 multiplies (IMAD.WIDE) and two three-input XORs whose key operand is
// a constant-bank word.
__host__ __device__ __forceinline__ void lmc_philox_round_keys(uint2 key, uint32_t rk[20]) {
  for (int r = 0; r < 10; ++r) {
    rk[2 * r] = key.x + (uint32_t)r * LMC_PHILOX_W0;
    rk[2 * r + 1] = key.y + (uint32_t)r * LMC_PHILOX_W1;
  }
}

__device__ __forceinline__ uint4 lmc_philox4x32_10_rk(uint4 ctr, const uint32_t* __restrict__ rk) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned long long p0 = (unsigned long long)LMC_PHILOX_M0 * ctr.x;
    const unsigned long long p1 = (unsigned long long)LMC_PHILOX_M1 * ctr.z;
    ctr = make_uint4((uint32_t)(p1 >> 32) ^ ctr.y ^ rk[2 * r], (uint32_t)p1,
                     (uint32_t)(p0 >> 32) ^ ctr.w ^ rk[2 * r + 1], (uint32_t)p0);
  }
  return ctr;
}

// ------------------------------------------------------------------ MUFU-backed fp32 primitives
// One SFU instruction each (relative error about 2^-22), no slow paths: the register-resident
// chains are bound by instruction issue, so the libdevice range-reduction sequences matter.
#define LMC_MUFU1(name, op)                                            \
  __device__ __forceinline__ float name(float x) {                     \
    float y;                                                           \
    asm(op " %0, %1;" : "=f"(y) : "f"(x));                             \
    return y;                                                          \
  }
LMC_MUFU1(lmc_ex2f, "ex2.approx.ftz.f32")
LMC_MUFU1(lmc_lg2f, "lg2.approx.ftz.f32")
LMC_MUFU1(lmc_rcpf, "rcp.approx.ftz.f32")
LMC_MUFU1(lmc_sqrtf, "sqrt.approx.ftz.f32")
LMC_MUFU1(lmc_rsqrtf, "rsqrt.approx.ftz.f32")
LMC_MUFU1(lmc_sinf, "sin.approx.ftz.f32")
LMC_MUFU1(lmc_cosf, "cos.approx.ftz.f32")
#undef LMC_MUFU1

// streams of one chain (second counter word)
#define LMC_STREAM_STEP 0u
#define LMC_STREAM_INIT 1u

// ------------------------------------------------------------------ real-type traits
template <typename T> struct Math;

template <> struct Math<float> {
  static constexpr bool is_double = false;
  __device__ __forceinline__ static float exp2(float x) { return lmc_ex2f(x); }
  __device__ __forceinline__ static float log2(float x) { return lmc_lg2f(x); }
  __device__ __forceinline__ static float exp(float x) { return lmc_ex2f(x * 1.4426950408889634f); }
  __device__ __forceinline__ static float log(float x) { return lmc_lg2f(x) * 0.6931471805599453f; }
  __device__ __forceinline__ static float sqrt(float x) { return lmc_sqrtf(x); }
  __device__ __forceinline__ static float rsqrt(float x) { return lmc_rsqrtf(x); }
  __device__ __forceinline__ static float rcp(float x) { return lmc_rcpf(x); }
  __device__ __forceinline__ static float div(float a, float b) { return a * lmc_rcpf(b); }
  // sinh(x) = (2^(x log2 e) - 2^(-x log2 e)) / 2; absolute error ~1e-7 cosh(x)
  __device__ __forceinline__ static float sinh(float x) {
    float a = x * 1.4426950408889634f;
    return 0.5f * (lmc_ex2f(a) - lmc_ex2f(-a));
  }
  // asinh(x) = sign(x) ln(|x| + sqrt(x^2 + 1))
  __device__ __forceinline__ static float asinh(float x) {
    float ax = fabsf(x);
    float l = lmc_lg2f(ax + lmc_sqrtf(fmaf(ax, ax, 1.0f))) * 0.6931471805599453f;
    return copysignf(l, x);
  }
  __device__ __forceinline__ static float abs(float x) { return fabsf(x); }
  __device__ __forceinline__ static float max(float a, float b) { return fmaxf(a, b); }
  __device__ __forceinline__ static float min(float a, float b) { return fminf(a, b); }
  __device__ __forceinline__ static float fma(float a, float b, float c) { return fmaf(a, b, c); }
  __device__ __forceinline__ static float floor(float a) { return floorf(a); }
  // two N(0,1) from two 32-bit words (Box-Muller) on the top 24 bits of each word:
  // u1 = (a24 + 1) 2^-24 in (0,1]; angle = 2 pi b24 2^-24 taken in [-pi, pi) (b24 read as a
  // signed 24-bit number: same cos/sin by periodicity, better conditioned for MUFU.SIN/COS).
  __device__ __forceinline__ static void normals(uint32_t a, uint32_t b, float& z0, float& z1) {
    float u1 = fmaf((float)(a >> 8), 5.9604644775390625e-8f, 5.9604644775390625e-8f);
    float ang = (float)((int32_t)b >> 8) * 3.7450703661175736e-7f;   // 2 pi 2^-24
    float r = lmc_sqrtf(lmc_lg2f(u1) * -1.3862943611198906f);        // sqrt(-2 ln u1)
    z0 = r * lmc_cosf(ang);
    z1 = r * lmc_sinf(ang);
  }
  __device__ __forceinline__ static float uniform(uint32_t a) {
    return (float)(a >> 8) * 5.9604644775390625e-8f;  // [0,1)
  }
};

template <> struct Math<double> {
  static constexpr bool is_double = true;
  __device__ __forceinline__ static double exp2(double x) { return ::exp2(x); }
  __device__ __forceinline__ static double log2(double x) { return ::log2(x); }
  __device__ __forceinline__ static double exp(double x) { return ::exp(x); }
  __device__ __forceinline__ static double log(double x) { return ::log(x); }
  __device__ __forceinline__ static double sqrt(double x) { return ::sqrt(x); }
  __device__ __forceinline__ static double rsqrt(double x) { return 1.0 / ::sqrt(x); }
  __device__ __forceinline__ static double rcp(double x) { return 1.0 / x; }
  __device__ __forceinline__ static double div(double a, double b) { return a / b; }
  __device__ __forceinline__ static double sinh(double x) { return ::sinh(x); }
  __device__ __forceinline__ static double asinh(double x) { return ::asinh(x); }
  __device__ __forceinline__ static double abs(double x) { return ::fabs(x); }
  __device__ __forceinline__ static double max(double a, double b) { return ::fmax(a, b); }
  __device__ __forceinline__ static double min(double a, double b) { return ::fmin(a, b); }
  __device__ __forceinline__ static double fma(double a, double b, double c) { return ::fma(a, b, c); }
  __device__ __forceinline__ static double floor(double a) { return ::floor(a); }
  __device__ __forceinline__ static void normals(uint32_t a, uint32_t b, double& z0, double& z1) {
    double u1 = ((double)a + 1.0) * 2.3283064365386963e-10;  // 2^-32, (0,1]
    double u2 = (double)b * 2.3283064365386963e-10;
    double r = ::sqrt(-2.0 * ::log(u1));
    double s, c;
    sincospi(2.0 * u2, &s, &c);
    z0 = r * c;
    z1 = r * s;
  }
  __device__ __forceinline__ static double uniform(uint32_t a) {
    return (double)a * 2.3283064365386963e-10;
  }
};

template <typename T>
__device__ __forceinline__ T lmc_soft(T x, T thr) {
  // prox.py:18-19: sign(x) * max(|x| - thr, 0)
  T m = Math<T>::max(Math<T>::abs(x) - thr, T(0));
  return copysign(m, x);
}

__device__ __forceinline__ double lmc_warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
