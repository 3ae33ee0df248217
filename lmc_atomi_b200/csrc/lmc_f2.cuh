// Packed fp32 pairs for the register-resident mixture chains (sm_100a): one thread advances TWO chains, the x and
// y halves of an `f2`, with Blackwell's packed FP32 instructions (PTX add/sub/mul/fma .f32x2 -> SASS FADD2 / FMUL2 /
// FFMA2).  A packed instruction retires two IEEE operations in one issue slot at the FLOP rate of the scalar forms
// (measured with scripts/ubench_issue.cu: FFMA2 0.5 warp-instructions per clock per sub-partition vs 1.0 for FFMA),
// so it frees issue slots -- the resource the chains are bound by (profiles/r01e: 78-84 % issue-active) -- for the
// integer (Philox), min/max and MUFU work that has no packed form.  Per-component constants live pre-duplicated in
// the kernel parameters, where ptxas feeds them to the packed instructions as 64-bit uniform-register operands;
// abs / neg of the halves fold into operand modifiers.
#pragma once
#include "lmc_common.cuh"

struct alignas(8) f2 {
  float x, y;
  __host__ __device__ __forceinline__ f2() {}
  __host__ __device__ __forceinline__ f2(float a) : x(a), y(a) {}
  __host__ __device__ __forceinline__ f2(double a) : x((float)a), y((float)a) {}
  __host__ __device__ __forceinline__ f2(int a) : x((float)a), y((float)a) {}
  __host__ __device__ __forceinline__ f2(float a, float b) : x(a), y(b) {}
};

struct b2 {   // per-half predicate
  bool x, y;
};
struct d2 {   // per-half fp64 accumulator
  double x, y;
};
struct u2 {   // per-half 32-bit word
  uint32_t x, y;
};

#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long f2_pack(f2 v) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(v.x), "f"(v.y));
  return r;
}
__device__ __forceinline__ f2 f2_unpack(unsigned long long v) {
  f2 r;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
  return r;
}
__device__ __forceinline__ f2 operator+(f2 a, f2 b) {
  unsigned long long d;
  asm("add.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)));
  return f2_unpack(d);
}
__device__ __forceinline__ f2 operator-(f2 a, f2 b) {
  unsigned long long d;
  asm("sub.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)));
  return f2_unpack(d);
}
__device__ __forceinline__ f2 operator*(f2 a, f2 b) {
  unsigned long long d;
  asm("mul.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)));
  return f2_unpack(d);
}
__device__ __forceinline__ f2 f2_fma(f2 a, f2 b, f2 c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)), "l"(f2_pack(c)));
  return f2_unpack(d);
}
__device__ __forceinline__ f2 operator-(f2 a) { return f2(-a.x, -a.y); }          // folds into a neg modifier
__device__ __forceinline__ f2& operator+=(f2& a, f2 b) { a = a + b; return a; }
__device__ __forceinline__ f2& operator-=(f2& a, f2 b) { a = a - b; return a; }

__device__ __forceinline__ f2 copysign(f2 m, f2 s) { return f2(copysignf(m.x, s.x), copysignf(m.y, s.y)); }

template <> struct Math<f2> {
  using F = Math<float>;
  static constexpr bool is_double = false;
#define LMC_F2_MAP1(name) \
  __device__ __forceinline__ static f2 name(f2 a) { return f2(F::name(a.x), F::name(a.y)); }
  LMC_F2_MAP1(exp2) LMC_F2_MAP1(log2) LMC_F2_MAP1(exp) LMC_F2_MAP1(log) LMC_F2_MAP1(sqrt) LMC_F2_MAP1(rsqrt)
  LMC_F2_MAP1(rcp) LMC_F2_MAP1(sinh) LMC_F2_MAP1(asinh) LMC_F2_MAP1(abs) LMC_F2_MAP1(floor)
#undef LMC_F2_MAP1
  __device__ __forceinline__ static f2 div(f2 a, f2 b) { return a * rcp(b); }
  __device__ __forceinline__ static f2 max(f2 a, f2 b) { return f2(fmaxf(a.x, b.x), fmaxf(a.y, b.y)); }
  __device__ __forceinline__ static f2 min(f2 a, f2 b) { return f2(fminf(a.x, b.x), fminf(a.y, b.y)); }
  __device__ __forceinline__ static f2 fma(f2 a, f2 b, f2 c) { return f2_fma(a, b, c); }
  __device__ __forceinline__ static void normals(u2 a, u2 b, f2& z0, f2& z1) {
    F::normals(a.x, b.x, z0.x, z1.x);
    F::normals(a.y, b.y, z0.y, z1.y);
  }
  __device__ __forceinline__ static f2 uniform(u2 a) { return f2(F::uniform(a.x), F::uniform(a.y)); }
};
#endif

// ------------------------------------------------------------------ lane traits: what differs between one chain
// per thread (float / double) and two chains per thread (f2)
template <typename real> struct Lanes {
  static constexpr int N = 1;
  using mask = bool;
  using dbl = double;
  using word = uint32_t;
  using count = uint32_t;
  struct cid { uint32_t lo, hi; };
  struct blk { uint32_t x, y, z, w; };
#ifdef __CUDACC__
  __device__ __forceinline__ static mask gt(real a, real b) { return a > b; }
  __device__ __forceinline__ static mask lt(real a, real b) { return a < b; }
  __device__ __forceinline__ static mask not_(mask m) { return !m; }
  __device__ __forceinline__ static bool any(mask m) { return m; }
  __device__ __forceinline__ static real select(mask m, real a, real b) { return m ? a : b; }
  __device__ __forceinline__ static void count_add(count& n, mask m) { n += m ? 1u : 0u; }
  __device__ __forceinline__ static void dbl_add(dbl& a, real b) { a += (double)b; }
  __device__ __forceinline__ static double dbl_total(dbl a, mask live) { return live ? a : 0.0; }
  __device__ __forceinline__ static dbl dbl_set(double v) { return v; }
  __device__ __forceinline__ static blk philox(uint32_t c0, uint32_t c1, cid c, const uint32_t* rk) {
    uint4 r = lmc_philox4x32_10_rk(make_uint4(c0, c1, c.lo, c.hi), rk);
    return blk{r.x, r.y, r.z, r.w};
  }
#endif
};

template <> struct Lanes<f2> {
  static constexpr int N = 2;
  using mask = b2;
  using dbl = d2;
  using word = u2;
  struct count { uint32_t x, y; };
  struct cid { uint32_t lo, hi, lo1, hi1; };
  struct blk { u2 x, y, z, w; };
#ifdef __CUDACC__
  __device__ __forceinline__ static mask gt(f2 a, f2 b) { return b2{a.x > b.x, a.y > b.y}; }
  __device__ __forceinline__ static mask lt(f2 a, f2 b) { return b2{a.x < b.x, a.y < b.y}; }
  __device__ __forceinline__ static mask not_(mask m) { return b2{!m.x, !m.y}; }
  __device__ __forceinline__ static bool any(mask m) { return m.x || m.y; }
  __device__ __forceinline__ static f2 select(mask m, f2 a, f2 b) { return f2(m.x ? a.x : b.x, m.y ? a.y : b.y); }
  __device__ __forceinline__ static void count_add(count& n, mask m) { n.x += m.x ? 1u : 0u; n.y += m.y ? 1u : 0u; }
  __device__ __forceinline__ static void dbl_add(dbl& a, f2 b) { a.x += (double)b.x; a.y += (double)b.y; }
  __device__ __forceinline__ static double dbl_total(dbl a, mask live) { return (live.x ? a.x : 0.0) + (live.y ? a.y : 0.0); }
  __device__ __forceinline__ static dbl dbl_set(double v) { return d2{v, v}; }
  __device__ __forceinline__ static blk philox(uint32_t c0, uint32_t c1, cid c, const uint32_t* rk) {
    uint4 a = lmc_philox4x32_10_rk(make_uint4(c0, c1, c.lo, c.hi), rk);
    uint4 b = lmc_philox4x32_10_rk(make_uint4(c0, c1, c.lo1, c.hi1), rk);
    return blk{u2{a.x, b.x}, u2{a.y, b.y}, u2{a.z, b.z}, u2{a.w, b.w}};
  }
#endif
};
