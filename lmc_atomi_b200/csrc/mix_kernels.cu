// Persistent register-resident Langevin chains on 2-D mixture targets.
//
// One thread = one chain; all K steps of the chain run inside ONE launch with the
// state (theta, Philox counter, running moments) held in registers.  A warp grabs
// 32 consecutive chains from a global queue (dynamic, warp-granular scheduling, so
// the 148 SMs stay balanced for any C), steps them K times and writes the final
// state, accept count and its moment partials.  Steady-state HBM traffic is zero:
// the kernel is bound by FP32/INT32 issue and the MUFU pipe (kernel M1).
//
// Reference loops replaced (file:line under /root/reference):
//   lmc.py:94-190, lmc_laplace.py:110-216, prox_lmc.py:99-255
// The target is evaluated in the log2 domain with a max-shift (log-sum-exp), which
// is algebraically identical to the reference's plain ratio -grad p / p
// (lmc.py:60-61) but cannot underflow in fp32.
#include <stdlib.h>

#include "lmc_common.cuh"
#include "lmc_f2.cuh"

namespace {

constexpr int NMAX = LMC_MAX_COMP;
constexpr int BLOCK = 256;

// build-time experiment switches (defaults = the shipped configuration; see profiles/r02g_mix_variants.txt)
// A/B on B200 at the benchmark shape (profiles/r02h_mix_variants.txt, G chain-steps/s, reproducible to 0.5 %):
//   shift + no hint 173.7 | shift + 3 CTAs/SM 176.7 | no-shift + no hint 176.9 | no-shift + 2 CTAs/SM 188.3 (shipped)
//   | no-shift + 3 CTAs/SM 186.5 | no-shift + 4 CTAs/SM 183.8        (scalar one-chain-per-thread kernels: 169.4)
#ifndef LMC_MIX_NOSHIFT
#define LMC_MIX_NOSHIFT 1
#endif
#ifndef LMC_MIX_MINBLOCKS
#define LMC_MIX_MINBLOCKS 2
#endif

template <typename real>
struct MixP {
  int n, t;
  // components
  real mu0[NMAX], mu1[NMAX];
  real p00[NMAX], p01[NMAX], p10[NMAX], p11[NMAX];  // Sigma^-1 (Gaussian kinds)
  real lc[NMAX];    // log2 of the component's constant factor (omega_i / norm_i or omega_i (alpha_i/2)^2)
  real al[NMAX];    // Laplacian scale alpha_i
  real thr[NMAX];   // lamda * alpha_i
  real c1[NMAX];    // -log2(e) alpha_i
  real c2;          // -log2(e) / (2 lamda)
  real lamda, inv_lamda;
  // l1 prior (GAUSS_L1)
  real pr_alpha, pr_thr, pr_mu0, pr_mu1;
  // step
  real ngs;   // -gamma (Gaussian mixture) or -gamma / lamda (Laplacian): scale from (gp0, gp1) / S to the ULA drift
  real gamma, s2g, inv4g, inv4g_l2e;   // inv4g_l2e = log2(e) / (4 gamma)
  real m00, m01, m10, m11;      // M
  real s00, s01, s10, s11;      // sqrtm(M)
  real gm00, gm01, gm10, gm11;  // -gamma M            (fp32 production PULA: the step as four FMAs per coordinate)
  real gs00, gs01, gs10, gs11;  // sqrt(2 gamma) sqrtm(M)
  real q00, q01, q10, q11;      // Q
  real qi00, qi01, qi10, qi11;  // Q^-1
  real eta, inv_eta, pp_thr;    // PPULA dual step, 1/eta, lamda/eta
  real b0, b1, sg0, sg1;        // beta, sigma
  real ib0, ib1, isg0, isg1;    // their reciprocals
  real hshift;
};

struct MixIo {
  void* theta;
  const void* noise;
  const void* unif;
  void* traj;
  uint8_t* accept;
  uint32_t* naccept;
  unsigned long long* hist;
  unsigned int* queue;   // workspace: work-item counter
  double* partial;       // workspace: (n_items, 6) moment partials
  long long C, K, burnin, hist_every;
  unsigned long long chain_offset;
  uint2 key;
  uint32_t rk[20];       // Philox round keys (host-computed from key)
  int init_from_rng, hist_bins;
  double hist_lo, hist_scale;  // bin = floor((x - lo) * scale)
  int n_items;
};

template <typename real>
struct Eval {
  real g0, g1;       // grad U
  real gp0, gp1, rS; // its pieces: grad U = (gp0, gp1) * rS [* 1/lamda for the Laplacian mixture]
  real logp2;        // log2 of the (smoothed) mixture density
  real h00, h01, h11;  // Hessian of U (only when requested)
};

constexpr float LOG2E_F = 1.4426950408889634f;

// ----------------------------------------------------------------------------------
// Gaussian mixture: lmc.py:39-75.  a_i = log2(omega_i N_i(theta)); softmax weights
// w_i = 2^(a_i - max a); grad U = sum w_i P_i (theta - mu_i) / sum w_i.
template <typename real, int NC, bool HESS>
__device__ __forceinline__ void eval_gauss(const MixP<real>& P, real t0, real t1, Eval<real>& e) {
  using M = Math<real>;
  using L = Lanes<real>;
  const int n = NC > 0 ? NC : P.n;
  constexpr bool NOSHIFT = LMC_MIX_NOSHIFT && !M::is_double && !HESS;   // see eval_laplace
  real a[NMAX], v0[NMAX], v1[NMAX];
  real amax = real(-1e30);
#pragma unroll
  for (int i = 0; i < NMAX; ++i) {
    if (i < n) {
      real d0 = t0 - P.mu0[i], d1 = t1 - P.mu1[i];
      v0[i] = M::fma(P.p00[i], d0, P.p01[i] * d1);
      v1[i] = M::fma(P.p10[i], d0, P.p11[i] * d1);
      real fac = M::fma(d0, v0[i], d1 * v1[i]);
      a[i] = M::fma(fac, real(-0.5 * 1.4426950408889634), P.lc[i]);
      if (!NOSHIFT) amax = M::max(amax, a[i]);
    }
  }
  real S, gp0, gp1, hp00 = real(0), hp01 = real(0), hp11 = real(0);
  auto sums = [&](real shift) {
    S = real(0); gp0 = real(0); gp1 = real(0);
#pragma unroll
    for (int i = 0; i < NMAX; ++i) {
      if (i < n) {
        real w = M::exp2(a[i] - shift);
        S += w;
        gp0 = M::fma(w, v0[i], gp0);  // note: this is -grad p (sign folded)
        gp1 = M::fma(w, v1[i], gp1);
        if (HESS) {  // lmc.py:63-65: N_i (P d d^T P - P)
          hp00 = M::fma(w, M::fma(v0[i], v0[i], -P.p00[i]), hp00);
          hp01 = M::fma(w, M::fma(v0[i], v1[i], -(real(0.5) * (P.p01[i] + P.p10[i]))), hp01);
          hp11 = M::fma(w, M::fma(v1[i], v1[i], -P.p11[i]), hp11);
        }
      }
    }
  };
  if (NOSHIFT) {
    amax = real(0);
    sums(real(0));
    if (L::any(L::lt(S, real(1e-30)))) {   // every component underflowed: redo with the max shift
      amax = real(-1e30);
#pragma unroll
      for (int i = 0; i < NMAX; ++i)
        if (i < n) amax = M::max(amax, a[i]);
      sums(amax);
    }
  } else {
    sums(amax);
  }
  real rS = M::rcp(S);
  e.gp0 = gp0; e.gp1 = gp1; e.rS = rS;
  e.g0 = gp0 * rS;
  e.g1 = gp1 * rS;
  e.logp2 = amax + M::log2(S);
  if (HESS) {  // lmc.py:71-75: grad p grad p^T / p^2 - hess p / p
    e.h00 = M::fma(e.g0, e.g0, -(hp00 * rS));
    e.h01 = M::fma(e.g0, e.g1, -(hp01 * rS));
    e.h11 = M::fma(e.g1, e.g1, -(hp11 * rS));
  }
}

// Moreau-smoothed Laplacian mixture: lmc_laplace.py:43-92.  With d = theta - mu_i and
// r = clamp(d, -lamda alpha_i, lamda alpha_i) = theta - prox (lmc_laplace.py:53-54), s = d - r = prox - mu_i:
// envelope = alpha_i |s|_1 + |r|^2 / (2 lamda) (56-58), component gradient direction r / lamda (70-71).
template <typename real, int NC, bool HESS>
__device__ __forceinline__ void eval_laplace(const MixP<real>& P, real t0, real t1, Eval<real>& e) {
  using M = Math<real>;
  using L = Lanes<real>;
  const int n = NC > 0 ? NC : P.n;
  constexpr bool NOSHIFT = LMC_MIX_NOSHIFT && !M::is_double && !HESS;
  real a[NMAX], r0[NMAX], r1[NMAX];
  real amax = real(-1e30);
#pragma unroll
  for (int i = 0; i < NMAX; ++i) {
    if (i < n) {
      real d0 = t0 - P.mu0[i], d1 = t1 - P.mu1[i];
      r0[i] = M::max(M::min(d0, P.thr[i]), -P.thr[i]);
      r1[i] = M::max(M::min(d1, P.thr[i]), -P.thr[i]);
      real s1n = M::abs(d0 - r0[i]) + M::abs(d1 - r1[i]);
      real rr = M::fma(r0[i], r0[i], r1[i] * r1[i]);
      a[i] = M::fma(rr, P.c2, M::fma(s1n, P.c1[i], P.lc[i]));   // log2(omega_i pdf_lamda,i)
      if (!NOSHIFT) amax = M::max(amax, a[i]);
    }
  }
  real S, gp0, gp1, hp00 = real(0), hp01 = real(0), hp11 = real(0);
  const real l1 = M::abs(t0) + M::abs(t1);
  auto sums = [&](real shift) {
    S = real(0); gp0 = real(0); gp1 = real(0);
#pragma unroll
    for (int i = 0; i < NMAX; ++i) {
      if (i < n) {
        real w = M::exp2(a[i] - shift);
        S += w;
        gp0 = M::fma(w, r0[i], gp0);
        gp1 = M::fma(w, r1[i], gp1);
        if (HESS) {
          // lmc_laplace.py:80-82 (sic): weight (alpha/2)^2 exp(-alpha ||theta||_1), no mu,
          // unsmoothed exponent; matrix (-I + g g^T) with g = r / lamda.
          real v0 = r0[i] * P.inv_lamda, v1 = r1[i] * P.inv_lamda;
          real wq = M::exp2(M::fma(l1, P.c1[i], P.lc[i]) - shift);
          hp00 = M::fma(wq, M::fma(v0, v0, real(-1)), hp00);
          hp01 = M::fma(wq, v0 * v1, hp01);
          hp11 = M::fma(wq, M::fma(v1, v1, real(-1)), hp11);
        }
      }
    }
  };
  if (NOSHIFT) {
    // fp32: every exponent a_i is <= lc_i = log2(omega_i x normalising constant), a finite per-component constant,
    // so 2^a_i cannot overflow unless the constant itself does; the log-sum-exp shift is only needed when EVERY
    // component underflows (a chain ~170 / alpha away from all modes), handled by the rare second evaluation below.
    // Saves the running max and the n subtractions of every evaluation (profiles/r02h_mix_variants.txt).
    amax = real(0);
    sums(real(0));
    if (L::any(L::lt(S, real(1e-30)))) {
      amax = real(-1e30);
#pragma unroll
      for (int i = 0; i < NMAX; ++i)
        if (i < n) amax = M::max(amax, a[i]);
      sums(amax);
    }
  } else {
    sums(amax);
  }
  const real rS = M::rcp(S), gS = rS * P.inv_lamda;
  e.gp0 = gp0; e.gp1 = gp1; e.rS = rS;
  e.g0 = gp0 * gS;
  e.g1 = gp1 * gS;
  e.logp2 = amax + M::log2(S);
  if (HESS) {
    e.h00 = M::fma(e.g0, e.g0, -(hp00 * rS));
    e.h01 = M::fma(e.g0, e.g1, -(hp01 * rS));
    e.h11 = M::fma(e.g1, e.g1, -(hp11 * rS));
  }
}

template <typename real, int TARGET, int NC, bool HESS>
__device__ __forceinline__ void eval_target(const MixP<real>& P, real t0, real t1, Eval<real>& e) {
  if (TARGET == LMC_TARGET_LAPLACE) eval_laplace<real, NC, HESS>(P, t0, t1, e);
  else eval_gauss<real, NC, HESS>(P, t0, t1, e);
}

// closed-form inverse and principal square root of a symmetric positive definite 2x2
template <typename real>
__device__ __forceinline__ void spd_inv_sqrt(real a, real b, real d, real& i00, real& i01, real& i11,
                                             real& r00, real& r01, real& r11) {
  using M = Math<real>;
  real det = M::fma(a, d, -b * b);
  real rdet = M::rcp(det);
  i00 = d * rdet; i01 = -b * rdet; i11 = a * rdet;          // M = H^-1
  real sdet = M::sqrt(M::fma(i00, i11, -i01 * i01));        // sqrt(det M)
  real tr = i00 + i11;
  real rs = M::rsqrt(tr + real(2) * sdet);
  r00 = (i00 + sdet) * rs; r01 = i01 * rs; r11 = (i11 + sdet) * rs;  // sqrtm(M)
}

// asinh(t / b) from sq = sqrt(t^2 + b^2) and ib = 1 / b
template <typename real>
__device__ __forceinline__ real mirror_asinh(real t, real sq, real ib) {
  using M = Math<real>;
  return copysign(M::log((M::abs(t) + sq) * ib), t);
}

template <int ALG> struct AlgTraits {
  static constexpr bool adjusted = (ALG == LMC_MALA || ALG == LMC_MYMALA);
  static constexpr bool hess = (ALG == LMC_IHPULA || ALG == LMC_FBULA);
};

// mean of the Langevin proposal: "gd_update" (lmc.py:77-78) plus, for MYMALA/MYULA,
// the Moreau-Yosida drift "prox_update" (prox_lmc.py:114-118)
template <typename real, int ALG>
__device__ __forceinline__ void drift_centre(const MixP<real>& P, real t0, real t1, const Eval<real>& e,
                                             real& c0, real& c1) {
  using M = Math<real>;
  c0 = M::fma(-P.gamma, e.g0, t0);
  c1 = M::fma(-P.gamma, e.g1, t1);
  if (ALG == LMC_MYULA || ALG == LMC_MYMALA) {
    real y0 = (t0 - lmc_soft(t0, P.pr_thr)) * P.inv_lamda;
    real y1 = (t1 - lmc_soft(t1, P.pr_thr)) * P.inv_lamda;
    c0 = M::fma(-P.gamma, y0, c0);
    c1 = M::fma(-P.gamma, y1, c1);
  }
}

// One transition of an unadjusted sampler.
template <typename real, int TARGET, int ALG, int NC>
__device__ __forceinline__ void step_unadjusted(const MixP<real>& P, real& t0, real& t1, real x0, real x1) {
  using M = Math<real>;
  using L = Lanes<real>;
  Eval<real> e;
  if (ALG == LMC_PGLD) {  // prox_lmc.py:106
    t0 = lmc_soft(t0, P.pr_thr);
    t1 = lmc_soft(t1, P.pr_thr);
  }
  eval_target<real, TARGET, NC, AlgTraits<ALG>::hess>(P, t0, t1, e);
  if (ALG == LMC_ULA && !M::is_double) {
    // float32: -gamma [/ lamda] folded into the normalisation 1 / S, one multiply instead of three
    const real k = e.rS * P.ngs;
    t0 = M::fma(P.s2g, x0, M::fma(e.gp0, k, t0));
    t1 = M::fma(P.s2g, x1, M::fma(e.gp1, k, t1));
  } else if (ALG == LMC_ULA || ALG == LMC_PGLD || ALG == LMC_MYULA) {
    real c0, c1;
    drift_centre<real, ALG>(P, t0, t1, e, c0, c1);
    t0 = M::fma(P.s2g, x0, c0);
    t1 = M::fma(P.s2g, x1, c1);
  } else if (ALG == LMC_PULA && !M::is_double) {
    // theta - gamma M grad U + sqrt(2 gamma) sqrtm(M) xi with the scalars folded into the matrices on the host
    // (float32 only: float64 keeps the reference's order of operations below)
    t0 = M::fma(P.gs00, x0, M::fma(P.gs01, x1, M::fma(P.gm00, e.g0, M::fma(P.gm01, e.g1, t0))));
    t1 = M::fma(P.gs10, x0, M::fma(P.gs11, x1, M::fma(P.gm10, e.g0, M::fma(P.gm11, e.g1, t1))));
  } else if (ALG == LMC_PULA || ALG == LMC_PPULA) {
    real n0 = M::fma(P.m00, e.g0, P.m01 * e.g1), n1 = M::fma(P.m10, e.g0, P.m11 * e.g1);
    real z0 = M::fma(P.s00, x0, P.s01 * x1), z1 = M::fma(P.s10, x0, P.s11 * x1);
    real c0 = M::fma(-P.gamma, n0, t0), c1 = M::fma(-P.gamma, n1, t1);
    if (ALG == LMC_PPULA) {
      // prox_lmc.py:165-176: dual forward-backward iterations for prox in the Q metric,
      // called with threshold lamda (sic, not lamda*alpha); returns the last primal u.
      real w0 = 0, w1 = 0, u0 = t0, u1 = t1;
      for (int it = 0; it < P.t; ++it) {
        u0 = t0 - M::fma(P.q00, w0, P.q01 * w1);
        u1 = t1 - M::fma(P.q10, w0, P.q11 * w1);
        real y0 = M::fma(w0, P.inv_eta, u0), y1 = M::fma(w1, P.inv_eta, u1);
        w0 = M::fma(P.eta, u0, w0) - P.eta * lmc_soft(y0, P.pp_thr);
        w1 = M::fma(P.eta, u1, w1) - P.eta * lmc_soft(y1, P.pp_thr);
      }
      real r0 = t0 - u0, r1 = t1 - u1;
      real k0 = M::fma(P.qi00, r0, P.qi01 * r1) * P.inv_lamda;
      real k1 = M::fma(P.qi10, r0, P.qi11 * r1) * P.inv_lamda;
      c0 = M::fma(-P.gamma, k0, c0);
      c1 = M::fma(-P.gamma, k1, c1);
    }
    t0 = M::fma(P.s2g, z0, c0);
    t1 = M::fma(P.s2g, z1, c1);
  } else if (ALG == LMC_IHPULA) {
    // lmc.py:159-166: M = (H + (|lambda_min| + shift) I)^-1 (n>1) or H^-1 (n==1)
    real a = e.h00, b = e.h01, d = e.h11;
    if (P.n > 1) {
      real hm = real(0.5) * (a + d), hd = real(0.5) * (a - d);
      real emin = hm - M::sqrt(M::fma(hd, hd, b * b));
      real sh = M::abs(emin) + P.hshift;
      a += sh; d += sh;
    }
    real i00, i01, i11, r00, r01, r11;
    spd_inv_sqrt(a, b, d, i00, i01, i11, r00, r01, r11);
    real n0 = M::fma(i00, e.g0, i01 * e.g1), n1 = M::fma(i01, e.g0, i11 * e.g1);
    real z0 = M::fma(r00, x0, r01 * x1), z1 = M::fma(r01, x0, r11 * x1);
    t0 = M::fma(P.s2g, z0, M::fma(-P.gamma, n0, t0));
    t1 = M::fma(P.s2g, z1, M::fma(-P.gamma, n1, t1));
  } else if (ALG == LMC_MLA || ALG == LMC_LBMUMLA) {
    // lmc.py:186-187 / prox_lmc.py:251-252: mirror map grad = asinh(theta/beta), evaluated as
    // sign(theta) log((|theta| + sqrt(theta^2 + beta^2)) / beta): the root is shared with the
    // noise scale (theta^2 + beta^2)^(-1/4).
    real sq0 = M::sqrt(M::fma(t0, t0, P.b0 * P.b0)), sq1 = M::sqrt(M::fma(t1, t1, P.b1 * P.b1));
    real e0 = mirror_asinh(t0, sq0, P.ib0) - P.gamma * e.g0;
    real e1 = mirror_asinh(t1, sq1, P.ib1) - P.gamma * e.g1;
    if (ALG == LMC_LBMUMLA) {
      // prox_lmc.py:218-242 with beta := sigma (the reference passes sigma there)
      real thr = P.pr_thr;
      real sh = M::sinh(thr);
      {
        real rt = M::sqrt(M::fma(t0, t0, P.sg0 * P.sg0));
        real as = mirror_asinh(t0, rt, P.isg0);
        real lo = P.sg0 * M::sinh(as - thr), hi = P.sg0 * M::sinh(as + thr);
        real p = L::select(L::gt(t0, P.sg0 * sh), lo, rt - P.sg0);
        p = L::select(L::lt(t0, -(P.sg0 * sh)), hi, p);
        e0 -= P.gamma * (M::div(P.inv_lamda, rt) * (t0 - p));
      }
      {
        real rt = M::sqrt(M::fma(t1, t1, P.sg1 * P.sg1));
        real as = mirror_asinh(t1, rt, P.isg1);
        real lo = P.sg1 * M::sinh(as - thr), hi = P.sg1 * M::sinh(as + thr);
        real p = L::select(L::gt(t1, P.sg1 * sh), lo, rt - P.sg1);
        p = L::select(L::lt(t1, -(P.sg1 * sh)), hi, p);
        e1 -= P.gamma * (M::div(P.inv_lamda, rt) * (t1 - p));
      }
    }
    e0 = M::fma(P.s2g * M::rsqrt(sq0), x0, e0);
    e1 = M::fma(P.s2g * M::rsqrt(sq1), x1, e1);
    t0 = P.b0 * M::sinh(e0);
    t1 = P.b1 * M::sinh(e1);
  } else if (ALG == LMC_FBULA) {
    // prox_lmc.py:192-196: (I - lamda H)(theta - soft(theta - lamda grad U, lamda alpha))/lamda
    real f0 = M::fma(-P.lamda, e.g0, t0), f1 = M::fma(-P.lamda, e.g1, t1);
    real r0 = t0 - lmc_soft(f0, P.pr_thr), r1 = t1 - lmc_soft(f1, P.pr_thr);
    real a = real(1) - P.lamda * e.h00, b = -P.lamda * e.h01, d = real(1) - P.lamda * e.h11;
    real k0 = M::fma(a, r0, b * r1) * P.inv_lamda, k1 = M::fma(b, r0, d * r1) * P.inv_lamda;
    t0 = M::fma(P.s2g, x0, M::fma(-P.gamma, k0, t0));
    t1 = M::fma(P.s2g, x1, M::fma(-P.gamma, k1, t1));
  }
}

// log2 of the density entering the MH ratio (lmc.py:113, lmc_laplace.py:131, prox_lmc.py:139-140)
template <typename real, int ALG>
__device__ __forceinline__ real mh_logdens2(const MixP<real>& P, real t0, real t1, const Eval<real>& e) {
  using M = Math<real>;
  real l = e.logp2;
  if (ALG == LMC_MYMALA)
    l -= real(1.4426950408889634) * P.pr_alpha * (M::abs(t0 - P.pr_mu0) + M::abs(t1 - P.pr_mu1));
  return l;
}

// State of one chain (registers of one thread) + the per-step arithmetic shared by the
// validation loop (host-supplied noise, trajectory written every step) and the production
// loop (Philox noise, nothing but registers touched inside the K-loop).
// MIRROR (production MLA in fp32): the chain is carried in mirror coordinates.  theta_{k+1} = beta sinh(y_{k+1}) with
// y_{k+1} = asinh(theta_k / beta) - gamma grad U(theta_k) + noise (lmc.py:186-187), so asinh(theta_{k+1} / beta) IS
// y_{k+1} and sqrt(theta_{k+1}^2 + beta^2) IS beta cosh(y_{k+1}): both come out of the two exponentials the sinh
// already costs.  Saves a square root and a logarithm per coordinate and step (4 of the ~20 MUFU operations of an
// MLA step, the pipe that bounds it: XU 74 % active in profiles/r02i) at the price of not re-rounding y through
// sinh/asinh every step -- a relative 1e-7 per step, the size of the fp32 arithmetic itself.  Validation mode and
// float64 keep the reference's evaluation order.
#ifndef LMC_MIX_MLA_MIRROR
#define LMC_MIX_MLA_MIRROR 1
#endif

template <typename real, int TARGET, int ALG, int NC, bool MIRROR = false>
struct Chain {
  using M = Math<real>;
  using L = Lanes<real>;
  using mask = typename L::mask;
  static constexpr bool ADJ = AlgTraits<ALG>::adjusted;
  real t0, t1;
  real my0, my1, msq0, msq1;   // MIRROR: y = asinh(theta / beta) and sqrt(theta^2 + beta^2) of the current state
  real cc0, cc1, clog;   // MH chains: proposal centre and log2-density at the current state
  typename L::count nacc;
  real b_n, b_s0, b_s1, b_s00, b_s01, b_s11;        // fp32/fp64 partial sums of <= 32 steps
  typename L::dbl m_n, m_s0, m_s1, m_s00, m_s01, m_s11;      // fp64 totals (per lane)

  __device__ __forceinline__ void init(const MixP<real>& P, real a0, real a1) {
    t0 = a0; t1 = a1;
    cc0 = cc1 = clog = real(0);
    nacc = typename L::count{};
    b_n = b_s0 = b_s1 = b_s00 = b_s01 = b_s11 = real(0);
    m_n = m_s0 = m_s1 = m_s00 = m_s01 = m_s11 = L::dbl_set(0.0);
    if (ADJ) {
      Eval<real> cur;
      eval_target<real, TARGET, NC, false>(P, t0, t1, cur);
      drift_centre<real, ALG>(P, t0, t1, cur, cc0, cc1);
      clog = mh_logdens2<real, ALG>(P, t0, t1, cur);
    }
    my0 = my1 = msq0 = msq1 = real(0);
    if (MIRROR) {
      msq0 = M::sqrt(M::fma(t0, t0, P.b0 * P.b0));
      msq1 = M::sqrt(M::fma(t1, t1, P.b1 * P.b1));
      my0 = mirror_asinh(t0, msq0, P.ib0);
      my1 = mirror_asinh(t1, msq1, P.ib1);
    }
  }

  // MLA step in mirror coordinates (see MIRROR above); same operations as step_unadjusted otherwise
  __device__ __forceinline__ void step_mirror(const MixP<real>& P, real x0, real x1) {
    Eval<real> e;
    eval_target<real, TARGET, NC, false>(P, t0, t1, e);
    real y0 = my0 - P.gamma * e.g0, y1 = my1 - P.gamma * e.g1;
    y0 = M::fma(P.s2g * M::rsqrt(msq0), x0, y0);
    y1 = M::fma(P.s2g * M::rsqrt(msq1), x1, y1);
    const real a0 = y0 * real(1.4426950408889634), a1 = y1 * real(1.4426950408889634);
    const real E0 = M::exp2(a0), F0 = M::exp2(-a0), E1 = M::exp2(a1), F1 = M::exp2(-a1);
    t0 = P.b0 * (real(0.5) * (E0 - F0));    // beta sinh(y), as Math<float>::sinh evaluates it
    t1 = P.b1 * (real(0.5) * (E1 - F1));
    msq0 = P.b0 * (real(0.5) * (E0 + F0));  // beta cosh(y) = sqrt(theta^2 + beta^2)
    msq1 = P.b1 * (real(0.5) * (E1 + F1));
    my0 = y0;
    my1 = y1;
  }

  // one transition; returns whether a sample was emitted (always, for unadjusted samplers)
  __device__ __forceinline__ mask step(const MixP<real>& P, real x0, real x1, real u) {
    if (!ADJ) {
      if (MIRROR) step_mirror(P, x0, x1);
      else step_unadjusted<real, TARGET, ALG, NC>(P, t0, t1, x0, x1);
      return L::not_(L::gt(real(0), real(0)));   // all lanes emit
    }
    // lmc.py:124-130: propose, MH ratio p(new) q(old|new) / (p(old) q(new|old)),
    // accept iff u <= min(1, ratio); a NaN ratio accepts (Python's min(1, nan) == 1).
    real n0 = M::fma(P.s2g, x0, cc0), n1 = M::fma(P.s2g, x1, cc1);
    Eval<real> prop;
    eval_target<real, TARGET, NC, false>(P, n0, n1, prop);
    real pc0, pc1;
    drift_centre<real, ALG>(P, n0, n1, prop, pc0, pc1);
    real plog = mh_logdens2<real, ALG>(P, n0, n1, prop);
    real f0 = t0 - pc0, f1 = t1 - pc1;   // old | new
    real r0 = n0 - cc0, r1 = n1 - cc1;   // new | old
    // log2 of the ratio: (log2 p(new) - log2 p(old)) + log2(e) (|new - c_old|^2 - |old - c_new|^2) / (4 gamma)
    real lr2 = M::fma(M::fma(r0, r0, r1 * r1) - M::fma(f0, f0, f1 * f1), P.inv4g_l2e, plog - clog);
    real ratio = M::exp2(lr2);
    // u < 1, so u > min(1, ratio) <=> u > ratio; a NaN ratio compares false and accepts, like Python's min(1, nan) == 1
    const mask emitted = L::not_(L::gt(u, ratio));
    t0 = L::select(emitted, n0, t0);
    t1 = L::select(emitted, n1, t1);
    cc0 = L::select(emitted, pc0, cc0);
    cc1 = L::select(emitted, pc1, cc1);
    clog = L::select(emitted, plog, clog);
    L::count_add(nacc, emitted);
    return emitted;
  }

  // running sums over the emitted samples
  __device__ __forceinline__ void accumulate(mask emitted) {
    if (ADJ) {
      real f = L::select(emitted, real(1), real(0));
      b_n += f;
      b_s0 = M::fma(f, t0, b_s0);
      b_s1 = M::fma(f, t1, b_s1);
      b_s00 = M::fma(f * t0, t0, b_s00);
      b_s01 = M::fma(f * t0, t1, b_s01);
      b_s11 = M::fma(f * t1, t1, b_s11);
    } else {
      b_s0 += t0;
      b_s1 += t1;
      b_s00 = M::fma(t0, t0, b_s00);
      b_s01 = M::fma(t0, t1, b_s01);
      b_s11 = M::fma(t1, t1, b_s11);
    }
  }

  __device__ __forceinline__ void flush() {
    L::dbl_add(m_n, b_n); L::dbl_add(m_s0, b_s0); L::dbl_add(m_s1, b_s1);
    L::dbl_add(m_s00, b_s00); L::dbl_add(m_s01, b_s01); L::dbl_add(m_s11, b_s11);
    b_n = b_s0 = b_s1 = b_s00 = b_s01 = b_s11 = real(0);
  }
};

__device__ __forceinline__ void hist_add(const MixIo& io, double a0, double a1) {
  int ix = (int)floor((a0 - io.hist_lo) * io.hist_scale);
  int iy = (int)floor((a1 - io.hist_lo) * io.hist_scale);
  if (ix >= 0 && ix < io.hist_bins && iy >= 0 && iy < io.hist_bins)
    atomicAdd(&io.hist[ix * io.hist_bins + iy], 1ull);
}

// histogram of the emitted, live lanes of a thread's state
template <typename real>
__device__ __forceinline__ void hist_lanes(const MixIo& io, real t0, real t1, typename Lanes<real>::mask emitted,
                                           typename Lanes<real>::mask live) {
  if constexpr (Lanes<real>::N == 2) {
    if (emitted.x && live.x) hist_add(io, (double)t0.x, (double)t1.x);
    if (emitted.y && live.y) hist_add(io, (double)t0.y, (double)t1.y);
  } else {
    if (emitted && live) hist_add(io, (double)t0, (double)t1);
  }
}

// Production loop over steps [kb, ke): Philox noise, registers only.  STATS = accumulate
// moments (steps past the burn-in), HIST = bin every io.hist_every-th emitted step.
// Unadjusted samplers use one Philox block per TWO steps (words x,y for even k, z,w for odd
// k, block counter k >> 1); adjusted samplers three words per step (normals, normals, uniform) taken consecutively
// from the blocks 0, 1, 2, ... of the chain: word 3k + i lives in block (3k + i) >> 2, three blocks per four steps.
template <typename real, int TARGET, int ALG, int NC, bool STATS, bool HIST, bool MIRROR>
__device__ __forceinline__ void run_production(Chain<real, TARGET, ALG, NC, MIRROR>& ch, const MixP<real>& P,
                                               const MixIo& io, typename Lanes<real>::cid cid,
                                               typename Lanes<real>::mask live, long long kb, long long ke) {
  using M = Math<real>;
  using L = Lanes<real>;
  constexpr bool ADJ = AlgTraits<ALG>::adjusted;
  long long hcount = 0;
  if (HIST) hcount = io.hist_every - (kb % io.hist_every);
  auto after = [&](typename L::mask emitted) {
    if (STATS) ch.accumulate(emitted);
    if (HIST) {
      if (--hcount == 0) {
        hcount = io.hist_every;
        hist_lanes<real>(io, ch.t0, ch.t1, emitted, live);
      }
    }
  };
  auto block = [&](long long ctr) {
    return L::philox((uint32_t)ctr, LMC_STREAM_STEP | ((uint32_t)(ctr >> 32) << 8), cid, io.rk);
  };
  long long k = kb;
  if (ADJ) {
    // step k consumes words 3k, 3k + 1 (normals) and 3k + 2 (uniform) of the chain's stream: 3 blocks per 4 steps
    using blk = decltype(block(0));
    auto pick = [](const blk& a, const blk& b, int i) {
      switch (i) {
        case 0: return a.x;
        case 1: return a.y;
        case 2: return a.z;
        case 3: return a.w;
        case 4: return b.x;
        case 5: return b.y;
        default: return b.z;
      }
    };
    auto one = [&](long long kk) {   // unaligned head / tail steps
      const long long w0 = 3 * kk;
      const int o = (int)(w0 & 3);
      const blk a = block(w0 >> 2), b = block((w0 >> 2) + 1);
      real x0, x1;
      M::normals(pick(a, b, o), pick(a, b, o + 1), x0, x1);
      after(ch.step(P, x0, x1, M::uniform(pick(a, b, o + 2))));
    };
    for (; k < ke && (k & 3); ++k) one(k);
    while (k + 4 <= ke) {
      const long long groups = (ke - k) >> 2;
      const int n = (int)(groups < 8 ? groups : 8);
      const long long b0 = 3 * (k >> 2);
      for (int j = 0; j < n; ++j) {
        real x0, x1;
        const blk a = block(b0 + 3 * j);
        M::normals(a.x, a.y, x0, x1);
        after(ch.step(P, x0, x1, M::uniform(a.z)));
        const blk b = block(b0 + 3 * j + 1);
        M::normals(a.w, b.x, x0, x1);
        after(ch.step(P, x0, x1, M::uniform(b.y)));
        const blk c = block(b0 + 3 * j + 2);
        M::normals(b.z, b.w, x0, x1);
        after(ch.step(P, x0, x1, M::uniform(c.x)));
        M::normals(c.y, c.z, x0, x1);
        after(ch.step(P, x0, x1, M::uniform(c.w)));
      }
      if (STATS) ch.flush();
      k += 4 * n;
    }
    for (; k < ke; ++k) one(k);
    if (STATS) ch.flush();
    return;
  }
  if ((k & 1) && k < ke) {   // odd start (odd burn-in): second half of block k >> 1
    auto r = block(k >> 1);
    real x0, x1;
    M::normals(r.z, r.w, x0, x1);
    after(ch.step(P, x0, x1, real(0)));
    ++k;
  }
  while (k + 1 < ke) {
    const long long left = (ke - k) >> 1;
    const int n = (int)(left < 16 ? left : 16);
    const long long c0 = k >> 1;
    for (int j = 0; j < n; ++j) {
      auto r = block(c0 + j);
      real x0, x1;
      M::normals(r.x, r.y, x0, x1);
      after(ch.step(P, x0, x1, real(0)));
      M::normals(r.z, r.w, x0, x1);
      after(ch.step(P, x0, x1, real(0)));
    }
    if (STATS) ch.flush();
    k += 2 * n;
  }
  if (k < ke) {
    auto r = block(k >> 1);
    real x0, x1;
    M::normals(r.x, r.y, x0, x1);
    after(ch.step(P, x0, x1, real(0)));
  }
  if (STATS) ch.flush();
}

template <typename real> struct ScalarOf { using type = real; };
template <> struct ScalarOf<f2> { using type = float; };

// Occupancy experiment (r01e, K=20000, 2^20 chains, fp32 production chains): a minimum of 5 resident CTAs per SM
// (48 registers) gave 167.9, 6 (40 registers, spills) 164.1, and even the hint "4" (which only changes the
// register allocator's target) 165.8 against 170.8 G chain-steps/s without any hint: the chains are bound by the
// MUFU/FMA pipes (math-pipe throttle + fixed-latency waits), not by occupancy.  No hint.
//
// real = f2 (production fp32): every thread advances TWO chains, lanes x / y of the packed type, chains
// item*64 + lane and item*64 + 32 + lane (both halves of a warp's loads and stores stay coalesced).
template <typename real, int TARGET, int ALG, int NC, bool PROD>
__global__ void __launch_bounds__(BLOCK, (Lanes<real>::N == 2 ? LMC_MIX_MINBLOCKS : 1)) mix_chain_kernel(const __grid_constant__ MixP<real> P,
                                                          const __grid_constant__ MixIo io) {
  using M = Math<real>;
  using L = Lanes<real>;
  using sreal = typename ScalarOf<real>::type;
  constexpr bool ADJ = AlgTraits<ALG>::adjusted;
  constexpr int PER_ITEM = 32 * L::N;
  const int lane = threadIdx.x & 31;
  sreal* theta = (sreal*)io.theta;
  const long long K = io.K;

  for (;;) {
    unsigned int item = 0;
    if (lane == 0) item = atomicAdd(io.queue, 1u);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= (unsigned)io.n_items) break;
    const long long c = (long long)item * PER_ITEM + lane;
    typename L::mask live;
    typename L::cid cid;
    real a0 = real(0), a1 = real(0);
    if constexpr (L::N == 2) {
      const long long cy = c + 32;
      live = b2{c < io.C, cy < io.C};
      const unsigned long long gx = io.chain_offset + (unsigned long long)(live.x ? c : 0);
      const unsigned long long gy = io.chain_offset + (unsigned long long)(live.y ? cy : 0);
      cid = typename L::cid{(uint32_t)gx, (uint32_t)(gx >> 32), (uint32_t)gy, (uint32_t)(gy >> 32)};
      if (!io.init_from_rng) {
        if (live.x) { a0.x = theta[2 * c]; a1.x = theta[2 * c + 1]; }
        if (live.y) { a0.y = theta[2 * cy]; a1.y = theta[2 * cy + 1]; }
      }
    } else {
      live = c < io.C;
      const unsigned long long gc = io.chain_offset + (unsigned long long)(live ? c : 0);
      cid = typename L::cid{(uint32_t)gc, (uint32_t)(gc >> 32)};
      if (!io.init_from_rng && live) {
        a0 = theta[2 * c];
        a1 = theta[2 * c + 1];
      }
    }
    if (io.init_from_rng) {
      auto r = L::philox(0u, LMC_STREAM_INIT, cid, io.rk);
      M::normals(r.x, r.y, a0, a1);
    }
    constexpr bool MIRROR = LMC_MIX_MLA_MIRROR && PROD && ALG == LMC_MLA && !M::is_double;
    Chain<real, TARGET, ALG, NC, MIRROR> ch;
    ch.init(P, a0, a1);

    if constexpr (PROD) {
      const long long kb = io.burnin < K ? (io.burnin > 0 ? io.burnin : 0) : K;
      const bool hist = io.hist != nullptr && io.hist_every > 0;
      run_production<real, TARGET, ALG, NC, false, false, MIRROR>(ch, P, io, cid, live, 0, kb);
      if (hist) run_production<real, TARGET, ALG, NC, true, true, MIRROR>(ch, P, io, cid, live, kb, K);
      else run_production<real, TARGET, ALG, NC, true, false, MIRROR>(ch, P, io, cid, live, kb, K);
    } else {
      // validation mode (one chain per thread): the reference's own draws come from HBM, every state goes back
      static_assert(PROD || L::N == 1, "validation mode runs one chain per thread");
      const real* noise = (const real*)io.noise;
      const real* unif = (const real*)io.unif;
      real* traj = (real*)io.traj;
      typename L::blk rnd{};
      for (long long k = 0; k < K; ++k) {
        real x0, x1, u = 0;
        if (noise != nullptr) {
          x0 = live ? noise[(c * K + k) * 2] : real(0);
          x1 = live ? noise[(c * K + k) * 2 + 1] : real(0);
          if (ADJ) u = live ? unif[c * K + k] : real(0);
        } else if (ADJ) {
          rnd = L::philox((uint32_t)k, LMC_STREAM_STEP | ((uint32_t)(k >> 32) << 8), cid, io.rk);
          M::normals(rnd.x, rnd.y, x0, x1);
          u = M::uniform(rnd.z);
        } else if ((k & 1) == 0) {
          long long kb = k >> 1;
          rnd = L::philox((uint32_t)kb, LMC_STREAM_STEP | ((uint32_t)(kb >> 32) << 8), cid, io.rk);
          M::normals(rnd.x, rnd.y, x0, x1);
        } else {
          M::normals(rnd.z, rnd.w, x0, x1);
        }
        const bool emitted = ch.step(P, x0, x1, u);
        if (ADJ && io.accept != nullptr && live) io.accept[c * K + k] = emitted ? 1 : 0;
        if (traj != nullptr && live) {
          traj[(c * K + k) * 2] = ch.t0;
          traj[(c * K + k) * 2 + 1] = ch.t1;
        }
        if (k >= io.burnin) {
          ch.accumulate(emitted);
          if ((k & 31) == 31) ch.flush();
          if (io.hist != nullptr && io.hist_every > 0 && emitted && live && ((k + 1) % io.hist_every == 0))
            hist_add(io, (double)ch.t0, (double)ch.t1);
        }
      }
      ch.flush();
    }

    // unadjusted samplers emit every step: the count is known without a per-step add
    if (!ADJ) ch.m_n = L::dbl_set((double)(K - (io.burnin < K ? (io.burnin > 0 ? io.burnin : 0) : K)));
    double m_n = L::dbl_total(ch.m_n, live), m_s0 = L::dbl_total(ch.m_s0, live), m_s1 = L::dbl_total(ch.m_s1, live);
    double m_s00 = L::dbl_total(ch.m_s00, live), m_s01 = L::dbl_total(ch.m_s01, live),
           m_s11 = L::dbl_total(ch.m_s11, live);
    if constexpr (L::N == 2) {
      const long long cy = c + 32;
      if (live.x) {
        theta[2 * c] = ch.t0.x;
        theta[2 * c + 1] = ch.t1.x;
        if (io.naccept != nullptr) io.naccept[c] = ADJ ? ch.nacc.x : (uint32_t)K;
      }
      if (live.y) {
        theta[2 * cy] = ch.t0.y;
        theta[2 * cy + 1] = ch.t1.y;
        if (io.naccept != nullptr) io.naccept[cy] = ADJ ? ch.nacc.y : (uint32_t)K;
      }
      if (io.hist != nullptr && io.hist_every == 0) hist_lanes<real>(io, ch.t0, ch.t1, b2{true, true}, live);
    } else {
      if (live) {
        theta[2 * c] = ch.t0;
        theta[2 * c + 1] = ch.t1;
        if (io.naccept != nullptr) io.naccept[c] = ADJ ? ch.nacc : (uint32_t)K;
        if (io.hist != nullptr && io.hist_every == 0) hist_add(io, (double)ch.t0, (double)ch.t1);
      }
    }
    m_n = lmc_warp_sum(m_n); m_s0 = lmc_warp_sum(m_s0); m_s1 = lmc_warp_sum(m_s1);
    m_s00 = lmc_warp_sum(m_s00); m_s01 = lmc_warp_sum(m_s01); m_s11 = lmc_warp_sum(m_s11);
    if (lane == 0) {
      double* p = io.partial + (size_t)item * 6;
      p[0] = m_n; p[1] = m_s0; p[2] = m_s1; p[3] = m_s00; p[4] = m_s01; p[5] = m_s11;
    }
  }
}

// fixed-order reduction of the per-item partials -> moments (deterministic)
__global__ void __launch_bounds__(256) mix_reduce_moments(const double* __restrict__ partial, int n_items,
                                                          double* __restrict__ out) {
  __shared__ double sm[256][6];
  double acc[6] = {0, 0, 0, 0, 0, 0};
  for (int i = threadIdx.x; i < n_items; i += 256)
#pragma unroll
    for (int j = 0; j < 6; ++j) acc[j] += partial[(size_t)i * 6 + j];
#pragma unroll
  for (int j = 0; j < 6; ++j) sm[threadIdx.x][j] = acc[j];
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s)
#pragma unroll
      for (int j = 0; j < 6; ++j) sm[threadIdx.x][j] += sm[threadIdx.x + s][j];
    __syncthreads();
  }
  if (threadIdx.x < 6) out[threadIdx.x] += sm[0][threadIdx.x];
}

template <typename real>
__global__ void philox_draws_kernel(int adjusted, long long C, long long K, uint2 key,
                                    unsigned long long chain_offset, real* normals, real* unif, real* theta0) {
  using M = Math<real>;
  uint32_t rk[20];
  lmc_philox_round_keys(key, rk);
  long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  unsigned long long gc = chain_offset + (unsigned long long)c;
  uint32_t c_lo = (uint32_t)gc, c_hi = (uint32_t)(gc >> 32);
  if (theta0 != nullptr) {
    uint4 r = lmc_philox4x32_10_rk(make_uint4(0u, LMC_STREAM_INIT, c_lo, c_hi), rk);
    M::normals(r.x, r.y, theta0[2 * c], theta0[2 * c + 1]);
  }
  uint4 rnd = make_uint4(0, 0, 0, 0);
  for (long long k = 0; k < K; ++k) {
    real x0, x1;
    if (adjusted) {   // words 3k, 3k + 1, 3k + 2 of the chain's stream (run_production)
      const long long w0 = 3 * k, bi = w0 >> 2;
      const int o = (int)(w0 & 3);
      const uint4 a = lmc_philox4x32_10_rk(make_uint4((uint32_t)bi, LMC_STREAM_STEP | ((uint32_t)(bi >> 32) << 8), c_lo, c_hi), rk);
      const uint4 b = lmc_philox4x32_10_rk(make_uint4((uint32_t)(bi + 1), LMC_STREAM_STEP | ((uint32_t)((bi + 1) >> 32) << 8), c_lo, c_hi), rk);
      const uint32_t wd[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      M::normals(wd[o], wd[o + 1], x0, x1);
      if (unif != nullptr) unif[c * K + k] = M::uniform(wd[o + 2]);
    } else if ((k & 1) == 0) {
      long long kb = k >> 1;
      rnd = lmc_philox4x32_10_rk(make_uint4((uint32_t)kb, LMC_STREAM_STEP | ((uint32_t)(kb >> 32) << 8), c_lo, c_hi), rk);
      M::normals(rnd.x, rnd.y, x0, x1);
    } else {
      M::normals(rnd.z, rnd.w, x0, x1);
    }
    normals[(c * K + k) * 2] = x0;
    normals[(c * K + k) * 2 + 1] = x1;
  }
}

// ----------------------------------------------------------------------------------
// host side
template <typename real>
void fill_params(const lmc_mix_target& T, const lmc_mix_step& S, MixP<real>& P) {
  memset(&P, 0, sizeof(P));
  P.n = T.n;
  P.t = S.t;
  for (int i = 0; i < T.n; ++i) {
    P.mu0[i] = (real)T.mu[i][0];
    P.mu1[i] = (real)T.mu[i][1];
    P.p00[i] = (real)T.prec[i][0]; P.p01[i] = (real)T.prec[i][1];
    P.p10[i] = (real)T.prec[i][2]; P.p11[i] = (real)T.prec[i][3];
    P.al[i] = (real)T.alpha[i];
    P.thr[i] = (real)(T.lamda * T.alpha[i]);
    P.c1[i] = (real)(-1.4426950408889634 * T.alpha[i]);
    double cst = (T.kind == LMC_TARGET_LAPLACE) ? T.omega[i] * (T.alpha[i] / 2) * (T.alpha[i] / 2)
                                                : T.omega[i] / T.norm[i];
    P.lc[i] = (real)log2(cst);
  }
  P.lamda = (real)T.lamda;
  P.c2 = (real)(T.lamda != 0 ? -1.4426950408889634 / (2 * T.lamda) : 0.0);
  P.inv_lamda = (real)(T.lamda != 0 ? 1.0 / T.lamda : 0.0);
  P.pr_alpha = (real)T.prior_alpha;
  P.pr_thr = (real)(T.lamda * T.prior_alpha);
  P.pr_mu0 = (real)T.prior_mu[0];
  P.pr_mu1 = (real)T.prior_mu[1];
  P.gamma = (real)S.gamma;
  P.s2g = (real)sqrt(2 * S.gamma);
  P.inv4g = (real)(1.0 / (4 * S.gamma));
  P.inv4g_l2e = (real)(1.4426950408889634 / (4 * S.gamma));
  P.ngs = (real)(T.kind == LMC_TARGET_LAPLACE ? (T.lamda != 0 ? -S.gamma / T.lamda : 0.0) : -S.gamma);
  P.m00 = (real)S.M[0]; P.m01 = (real)S.M[1]; P.m10 = (real)S.M[2]; P.m11 = (real)S.M[3];
  P.s00 = (real)S.sqrtM[0]; P.s01 = (real)S.sqrtM[1]; P.s10 = (real)S.sqrtM[2]; P.s11 = (real)S.sqrtM[3];
  {
    const double s2g = sqrt(2.0 * S.gamma);
    P.gm00 = (real)(-S.gamma * S.M[0]); P.gm01 = (real)(-S.gamma * S.M[1]);
    P.gm10 = (real)(-S.gamma * S.M[2]); P.gm11 = (real)(-S.gamma * S.M[3]);
    P.gs00 = (real)(s2g * S.sqrtM[0]); P.gs01 = (real)(s2g * S.sqrtM[1]);
    P.gs10 = (real)(s2g * S.sqrtM[2]); P.gs11 = (real)(s2g * S.sqrtM[3]);
  }
  P.q00 = (real)S.Q[0]; P.q01 = (real)S.Q[1]; P.q10 = (real)S.Q[2]; P.q11 = (real)S.Q[3];
  P.qi00 = (real)S.Qinv[0]; P.qi01 = (real)S.Qinv[1]; P.qi10 = (real)S.Qinv[2]; P.qi11 = (real)S.Qinv[3];
  P.eta = (real)S.eta;
  P.inv_eta = (real)(S.eta != 0 ? 1.0 / S.eta : 0.0);
  P.pp_thr = (real)(S.eta != 0 ? T.lamda / S.eta : 0.0);
  P.b0 = (real)S.beta[0]; P.b1 = (real)S.beta[1];
  P.sg0 = (real)S.sigma[0]; P.sg1 = (real)S.sigma[1];
  P.ib0 = (real)(S.beta[0] != 0 ? 1.0 / S.beta[0] : 0.0); P.ib1 = (real)(S.beta[1] != 0 ? 1.0 / S.beta[1] : 0.0);
  P.isg0 = (real)(S.sigma[0] != 0 ? 1.0 / S.sigma[0] : 0.0); P.isg1 = (real)(S.sigma[1] != 0 ? 1.0 / S.sigma[1] : 0.0);
  P.hshift = (real)S.hess_shift;
}

template <typename real, int TARGET, int ALG, int NC, bool PROD>
int launch_one(const MixP<real>& P, const MixIo& io, cudaStream_t st) {
  auto kern = mix_chain_kernel<real, TARGET, ALG, NC, PROD>;
  int dev = 0, sms = 0, occ = 0;
  LMC_CUDA_OK(cudaGetDevice(&dev));
  LMC_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  LMC_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, BLOCK, 0));
  if (occ < 1) occ = 1;
  long long want = (io.n_items + (BLOCK / 32) - 1) / (BLOCK / 32);
  long long cap = (long long)sms * occ;
  int grid = (int)(want < cap ? want : cap);
  if (grid < 1) grid = 1;
  kern<<<grid, BLOCK, 0, st>>>(P, io);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}

// production = nothing but registers inside the K-loop (Philox noise, no per-step outputs)
inline bool is_production(const MixIo& io) { return io.noise == nullptr && io.traj == nullptr && io.accept == nullptr; }

// LMC_MIX_SCALAR=1 keeps fp32 production chains on the one-chain-per-thread kernel (A/B runs of the packed path)
inline bool packed_enabled();
inline long long items_for(int dtype, const MixIo& io, long long C) {
  const bool packed = dtype == LMC_F32 && is_production(io) && packed_enabled();
  return packed ? (C + 63) / 64 : (C + 31) / 32;
}
inline bool packed_enabled() {
  const char* e = getenv("LMC_MIX_SCALAR");
  return !(e != nullptr && e[0] == '1');
}

template <typename real, int TARGET, int ALG>
int launch_nc(const lmc_mix_target& T, const lmc_mix_step& S, MixIo io, cudaStream_t st) {
  const bool prod = is_production(io);
  if constexpr (sizeof(real) == 4) {
    if (prod && packed_enabled()) {   // fp32 production: two chains per thread, packed FP32 instructions
      MixP<f2> P2;
      fill_params(T, S, P2);
      // io.n_items was sized for 64 chains per warp by lmc_mix_run (items_for)
      return P2.n == 5 ? launch_one<f2, TARGET, ALG, 5, true>(P2, io, st) : launch_one<f2, TARGET, ALG, 0, true>(P2, io, st);
    }
  }
  MixP<real> P;
  fill_params(T, S, P);
  if (P.n == 5) return prod ? launch_one<real, TARGET, ALG, 5, true>(P, io, st)
                            : launch_one<real, TARGET, ALG, 5, false>(P, io, st);
  return prod ? launch_one<real, TARGET, ALG, 0, true>(P, io, st) : launch_one<real, TARGET, ALG, 0, false>(P, io, st);
}

// The 16 sampler x target instantiations (x {f2, float, double} x {production, validation}) are compiled in three
// translation units, one per target kind (mix_gauss.cu, mix_laplace.cu, mix_prox.cu define LMC_MIX_TU and include
// this file), so that nvcc builds them in parallel; this file alone (LMC_MIX_TU 0) is the API + dispatch.
#ifndef LMC_MIX_TU
#define LMC_MIX_TU 0
#endif
}  // namespace
int lmc_mix_dispatch_gauss(const lmc_mix_target* T, const lmc_mix_step* S, int dtype, const void* io, cudaStream_t st);
int lmc_mix_dispatch_laplace(const lmc_mix_target* T, const lmc_mix_step* S, int dtype, const void* io, cudaStream_t st);
int lmc_mix_dispatch_prox(const lmc_mix_target* T, const lmc_mix_step* S, int dtype, const void* io, cudaStream_t st);
namespace {

#define LMC_CASE(TG, AL) \
  if (S.alg == AL) return dtype == LMC_F64 ? launch_nc<double, TG, AL>(T, S, io, st) : launch_nc<float, TG, AL>(T, S, io, st);
#if LMC_MIX_TU == 1 || defined(LMC_MIX_BENCH_ONLY)
int dispatch_gauss(const lmc_mix_target& T, const lmc_mix_step& S, int dtype, const MixIo& io, cudaStream_t st) {
#ifndef LMC_MIX_BENCH_ONLY
  LMC_CASE(LMC_TARGET_GAUSS, LMC_ULA)
  LMC_CASE(LMC_TARGET_GAUSS, LMC_MALA)
  LMC_CASE(LMC_TARGET_GAUSS, LMC_PULA)
  LMC_CASE(LMC_TARGET_GAUSS, LMC_IHPULA)
  LMC_CASE(LMC_TARGET_GAUSS, LMC_MLA)
#endif
  lmc_set_error("lmc_mix_run: sampler %d is not defined for the Gaussian mixture", S.alg);
  return -1;
}
#endif
#if LMC_MIX_TU == 2 || defined(LMC_MIX_BENCH_ONLY)
int dispatch_laplace(const lmc_mix_target& T, const lmc_mix_step& S, int dtype, const MixIo& io, cudaStream_t st) {
  LMC_CASE(LMC_TARGET_LAPLACE, LMC_ULA)
  LMC_CASE(LMC_TARGET_LAPLACE, LMC_MALA)
  LMC_CASE(LMC_TARGET_LAPLACE, LMC_PULA)
  LMC_CASE(LMC_TARGET_LAPLACE, LMC_MLA)
#ifndef LMC_MIX_BENCH_ONLY
  LMC_CASE(LMC_TARGET_LAPLACE, LMC_IHPULA)
#endif
  lmc_set_error("lmc_mix_run: sampler %d is not defined for the Laplacian mixture", S.alg);
  return -1;
}
#endif
#if LMC_MIX_TU == 3 || defined(LMC_MIX_BENCH_ONLY)
int dispatch_prox(const lmc_mix_target& T, const lmc_mix_step& S, int dtype, const MixIo& io, cudaStream_t st) {
#ifndef LMC_MIX_BENCH_ONLY
  LMC_CASE(LMC_TARGET_GAUSS_L1, LMC_PGLD)
  LMC_CASE(LMC_TARGET_GAUSS_L1, LMC_MYULA)
  LMC_CASE(LMC_TARGET_GAUSS_L1, LMC_MYMALA)
  LMC_CASE(LMC_TARGET_GAUSS_L1, LMC_PPULA)
  LMC_CASE(LMC_TARGET_GAUSS_L1, LMC_FBULA)
  LMC_CASE(LMC_TARGET_GAUSS_L1, LMC_LBMUMLA)
#endif
  lmc_set_error("lmc_mix_run: sampler %d is not defined for the Gaussian mixture x Laplacian prior", S.alg);
  return -1;
}
#endif
#undef LMC_CASE

}  // namespace

#if LMC_MIX_TU == 1 || defined(LMC_MIX_BENCH_ONLY)
int lmc_mix_dispatch_gauss(const lmc_mix_target* T, const lmc_mix_step* S, int dtype, const void* io, cudaStream_t st) {
  return dispatch_gauss(*T, *S, dtype, *static_cast<const MixIo*>(io), st);
}
#endif
#if LMC_MIX_TU == 2 || defined(LMC_MIX_BENCH_ONLY)
int lmc_mix_dispatch_laplace(const lmc_mix_target* T, const lmc_mix_step* S, int dtype, const void* io, cudaStream_t st) {
  return dispatch_laplace(*T, *S, dtype, *static_cast<const MixIo*>(io), st);
}
#endif
#if LMC_MIX_TU == 3 || defined(LMC_MIX_BENCH_ONLY)
int lmc_mix_dispatch_prox(const lmc_mix_target* T, const lmc_mix_step* S, int dtype, const void* io, cudaStream_t st) {
  return dispatch_prox(*T, *S, dtype, *static_cast<const MixIo*>(io), st);
}
#endif

#if LMC_MIX_TU == 0
namespace {
inline uint2 make_key(uint64_t seed) { return make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)); }

}  // namespace

extern "C" size_t lmc_mix_workspace_bytes(int64_t C) {
  size_t items = (size_t)((C + 31) / 32);
  return 256 + items * 6 * sizeof(double);
}

extern "C" int lmc_mix_run(const lmc_mix_target* target, const lmc_mix_step* step, int dtype, int64_t C,
                           int64_t K, uint64_t seed, uint64_t chain_offset, const lmc_mix_io* uio,
                           void* stream) {
  LMC_REQUIRE(target && step && uio, "lmc_mix_run: null descriptor");
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_mix_run: dtype must be LMC_F32 or LMC_F64");
  LMC_REQUIRE(target->n >= 1 && target->n <= LMC_MAX_COMP, "lmc_mix_run: n=%d outside 1..%d", target->n,
              LMC_MAX_COMP);
  LMC_REQUIRE(C >= 0 && K >= 0, "lmc_mix_run: negative C or K");
  LMC_REQUIRE(C <= ((int64_t)1 << 36) - 64, "lmc_mix_run: C too large (work items are 32-bit)");
  LMC_REQUIRE(step->gamma > 0, "lmc_mix_run: gamma must be positive");
  if (C == 0) return 0;
  LMC_REQUIRE(uio->d_theta != nullptr, "lmc_mix_run: d_theta is required");
  LMC_REQUIRE(uio->d_workspace != nullptr && uio->workspace_bytes >= lmc_mix_workspace_bytes(C),
              "lmc_mix_run: workspace too small (%zu < %zu)", uio->workspace_bytes,
              lmc_mix_workspace_bytes(C));
  const bool adj = step->alg == LMC_MALA || step->alg == LMC_MYMALA;
  LMC_REQUIRE(!(uio->d_noise != nullptr && adj && uio->d_uniforms == nullptr),
              "lmc_mix_run: adjusted samplers in validation mode need d_uniforms");
  LMC_REQUIRE(!(uio->d_noise != nullptr && uio->init_from_rng),
              "lmc_mix_run: validation mode takes the start state from d_theta");
  LMC_REQUIRE(!(step->alg == LMC_PPULA && (step->t < 1 || !(step->eta > 0))),
              "lmc_mix_run: PPULA needs t >= 1 and eta > 0");
  LMC_REQUIRE(!(uio->d_hist != nullptr && (uio->hist_bins < 1 || !(uio->hist_hi > uio->hist_lo))),
              "lmc_mix_run: bad histogram range");
  cudaStream_t st = (cudaStream_t)stream;

  MixIo io;
  memset(&io, 0, sizeof(io));
  io.theta = uio->d_theta;
  io.noise = uio->d_noise;
  io.unif = uio->d_uniforms;
  io.traj = uio->d_traj;
  io.accept = uio->d_accept;
  io.naccept = uio->d_naccept;
  io.hist = uio->d_hist;
  io.queue = (unsigned int*)uio->d_workspace;
  io.partial = (double*)((char*)uio->d_workspace + 256);
  io.C = C;
  io.K = K;
  io.burnin = uio->burnin;
  io.hist_every = uio->hist_every;
  io.chain_offset = chain_offset;
  io.key = make_key(seed);
  lmc_philox_round_keys(io.key, io.rk);
  io.init_from_rng = uio->init_from_rng;
  io.hist_bins = uio->hist_bins;
  io.hist_lo = uio->hist_lo;
  io.hist_scale = uio->d_hist ? uio->hist_bins / (uio->hist_hi - uio->hist_lo) : 0.0;
  io.n_items = (int)items_for(dtype, io, C);

  LMC_CUDA_OK(cudaMemsetAsync(io.queue, 0, 256, st));
  int rc = -1;
  switch (target->kind) {
    case LMC_TARGET_GAUSS: rc = lmc_mix_dispatch_gauss(target, step, dtype, &io, st); break;
    case LMC_TARGET_LAPLACE: rc = lmc_mix_dispatch_laplace(target, step, dtype, &io, st); break;
    case LMC_TARGET_GAUSS_L1: rc = lmc_mix_dispatch_prox(target, step, dtype, &io, st); break;
    default: lmc_set_error("lmc_mix_run: unknown target kind %d", target->kind);
  }
  if (rc != 0) return rc;
  if (uio->d_moments != nullptr) {
    mix_reduce_moments<<<1, 256, 0, st>>>(io.partial, io.n_items, uio->d_moments);
    LMC_CUDA_OK(cudaGetLastError());
  }
  return 0;
}

extern "C" int lmc_mix_philox_draws(int dtype, int adjusted, int64_t C, int64_t K, uint64_t seed,
                                    uint64_t chain_offset, void* d_normals, void* d_uniforms, void* d_theta0,
                                    void* stream) {
  LMC_REQUIRE(d_normals != nullptr, "lmc_mix_philox_draws: d_normals is required");
  if (C <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  int grid = (int)((C + 127) / 128);
  if (dtype == LMC_F64)
    philox_draws_kernel<double><<<grid, 128, 0, st>>>(adjusted, C, K, make_key(seed), chain_offset,
                                                      (double*)d_normals, (double*)d_uniforms, (double*)d_theta0);
  else
    philox_draws_kernel<float><<<grid, 128, 0, st>>>(adjusted, C, K, make_key(seed), chain_offset,
                                                     (float*)d_normals, (float*)d_uniforms, (float*)d_theta0);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}
#endif  // LMC_MIX_TU == 0
