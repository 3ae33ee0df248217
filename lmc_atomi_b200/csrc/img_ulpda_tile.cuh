// Device side of the ULPDA kernels shared by the tile kernel (img_ulpda.cu) and the per-tap translation units of the
// band kernel (img_ulpda_band{5,6,7}.cu): argument block, LSQR state machine, scalar phases, tile phases and the
// in-kernel non-convex pre-step.  See img_ulpda.cu for the phase plan.
#pragma once
#include "img_common.cuh"
#include "img_myula_tile.cuh"   // the TV pipeline (ME-TV prelude of a non-convex data term) + MC-TV stencil

using namespace lmcimg;

namespace {

struct LsqrState {
  double alfa, beta, rhobar, phibar, anorm, ddnorm, xxnorm, z, cs2, sn2, bnorm, res2;
  double su, sv, sw;   // pending scales of the stored u, v, w
  double t1, t2;       // deferred x / w update coefficients
  double wnorm2;       // ||w||^2 of the (scaled) stored w
  int itn, istop, active, pending;
};

template <typename real>
struct UlpdaArgs {
  Geom<real> g;
  real* x;
  real* xhat;
  real* y;       // [C][2][P]
  real* warm;    // [C][P]  LSQR x / warm start
  const real* b;
  long long b_stride;
  const real* zlin;   // optional linear term (P,)
  real* optb;    // [Cb][P] sigma H^T b
  real* u;
  real* v;
  real* w;
  const real* noise;
  real* samples;
  real* ysamples;
  real* mean;
  real* m2;
  const real* in;   // l2-prox mode: argument; ULPDA mode: optional override of x - tau (G^T y + z)
  real* out;        // l2-prox mode: result
  double* partial;  // [3][C][tiles]
  LsqrState* st;
  int* n_active;
  unsigned long long* counters;
  long long C, N, burnin, stat_count0;
  unsigned long long chain_offset, iter_offset;
  uint2 key;
  real sigma_l2, tau, mu, theta, s2t, g_sigma, cA;  // cA = tau * sigma_l2
  int isotropic, gfirst, iter_lim, warm_on;
  double atol, btol, ctol;
  int mode;  // 0 = ULPDA, 1 = L2 prox only
  // non-convex data term evaluated in-kernel (algs.L2_ncvx_tv.prox, algs.py:201-223): every iteration
  // arg_w = x - tau (G^T y + z); `in` = arg_w + nc_scale * grad-like term(arg_w)
  NcvxK<real> nc;     // kind 1: MC-TV stencil on arg_w; kind 2: ME-TV (TV pipeline of the second kernel argument)
  real nc_scale;      // MC-TV: tau lamda (isotropic) or tau lamda / gamma (anisotropic)
  real* arg_w;        // (C,P) workspace
  real* arg2_w;       // (C,P) workspace = `in` when nc.kind != 0
  // band kernel only (img_ulpda_band.inl)
  int pitch;        // shared-memory row pitch in elements
  int* queue;       // 4 work-queue counters (BandSched)
  long long group;  // chains advanced together (working set kept near the L2 size)
};

__device__ __forceinline__ void sym_ortho(double a, double b, double& c, double& s, double& r) {
  // scipy.sparse.linalg._isolve.lsqr._sym_ortho
  if (b == 0) { c = (a > 0) - (a < 0); s = 0; r = fabs(a); }
  else if (a == 0) { c = 0; s = (b > 0) - (b < 0); r = fabs(b); }
  else if (fabs(b) > fabs(a)) {
    double tau = a / b;
    s = ((b > 0) - (b < 0)) / sqrt(1 + tau * tau);
    c = s * tau;
    r = b / s;
  } else {
    double tau = b / a;
    c = ((a > 0) - (a < 0)) / sqrt(1 + tau * tau);
    s = c * tau;
    r = a / c;
  }
}

// stage `scale * src` on the tile extended by e (zero outside the image)
template <typename real>
__device__ __forceinline__ void load_region_scaled(real* S, const real* img, const Geom<real>& g, int i0, int j0,
                                                   real scale) {
  const int e = g.e, rows = TH + 2 * e, cols = TW + 2 * e;
  for (int idx = threadIdx.x; idx < rows * cols; idx += NTHREADS) {
    int a = idx / cols, bcol = idx - a * cols, i = i0 - e + a, j = j0 - e + bcol;
    real val = real(0);
    if (i >= 0 && i < g.ny && j >= 0 && j < g.nx) val = scale * img[(long long)i * g.nx + j];
    S[a * RS + bcol] = val;
  }
}

template <typename real>
__device__ __forceinline__ void tile_of(const Geom<real>& g, long long w, long long& chain, int& i0, int& j0) {
  chain = w / g.tiles;
  int tile = (int)(w - chain * g.tiles);
  i0 = (tile / g.tiles_x) * TH;
  j0 = (tile % g.tiles_x) * TW;
}

// ---------------------------------------------------------------- P0: optb = sigma H^T b
template <typename real>
__device__ void phase_optb(const UlpdaArgs<real>& A, real* smem) {
  const Geom<real>& g = A.g;
  real *X = smem, *T1 = smem + RS * RS, *T2 = smem + 2 * RS * RS;
  const long long Cb = A.b_stride ? A.C : 1, items = Cb * g.tiles;
  const int e = g.e, hy = g.lo_y + g.ntap_y - 1;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain; int i0, j0;
    tile_of(g, w, chain, i0, j0);
    __syncthreads();
    load_region_scaled(X, A.b + chain * A.b_stride, g, i0, j0, real(1));
    __syncthreads();
    blur_pass<real, 0, -1>(X, T1, RS, e - hy, e + TH - g.lo_y, e, e + TW, g.wx, g.ntap_x, g.lo_x);
    __syncthreads();
    blur_pass<real, -1, 0>(T1, T2, RS, e, e + TH, e, e + TW, g.wy, g.ntap_y, g.lo_y);
    __syncthreads();
    for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
      int a = idx / TW, bcol = idx % TW, i = i0 + a, j = j0 + bcol;
      if (i < g.ny && j < g.nx) A.optb[chain * g.P + (long long)i * g.nx + j] = A.sigma_l2 * T2[(e + a) * RS + e + bcol];
    }
  }
}

// ---------------------------------------------------------------- P1
template <typename real>
__device__ void phase_primal_init(const UlpdaArgs<real>& A, real* smem, double* red) {
  const Geom<real>& g = A.g;
  real *X = smem, *T1 = smem + RS * RS, *T2 = smem + 2 * RS * RS;
  const long long items = A.C * g.tiles;
  const int e = g.e;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain; int i0, j0;
    tile_of(g, w, chain, i0, j0);
    real* zw = A.warm + chain * g.P;
    __syncthreads();
    if (A.warm_on) load_region_scaled(X, zw, g, i0, j0, real(1));
    else for (int idx = threadIdx.x; idx < RS * RS; idx += NTHREADS) X[idx] = real(0);
    __syncthreads();
    normal_stencil<real>(g, X, T1, T2, nullptr, i0, j0);
    const real* y0 = A.y + chain * 2 * g.P;
    const real* y1 = y0 + g.P;
    const real* ob = A.optb + (A.b_stride ? chain : 0) * g.P;
    double accb = 0, accu = 0;
    for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
      int a = idx / TW, bcol = idx % TW, i = i0 + a, j = j0 + bcol;
      if (i < g.ny && j < g.nx) {
        long long o = (long long)i * g.nx + j;
        real arg;
        if (A.in != nullptr) {   // L2-prox mode, or the non-convex data term's pre-processed argument
          arg = A.in[chain * g.P + o];
        } else {
          // G^T y in the accumulation order of pylops (axis 0 then axis 1); v[last] is never read
          real t = real(0);
          if (i < g.ny - 1) t -= y0[o];
          if (i >= 1) t += y0[o - g.nx];
          if (j < g.nx - 1) t -= y1[o];
          if (j >= 1) t += y1[o - 1];
          if (A.zlin != nullptr) t += A.zlin[o];
          arg = A.x[chain * g.P + o] - A.tau * t;   // algs.py:439/446: x - tau * ATy
        }
        real rhs = arg + A.tau * ob[o];             // pyproximal L2.prox: y = x + tau * OpTb
        real zc = X[(e + a) * RS + e + bcol];
        real u = rhs - (zc + A.cA * T2[(e + a) * RS + e + bcol]);   // u = b - A x0
        A.u[chain * g.P + o] = u;
        if (!A.warm_on) zw[o] = real(0);
        accb += (double)rhs * (double)rhs;
        accu += (double)u * (double)u;
      }
    }
    double sb = block_sum(accb, red);
    double su = block_sum(accu, red);
    if (threadIdx.x == 0) {
      A.partial[w] = sb;
      A.partial[items + w] = su;
    }
  }
}

template <typename real>
__device__ __forceinline__ double chain_partial(const UlpdaArgs<real>& A, int q, long long chain, int lane) {
  double s = 0;
  const double* p = A.partial + (long long)q * A.C * A.g.tiles + chain * A.g.tiles;
  for (int t = lane; t < A.g.tiles; t += 32) s += p[t];
  return lmc_warp_sum(s);
}

// one warp per chain of [c_lo, c_hi) (c_lo, c_hi: arguments of the enclosing scalar phase)
#define LMC_FOR_CHAINS_BY_WARP(chain)                                                   \
  const int lane = threadIdx.x & 31;                                                    \
  const long long warp_ = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;      \
  const long long nwarps_ = ((long long)gridDim.x * blockDim.x) >> 5;                   \
  for (long long chain = c_lo + warp_; chain < c_hi; chain += nwarps_)

// ---------------------------------------------------------------- S1
template <typename real>
__device__ void phase_s1(const UlpdaArgs<real>& A, long long c_lo, long long c_hi) {
  LMC_FOR_CHAINS_BY_WARP(chain) {
    double sb = chain_partial(A, 0, chain, lane), su = chain_partial(A, 1, chain, lane);
    if (lane == 0) {
      LsqrState s;
      memset(&s, 0, sizeof(s));
      s.bnorm = sqrt(sb);
      s.beta = sqrt(su);
      s.su = s.beta > 0 ? 1.0 / s.beta : 0.0;
      s.active = s.beta > 0 ? 1 : 0;   // beta == 0: alfa = 0 -> arnorm = 0 -> x = x0 returned
      s.cs2 = -1.0;
      A.st[chain] = s;
    }
  }
}

// ---------------------------------------------------------------- P2: v = Op1 (su * u); w = v
template <typename real>
__device__ void phase_v0(const UlpdaArgs<real>& A, real* smem, double* red) {
  const Geom<real>& g = A.g;
  real *X = smem, *T1 = smem + RS * RS, *T2 = smem + 2 * RS * RS;
  const long long items = A.C * g.tiles;
  const int e = g.e;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain; int i0, j0;
    tile_of(g, w, chain, i0, j0);
    if (!A.st[chain].active) continue;
    const real su = (real)A.st[chain].su;
    __syncthreads();
    load_region_scaled(X, A.u + chain * g.P, g, i0, j0, su);
    __syncthreads();
    normal_stencil<real>(g, X, T1, T2, nullptr, i0, j0);
    double acc = 0;
    for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
      int a = idx / TW, bcol = idx % TW, i = i0 + a, j = j0 + bcol;
      if (i < g.ny && j < g.nx) {
        long long o = chain * g.P + (long long)i * g.nx + j;
        real vv = X[(e + a) * RS + e + bcol] + A.cA * T2[(e + a) * RS + e + bcol];
        A.v[o] = vv;
        A.w[o] = vv;
        acc += (double)vv * (double)vv;
      }
    }
    double s = block_sum(acc, red);
    if (threadIdx.x == 0) A.partial[items + w] = s;
  }
}

// ---------------------------------------------------------------- S2
template <typename real>
__device__ void phase_s2(const UlpdaArgs<real>& A, long long c_lo, long long c_hi) {
  LMC_FOR_CHAINS_BY_WARP(chain) {
    LsqrState s = A.st[chain];
    if (!s.active) continue;
    double sv = chain_partial(A, 1, chain, lane);
    if (lane == 0) {
      s.alfa = sqrt(sv);
      s.sv = s.alfa > 0 ? 1.0 / s.alfa : 1.0;
      s.sw = s.sv;
      s.wnorm2 = sv * s.sv * s.sv;
      s.rhobar = s.alfa;
      s.phibar = s.beta;
      if (s.alfa * s.beta == 0) s.active = 0;   // arnorm == 0
      if (A.iter_lim <= 0) s.active = 0;
      if (s.active) atomicAdd(A.n_active, 1);
      A.st[chain] = s;
    }
  }
}

// ---------------------------------------------------------------- Ka
template <typename real>
__device__ void phase_ka(const UlpdaArgs<real>& A, real* smem, double* red) {
  const Geom<real>& g = A.g;
  real *X = smem, *T1 = smem + RS * RS, *T2 = smem + 2 * RS * RS;
  const long long items = A.C * g.tiles;
  const int e = g.e;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain; int i0, j0;
    tile_of(g, w, chain, i0, j0);
    const LsqrState s = A.st[chain];
    if (!s.active) continue;
    const real sv = (real)s.sv, sw = (real)s.sw, su = (real)s.su, t1 = (real)s.t1, t2 = (real)s.t2,
               alfa = (real)s.alfa;
    __syncthreads();
    load_region_scaled(X, A.v + chain * g.P, g, i0, j0, sv);
    __syncthreads();
    normal_stencil<real>(g, X, T1, T2, nullptr, i0, j0);
    double accu = 0, accw = 0;
    for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
      int a = idx / TW, bcol = idx % TW, i = i0 + a, j = j0 + bcol;
      if (i < g.ny && j < g.nx) {
        long long o = chain * g.P + (long long)i * g.nx + j;
        real vs = X[(e + a) * RS + e + bcol];
        if (s.pending) {   // deferred update of the previous LSQR iteration
          real wv = sw * A.w[o];
          A.warm[o] += t1 * wv;          // x = x + t1 * w
          wv = vs + t2 * wv;             // w = v + t2 * w
          A.w[o] = wv;
          accw += (double)wv * (double)wv;
        }
        real un = (vs + A.cA * T2[(e + a) * RS + e + bcol]) - alfa * (su * A.u[o]);   // u = A v - alfa u
        A.u[o] = un;
        accu += (double)un * (double)un;
      }
    }
    double s1 = block_sum(accu, red);
    double s2 = block_sum(accw, red);
    if (threadIdx.x == 0) {
      A.partial[items + w] = s1;
      A.partial[2 * items + w] = s2;
    }
  }
}

// ---------------------------------------------------------------- Sa
template <typename real>
__device__ void phase_sa(const UlpdaArgs<real>& A, long long c_lo, long long c_hi) {
  LMC_FOR_CHAINS_BY_WARP(chain) {
    LsqrState s = A.st[chain];
    if (!s.active) continue;
    double su = chain_partial(A, 1, chain, lane), sw = chain_partial(A, 2, chain, lane);
    if (lane == 0) {
      if (s.pending) { s.wnorm2 = sw; s.sw = 1.0; s.pending = 0; }
      s.beta = sqrt(su);
      if (s.beta > 0) {
        s.su = 1.0 / s.beta;
        s.anorm = sqrt(s.anorm * s.anorm + s.alfa * s.alfa + s.beta * s.beta);
      } else {
        s.su = 1.0;
      }
      A.st[chain] = s;
    }
  }
}

// ---------------------------------------------------------------- Kb
template <typename real>
__device__ void phase_kb(const UlpdaArgs<real>& A, real* smem, double* red) {
  const Geom<real>& g = A.g;
  real *X = smem, *T1 = smem + RS * RS, *T2 = smem + 2 * RS * RS;
  const long long items = A.C * g.tiles;
  const int e = g.e;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain; int i0, j0;
    tile_of(g, w, chain, i0, j0);
    const LsqrState s = A.st[chain];
    if (!s.active || !(s.beta > 0)) continue;
    const real su = (real)s.su, sv = (real)s.sv, beta = (real)s.beta;
    __syncthreads();
    load_region_scaled(X, A.u + chain * g.P, g, i0, j0, su);
    __syncthreads();
    normal_stencil<real>(g, X, T1, T2, nullptr, i0, j0);
    double acc = 0;
    for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
      int a = idx / TW, bcol = idx % TW, i = i0 + a, j = j0 + bcol;
      if (i < g.ny && j < g.nx) {
        long long o = chain * g.P + (long long)i * g.nx + j;
        real vn = (X[(e + a) * RS + e + bcol] + A.cA * T2[(e + a) * RS + e + bcol]) - beta * (sv * A.v[o]);
        A.v[o] = vn;
        acc += (double)vn * (double)vn;
      }
    }
    double s1 = block_sum(acc, red);
    if (threadIdx.x == 0) A.partial[items + w] = s1;
  }
}

// ---------------------------------------------------------------- Sb: scipy lsqr scalar block
template <typename real>
__device__ void phase_sb(const UlpdaArgs<real>& A, long long c_lo, long long c_hi) {
  const double eps = 2.220446049250313e-16;
  LMC_FOR_CHAINS_BY_WARP(chain) {
    LsqrState s = A.st[chain];
    if (!s.active) continue;
    double sv = (s.beta > 0) ? chain_partial(A, 1, chain, lane) : 0.0;
    if (lane == 0) {
      s.itn += 1;
      if (s.beta > 0) {
        s.alfa = sqrt(sv);
        s.sv = s.alfa > 0 ? 1.0 / s.alfa : 1.0;
      }
      // plane rotation eliminating beta (damp = 0: rhobar1 = rhobar, psi = 0)
      double cs, sn, rho;
      sym_ortho(s.rhobar, s.beta, cs, sn, rho);
      double theta = sn * s.alfa;
      s.rhobar = -cs * s.alfa;
      double phi = cs * s.phibar;
      s.phibar = sn * s.phibar;
      double tau = sn * phi;
      s.t1 = phi / rho;
      s.t2 = -theta / rho;
      s.ddnorm = s.ddnorm + s.wnorm2 / (rho * rho);   // ||(1/rho) w||^2
      s.pending = 1;
      // right rotation -> estimate of ||x||
      double delta = s.sn2 * rho, gambar = -s.cs2 * rho;
      double rhs = phi - delta * s.z, zbar = rhs / gambar;
      double xnorm = sqrt(s.xxnorm + zbar * zbar);
      double gamma = sqrt(gambar * gambar + theta * theta);
      s.cs2 = gambar / gamma;
      s.sn2 = theta / gamma;
      s.z = rhs / gamma;
      s.xxnorm = s.xxnorm + s.z * s.z;
      // convergence tests
      double acond = s.anorm * sqrt(s.ddnorm);
      double rnorm = sqrt(s.phibar * s.phibar + s.res2);
      double arnorm = s.alfa * fabs(tau);
      double test1 = rnorm / s.bnorm;
      double test2 = arnorm / (s.anorm * rnorm + eps);
      double test3 = 1 / (acond + eps);
      double tt1 = test1 / (1 + s.anorm * xnorm / s.bnorm);
      double rtol = A.btol + A.atol * s.anorm * xnorm / s.bnorm;
      int istop = 0;
      if (s.itn >= A.iter_lim) istop = 7;
      if (1 + test3 <= 1) istop = 6;
      if (1 + test2 <= 1) istop = 5;
      if (1 + tt1 <= 1) istop = 4;
      if (test3 <= A.ctol) istop = 3;
      if (test2 <= A.atol) istop = 2;
      if (test1 <= rtol) istop = 1;
      s.istop = istop;
      if (istop != 0) s.active = 0;
      if (s.active) atomicAdd(A.n_active, 1);
      atomicAdd(A.counters + 1, 1ull);
      A.st[chain] = s;
    }
  }
}

// ---------------------------------------------------------------- Pf
template <typename real>
__device__ void phase_primal_final(const UlpdaArgs<real>& A, long long n) {
  const Geom<real>& g = A.g;
  const long long items = A.C * g.tiles;
  const int quads_per_row = (g.nx + 3) / 4;
  const unsigned long long it_global = A.iter_offset + (unsigned long long)n;
  const bool stats = (A.mean != nullptr) && ((long long)it_global >= A.burnin);
  const real cntn = (real)(A.stat_count0 + ((long long)it_global - A.burnin) + 1);
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain; int i0, j0;
    tile_of(g, w, chain, i0, j0);
    const LsqrState s = A.st[chain];
    const real t1 = (real)s.t1, sw = (real)s.sw;
    for (int idx = threadIdx.x; idx < TH * (TW / 4); idx += NTHREADS) {
      int a = idx / (TW / 4), qd = idx % (TW / 4), i = i0 + a, jb = j0 + 4 * qd;
      if (i >= g.ny || jb >= g.nx) continue;
      real z[4] = {0, 0, 0, 0};
      if (A.mode == 0 && A.noise == nullptr)
        img_noise4<real>(A.key, A.chain_offset + (unsigned long long)chain, it_global, i, jb >> 2, quads_per_row, z);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        int j = jb + k;
        if (j >= g.nx) break;
        long long o = chain * g.P + (long long)i * g.nx + j;
        real zf = A.warm[o];
        if (s.pending) zf += t1 * (sw * A.w[o]);
        if (A.mode == 1) {
          A.out[o] = zf;
          A.warm[o] = zf;
          continue;
        }
        A.warm[o] = zf;
        real xi = (A.noise != nullptr) ? A.noise[(chain * A.N + n) * g.P + (long long)i * g.nx + j] : z[k];
        real xold = A.x[o];
        real xn = zf + A.s2t * xi;                     // algs.py:446
        A.x[o] = xn;
        A.xhat[o] = xn + A.theta * (xn - xold);        // algs.py:447
        if (A.samples != nullptr) A.samples[(chain * A.N + n) * g.P + (long long)i * g.nx + j] = xn;
        if (stats) {
          real m = A.mean[o];
          real d = xn - m;
          m += d / cntn;
          A.mean[o] = m;
          if (A.m2 != nullptr) A.m2[o] += d * (xn - m);
        }
      }
    }
  }
}

// ---------------------------------------------------------------- Pd
template <typename real>
__device__ void phase_dual(const UlpdaArgs<real>& A, long long n, real* smem) {
  const Geom<real>& g = A.g;
  constexpr int PT = TW + 1;
  real* XH = smem;  // (TH+1) x (TW+1)
  const long long items = A.C * g.tiles;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain; int i0, j0;
    tile_of(g, w, chain, i0, j0);
    __syncthreads();
    load_region(XH, PT, A.xhat + chain * g.P, g.ny, g.nx, i0, j0, TH + 1, TW + 1);
    __syncthreads();
    real* y0 = A.y + chain * 2 * g.P;
    real* y1 = y0 + g.P;
    for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
      int a = idx / TW, bcol = idx % TW, i = i0 + a, j = j0 + bcol;
      if (i < g.ny && j < g.nx) {
        long long o = (long long)i * g.nx + j;
        real xv = XH[a * PT + bcol];
        real d0 = (i < g.ny - 1) ? XH[(a + 1) * PT + bcol] - xv : real(0);
        real d1 = (j < g.nx - 1) ? XH[a * PT + bcol + 1] - xv : real(0);
        real v0 = y0[o] + A.mu * d0, v1 = y1[o] + A.mu * d1;   // y + mu * A xhat
        if (A.isotropic) {   // L21 dual prox: v / max(1, |v|_2 / sigma)
          real nrm = Math<real>::sqrt(v0 * v0 + v1 * v1);
          real den = Math<real>::max(real(1), nrm / A.g_sigma);
          v0 = v0 / den;
          v1 = v1 / den;
        } else {             // L1 dual prox: clip to [-sigma, sigma]
          v0 = Math<real>::max(-A.g_sigma, Math<real>::min(A.g_sigma, v0));
          v1 = Math<real>::max(-A.g_sigma, Math<real>::min(A.g_sigma, v1));
        }
        y0[o] = v0;
        y1[o] = v1;
        if (A.ysamples != nullptr) {
          real* ys = A.ysamples + (chain * A.N + n) * 2 * g.P;
          ys[o] = v0;
          ys[g.P + o] = v1;
        }
      }
    }
  }
}

// ---------------------------------------------------------------- non-convex data term (algs.py:201-223)
// arg_w = x - tau (G^T y + z) on chains [c_lo, c_hi): the argument ULPDA hands to proxf.prox (img_pd_arg_kernel)
template <typename real>
__device__ void phase_ncvx_arg(const UlpdaArgs<real>& A, long long c_lo, long long c_hi) {
  const Geom<real>& g = A.g;
  const long long total = (c_hi - c_lo) * g.P;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long cl = t / g.P, o = t - cl * g.P, chain = c_lo + cl;
    const int i = (int)(o / g.nx), j = (int)(o - (long long)i * g.nx);
    const real* y0 = A.y + chain * 2 * g.P;
    const real* y1 = y0 + g.P;
    real acc = real(0);
    if (i < g.ny - 1) acc -= y0[o];
    if (i >= 1) acc += y0[o - g.nx];
    if (j < g.nx - 1) acc -= y1[o];
    if (j >= 1) acc += y1[o - 1];
    if (A.zlin != nullptr) acc += A.zlin[o];
    A.arg_w[chain * g.P + o] = A.x[chain * g.P + o] - A.tau * acc;
  }
}

// MC-TV: arg2_w = arg_w + nc_scale * (G^T(w(G arg) G arg) / post_div)   (img_mctv_grad_kernel with base = 1)
template <typename real>
__device__ void phase_ncvx_mc(const UlpdaArgs<real>& A, long long c_lo, long long c_hi) {
  const Geom<real>& g = A.g;
  const long long total = (c_hi - c_lo) * g.P;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long cl = t / g.P, o = t - cl * g.P, chain = c_lo + cl;
    const int i = (int)(o / g.nx), j = (int)(o - (long long)i * g.nx);
    const real* x = A.arg_w + chain * g.P;
    const real acc = mctv_at(A.nc.W, i, j, g.ny, g.nx, [&](int di, int dj) { return x[o + di * g.nx + dj]; });
    const real m = A.nc_scale * (acc / A.nc.post_div);
    A.arg2_w[chain * g.P + o] = real(1) * x[o] + m;
  }
}

// the pre-step of L2_ncvx_tv.prox for chains [c_lo, c_hi) (ME-TV: all chains; B is the TV pipeline on arg_w)
template <typename real>
__device__ __forceinline__ void ncvx_prestep(const UlpdaArgs<real>& A, const MyulaArgs<real>& B, long long c_lo,
                                             long long c_hi, long long n, real* smem, double* red,
                                             cg::grid_group& grid) {
  phase_ncvx_arg(A, c_lo, c_hi);
  grid.sync();
  if (A.nc.kind == 1) {
    phase_ncvx_mc(A, c_lo, c_hi);
    grid.sync();
  } else {
    myula_tile_iteration(B, n, smem, red, grid);   // arg2_w = arg_w + tau lamda / gamma (arg_w - TV_1.prox(arg_w, gamma))
  }
}

}  // namespace
