// Tile (32 x 32) phases of the MYULA / TV-prox state machine, shared by img_myula.cu (the sampler) and img_ulpda.cu
// (which runs the same TV pipeline as the ME-TV prelude of a non-convex data term).  See img_myula.cu for the design.
#pragma once
#include "img_common.cuh"
#include "img_ncvx.cuh"

using namespace lmcimg;

namespace {

struct TvState {
  double prev_obj;
  int active;   // still iterating
  int kf;       // iteration whose `sol` is the result
};

template <typename real>
struct MyulaArgs {
  Geom<real> g;
  real* x;
  const real* b;
  long long b_stride;
  real* grad;
  const real* gextra;  // optional (C,P): grad f += gscale * gextra (non-convex data term)
  real gscale;
  real* D;            // [3][2][C][P]
  const real* noise;  // [C][N][P] or null
  real* samples;      // [C][N][P] or null
  real* mean;
  real* m2;
  real* out;          // tv-prox mode: result
  double* partial;    // [C][tiles]
  TvState* st;
  int* n_active;
  unsigned long long* counters;  // [0] executed TV sweeps (all chains, all iterations)
  long long C, N, burnin, stat_count0;
  unsigned long long chain_offset, iter_offset;
  uint2 key;
  real sigma_l2, tau, a_x, tog, s2t, tv_gamma;
  int tv_passes;     // executed passes of the TV loop at most: niter + LMC_TV_EXTRA_PASS (`while iter <= niter`)
  double tv_rtol;
  double tv_objw;    // weight of sum |grad sol| in the objective of the exit test: gamma * sigma_TV (see lmc_b200.h)
  int mode;  // 0 = MYULA, 1 = TV prox only (out = prox_{tv_gamma TV}(x)), 2 = ME-TV prelude of MYULA (out = (x - prox)
             // / me_den), 3 = ME-TV prelude of ULPDA (out = x + me_scale (x - prox), the in-place pre-step of
             // L2_ncvx_tv.prox, algs.py:221-223), -1 = disabled (second argument block of a launch without ME-TV)
  real me_den;        // mode 2: gamma of the Moreau envelope (algs.py:280-282)
  real me_scale;      // mode 3: tau * lamda / gamma
  NcvxK<real> nc;     // mode 0: in-kernel MC-TV term (kind 1): gextra_w <- G^T(w(Gx) Gx) / post_div in phase G
  real* gextra_w;     // (C,P) workspace the in-kernel non-convex terms write (read back through `gextra`)
  // band kernel only (img_myula_band.inl)
  int pitch;        // shared-memory row pitch in elements
  int* queue;       // 4 work-queue counters (BandSched)
};

__device__ __forceinline__ double fista_c(int it) {
  // t_k = (1 + sqrt(4 t_{k-1}^2)) / 2 = t_{k-1} + 1/2 (sic), t_0 = 1;
  // c_it = (t_{it-1} - 1) / t_it multiplies (p_it - p_{it-1}) in rr_it.
  if (it < 1) return 0.0;
  double told = 1.0 + 0.5 * (it - 1), t = (1.0 + sqrt(4.0 * told * told)) / 2.0;
  return (told - 1.0) / t;
}

template <typename real>
__device__ __forceinline__ real* dual_ptr(const MyulaArgs<real>& A, int buf, int comp, long long chain) {
  return A.D + (((long long)(buf * 2 + comp) * A.C + chain) * A.g.P);
}

// projected dual update of one pixel: r = rr - dx/(8 gamma), p = r / max(1, |(r,s)|)
template <typename real>
__device__ __forceinline__ void dual_update(real rr, real ss, real dx, real dy, real g8, real& p, real& q) {
  real r = rr - dx / g8, s = ss - dy / g8;
  real w = Math<real>::max(real(1), Math<real>::sqrt(r * r + s * s));
  p = r / w;
  q = s / w;
}

// ---------------------------------------------------------------- phase G (+ sweep 0)
template <typename real>
__device__ void phase_grad_sweep0(const MyulaArgs<real>& A, real* smem, double* red) {
  const Geom<real>& g = A.g;
  real* X = smem;
  real* T1 = smem + RS * RS;
  real* T2 = smem + 2 * RS * RS;
  const int e = g.e > 1 ? g.e : 1;   // sweep 0 needs a forward halo of 1
  const long long items = A.C * g.tiles;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    const long long chain = w / g.tiles;
    const int tile = (int)(w - chain * g.tiles);
    const int i0 = (tile / g.tiles_x) * TH, j0 = (tile % g.tiles_x) * TW;
    const real* xc = A.x + chain * g.P;
    __syncthreads();
    load_region(X, RS, xc, g.ny, g.nx, i0 - e, j0 - e, TH + 2 * e, TW + 2 * e);
    __syncthreads();
    if (A.mode == 0) {
      // normal_stencil assumes the region origin is (i0 - g.e, j0 - g.e); when e > g.e
      // (identity blur) shift the pointer so that local (0,0) is that origin.
      const real* Xo = X + (e - g.e) * RS + (e - g.e);
      normal_stencil(g, Xo, T1, T2, A.b + chain * A.b_stride, i0, j0);
      real* gout = A.grad + chain * g.P;
      for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
        int a = idx / TW, bcol = idx % TW, i = i0 + a, j = j0 + bcol;
        if (i < g.ny && j < g.nx) {
          gout[(long long)i * g.nx + j] = A.sigma_l2 * T2[(g.e + a) * RS + g.e + bcol];
          if (A.nc.kind == 1) {   // MC-TV: 7-point stencil out of the staged x (halo e >= 1)
            const real* xp = X + (e + a) * RS + e + bcol;
            const real m = mctv_at(A.nc.W, i, j, g.ny, g.nx, [&](int di, int dj) { return xp[di * RS + dj]; });
            A.gextra_w[chain * g.P + (long long)i * g.nx + j] = m / A.nc.post_div;
          }
        }
      }
    }
    // TV sweep 0: rr = ss = 0 -> sol = x, obj = gamma * sum |grad x|, p_1 = proj(-grad x / (8 gamma))
    real* p1 = dual_ptr(A, 1, 0, chain);
    real* q1 = dual_ptr(A, 1, 1, chain);
    const real g8 = real(8) * A.tv_gamma;
    double acc = 0;
    for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
      int a = idx / TW, bcol = idx % TW, i = i0 + a, j = j0 + bcol;
      if (i < g.ny && j < g.nx) {
        real xv = X[(e + a) * RS + e + bcol];
        real dx = (i < g.ny - 1) ? X[(e + a + 1) * RS + e + bcol] - xv : real(0);
        real dy = (j < g.nx - 1) ? X[(e + a) * RS + e + bcol + 1] - xv : real(0);
        acc += A.tv_objw * (double)Math<real>::sqrt(dx * dx + dy * dy);
        real p, q;
        dual_update(real(0), real(0), dx, dy, g8, p, q);
        p1[(long long)i * g.nx + j] = p;
        q1[(long long)i * g.nx + j] = q;
      }
    }
    double s = block_sum(acc, red);
    if (threadIdx.x == 0) A.partial[w] = s;
  }
}

// ---------------------------------------------------------------- phase S
template <typename real>
__device__ void phase_tv_scalar(const MyulaArgs<real>& A, int it) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * NTHREADS + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * NTHREADS) >> 5;
  for (long long chain = warp; chain < A.C; chain += nwarps) {
    TvState st = A.st[chain];
    if (it > 0 && !st.active) continue;
    double s = 0;
    for (int t = lane; t < A.g.tiles; t += 32) s += A.partial[chain * A.g.tiles + t];
    s = lmc_warp_sum(s);
    if (lane == 0) {
      double prev = it == 0 ? 0.0 : st.prev_obj;
      double rel = s > 1e-10 ? fabs(s - prev) / s : 2.0;
      st.prev_obj = s;
      st.kf = it;
      st.active = (rel < A.tv_rtol || it == A.tv_passes - 1) ? 0 : 1;
      A.st[chain] = st;
      if (st.active) atomicAdd(A.n_active, 1);
      atomicAdd(A.counters, 1ull);
    }
  }
}

// stage rr = p_it + c (p_it - p_{it-1}) for both components on rows [i0-1, i0+TH], cols [j0-1, j0+TW]
template <typename real>
__device__ __forceinline__ void stage_rr(const MyulaArgs<real>& A, long long chain, int it, int i0, int j0,
                                         real* RR, real* SS, int pitch) {
  const Geom<real>& g = A.g;
  const int rows = TH + 2, cols = TW + 2;
  if (it == 0) {
    for (int idx = threadIdx.x; idx < rows * cols; idx += NTHREADS) {
      RR[(idx / cols) * pitch + idx % cols] = real(0);
      SS[(idx / cols) * pitch + idx % cols] = real(0);
    }
    return;
  }
  const real c = (real)fista_c(it);
  const real* pa = dual_ptr(A, it % 3, 0, chain);
  const real* qa = dual_ptr(A, it % 3, 1, chain);
  const real* pb = dual_ptr(A, (it + 2) % 3, 0, chain);
  const real* qb = dual_ptr(A, (it + 2) % 3, 1, chain);
  for (int idx = threadIdx.x; idx < rows * cols; idx += NTHREADS) {
    int a = idx / cols, bcol = idx - a * cols, i = i0 - 1 + a, j = j0 - 1 + bcol;
    real r = real(0), s = real(0);
    if (i >= 0 && i < g.ny && j >= 0 && j < g.nx) {
      long long o = (long long)i * g.nx + j;
      real p = pa[o], q = qa[o];
      if (it >= 2) {
        r = p + c * (p - pb[o]);
        s = q + c * (q - qb[o]);
      } else {
        r = p;   // c_1 = 0 and p_0 = 0: p + 0 * (p - 0)
        s = q;
      }
    }
    RR[a * pitch + bcol] = r;
    SS[a * pitch + bcol] = s;
  }
}

// sol = x - gamma * div(rr, ss) at local (a, bcol) of the (TH+2)x(TW+2) staged duals
// (origin i0-1, j0-1); div = -grad^T, accumulated in the order of the oracle.
template <typename real>
__device__ __forceinline__ real sol_at(const real* RR, const real* SS, int pitch, int a, int bcol, real xv,
                                       real gamma) {
  real t = ((-RR[a * pitch + bcol] + RR[(a - 1) * pitch + bcol]) - SS[a * pitch + bcol]) + SS[a * pitch + bcol - 1];
  return xv - gamma * (-t);
}

// ---------------------------------------------------------------- phase T_k
template <typename real>
__device__ void phase_tv_sweep(const MyulaArgs<real>& A, int it, real* smem, double* red) {
  const Geom<real>& g = A.g;
  constexpr int PT = TW + 2;
  real* X = smem;                 // (TH+1) x (TW+1), pitch PT
  real* RR = X + (TH + 2) * PT;
  real* SS = RR + (TH + 2) * PT;
  real* SOL = SS + (TH + 2) * PT;  // (TH+1) x (TW+1)
  const long long items = A.C * g.tiles;
  const real g8 = real(8) * A.tv_gamma;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    const long long chain = w / g.tiles;
    if (!A.st[chain].active) continue;
    const int tile = (int)(w - chain * g.tiles);
    const int i0 = (tile / g.tiles_x) * TH, j0 = (tile % g.tiles_x) * TW;
    __syncthreads();
    load_region(X, PT, A.x + chain * g.P, g.ny, g.nx, i0, j0, TH + 1, TW + 1);
    stage_rr(A, chain, it, i0, j0, RR, SS, PT);
    __syncthreads();
    for (int idx = threadIdx.x; idx < (TH + 1) * (TW + 1); idx += NTHREADS) {
      int a = idx / (TW + 1), bcol = idx % (TW + 1);
      SOL[a * PT + bcol] = sol_at(RR, SS, PT, a + 1, bcol + 1, X[a * PT + bcol], A.tv_gamma);
    }
    __syncthreads();
    real* pn = dual_ptr(A, (it + 1) % 3, 0, chain);
    real* qn = dual_ptr(A, (it + 1) % 3, 1, chain);
    double acc = 0;
    for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
      int a = idx / TW, bcol = idx % TW, i = i0 + a, j = j0 + bcol;
      if (i < g.ny && j < g.nx) {
        real sv = SOL[a * PT + bcol];
        real dx = (i < g.ny - 1) ? SOL[(a + 1) * PT + bcol] - sv : real(0);
        real dy = (j < g.nx - 1) ? SOL[a * PT + bcol + 1] - sv : real(0);
        real df = X[a * PT + bcol] - sv;
        acc += 0.5 * (double)df * (double)df + A.tv_objw * (double)Math<real>::sqrt(dx * dx + dy * dy);
        real p, q;
        dual_update(RR[(a + 1) * PT + bcol + 1], SS[(a + 1) * PT + bcol + 1], dx, dy, g8, p, q);
        pn[(long long)i * g.nx + j] = p;
        qn[(long long)i * g.nx + j] = q;
      }
    }
    double s = block_sum(acc, red);
    if (threadIdx.x == 0) A.partial[w] = s;
  }
}

// ---------------------------------------------------------------- phase U
template <typename real>
__device__ void phase_update(const MyulaArgs<real>& A, long long n, real* smem) {
  const Geom<real>& g = A.g;
  constexpr int PT = TW + 2;
  real* RR = smem;
  real* SS = RR + (TH + 2) * PT;
  const long long items = A.C * g.tiles;
  const int quads_per_row = (g.nx + 3) / 4;
  const unsigned long long it_global = A.iter_offset + (unsigned long long)n;
  const bool stats = (A.mean != nullptr) && ((long long)it_global >= A.burnin);
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    const long long chain = w / g.tiles;
    const int tile = (int)(w - chain * g.tiles);
    const int i0 = (tile / g.tiles_x) * TH, j0 = (tile % g.tiles_x) * TW;
    const int kf = A.st[chain].kf;
    __syncthreads();
    stage_rr(A, chain, kf, i0, j0, RR, SS, PT);
    __syncthreads();
    real* xc = A.x + chain * g.P;
    // each thread: 4 consecutive pixels of one row (one Philox block of normals)
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
        long long o = (long long)i * g.nx + j;
        real xv = xc[o];
        real sol = sol_at(RR, SS, PT, a + 1, 4 * qd + k + 1, xv, A.tv_gamma);
        if (A.mode == 1) {
          A.out[chain * g.P + o] = sol;
          continue;
        }
        if (A.mode == 2) {   // ME-TV: grad of the Moreau envelope (x - prox) / gamma  (img_metv_combine_kernel)
          A.out[chain * g.P + o] = (xv - sol) / A.me_den;
          continue;
        }
        if (A.mode == 3) {   // ME-TV pre-step of the L2 prox: x + tau lamda / gamma (x - prox)  (img_metv_combine_kernel)
          A.out[chain * g.P + o] = xv + A.me_scale * (xv - sol);
          continue;
        }
        real xi = (A.noise != nullptr) ? A.noise[(chain * A.N + n) * g.P + o] : z[k];
        // algs.py:569: (1 - tau/gamma) x - tau grad f(x) + tau/gamma prox_g(x) + sqrt(2 tau) xi
        real gr = A.grad[chain * g.P + o];
        if (A.gextra != nullptr) gr = gr + A.gscale * A.gextra[chain * g.P + o];   // algs.py:291: g - lamda * grad_moreau
        real xn = A.a_x * xv - A.tau * gr + A.tog * sol + A.s2t * xi;
        xc[o] = xn;
        if (A.samples != nullptr) A.samples[(chain * A.N + n) * g.P + o] = xn;
        if (stats) {
          real cntn = (real)(A.stat_count0 + ((long long)it_global - A.burnin) + 1);
          real m = A.mean[chain * g.P + o];
          real d = xn - m;
          m += d / cntn;
          A.mean[chain * g.P + o] = m;
          if (A.m2 != nullptr) A.m2[chain * g.P + o] += d * (xn - m);
        }
      }
    }
  }
}

// one sampler iteration (mode 0) / one TV prox (modes 1, 2) of all chains
template <typename real>
__device__ __forceinline__ void myula_tile_iteration(const MyulaArgs<real>& A, long long n, real* smem, double* red,
                                                     cg::grid_group& grid) {
  if (blockIdx.x == 0 && threadIdx.x == 0) *A.n_active = 0;
  phase_grad_sweep0(A, smem, red);
  grid.sync();
  phase_tv_scalar(A, 0);
  grid.sync();
  for (int it = 1; it < A.tv_passes; ++it) {
    if (*(volatile int*)A.n_active == 0) break;
    grid.sync();   // everyone has read n_active before it is reset
    if (blockIdx.x == 0 && threadIdx.x == 0) *A.n_active = 0;
    phase_tv_sweep(A, it, smem, red);
    grid.sync();
    phase_tv_scalar(A, it);
    grid.sync();
  }
  phase_update(A, n, smem);
  grid.sync();
}


}  // namespace
