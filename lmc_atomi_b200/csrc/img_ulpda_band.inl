// ULPDA, row-band formulation (included by img_ulpda.cu; same arguments, LSQR state machine and
// scalar phases as the tile kernel above, different data movement).
//
// A work item is a band of RB full image rows of one chain.  The band plus its 2(NT-1) halo rows
// is staged once in shared memory (coalesced rows, zero pad columns = the zero boundary of
// Convolve2D); Op1 = I + tau sigma H^T H is then two fused separable passes with compile-time
// tap counts, each a horizontal filter out of shared memory feeding a vertical sliding window in
// registers (thread = image column), and the pointwise LSQR update of the phase runs in the
// epilogue of the second pass.  Chains are advanced in groups whose LSQR vectors (u, v, w, z)
// fit in the 126 MB L2, each group running all N sampler iterations before the next starts, so
// the inner LSQR loop is served from L2 instead of HBM.
namespace {

constexpr int BAND_PAD = 8;     // zero columns each side of a staged row (>= max taps - 1)
constexpr int BAND_BT = 256;    // threads per CTA

template <typename real, int NT, int RB>
struct Band {
  static constexpr int E = NT - 1;
  static constexpr int XROWS = RB + 2 * E;    // staged input rows
  static constexpr int TROWS = RB + E;        // intermediate (H x) rows
  __host__ __device__ static size_t smem_bytes(int pitch) { return (size_t)(XROWS + TROWS) * pitch * sizeof(real); }
};

__host__ __device__ inline int band_pitch(int nx) { return ((nx + 2 * BAND_PAD + 3) / 4) * 4; }

__device__ __forceinline__ double band_block_sum(double v, double* red) {
  v = lmc_warp_sum(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  double s = 0;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < BAND_BT / 32; ++i) s += red[i];
  }
  return s;
}

// rows [ifirst, ifirst + rows) of src -> S (zero outside the image), as 16-byte cp.async copies when the
// rows are 16-byte aligned.  Ends with the data visible to the whole CTA.
template <typename real>
__device__ __forceinline__ void band_stage(real* S, int pitch, const real* __restrict__ src, int ny, int nx, int ifirst,
                                           int rows) {
  constexpr int EPV = 16 / sizeof(real);
  const bool vec = (nx % EPV == 0) && ((reinterpret_cast<size_t>(src) & 15) == 0);
  for (int r = 0; r < rows; ++r) {
    const int i = ifirst + r;
    real* dst = S + r * pitch + BAND_PAD;
    if (i >= 0 && i < ny) {
      const real* s = src + (long long)i * nx;
      if (vec) {
        for (int c = threadIdx.x * EPV; c < nx; c += BAND_BT * EPV) {
          const unsigned d = (unsigned)__cvta_generic_to_shared(dst + c);
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(s + c) : "memory");
        }
      } else {
        for (int c = threadIdx.x; c < nx; c += BAND_BT) dst[c] = s[c];
      }
    } else {
      for (int c = threadIdx.x; c < nx; c += BAND_BT) dst[c] = real(0);
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
}

// second half of the operator: T (rows i0 - hi_y .. of a row-masked image) -> H^T T on the band.
// pre(i, c, v) loads the NPRE global operands of pixel (i, c) of the phase's pointwise update; they are
// fetched PFD rows ahead (all loads of a block in flight together), epi(i, c, k, value, v) consumes them
// (k = row within the band).  Both only for in-image pixels.
template <typename real, int NT, int RB, int NPRE, int PFD, class Pre, class Epi>
__device__ __forceinline__ void band_adjoint(const Geom<real>& g, const real* T, int pitch, int i0, const real (&wx)[NT],
                                             const real (&wy)[NT], Pre&& pre, Epi&& epi) {
  static_assert(RB % PFD == 0, "prefetch depth must divide the band height");
  constexpr int NBLK = RB / PFD;
  const int lox = g.lo_x;
  for (int c = threadIdx.x; c < g.nx; c += BAND_BT) {
    const real* trow = T + BAND_PAD + c - lox;
    real win[NT];
    real pf[2][PFD][NPRE > 0 ? NPRE : 1];   // two blocks of PFD rows in flight
    if (NPRE > 0) {
#pragma unroll
      for (int bl = 0; bl < 2 && bl < NBLK; ++bl)
#pragma unroll
        for (int kk = 0; kk < PFD; ++kk)
          if (i0 + bl * PFD + kk < g.ny) pre(i0 + bl * PFD + kk, c, pf[bl][kk]);
    }
#pragma unroll
    for (int jj = 0; jj < RB + NT - 1; ++jj) {
      real u = real(0);
#pragma unroll
      for (int s = 0; s < NT; ++s) u = Math<real>::fma(wx[s], trow[jj * pitch - s], u);
      win[jj % NT] = u;
      if (jj >= NT - 1) {
        const int k = jj - (NT - 1), i = i0 + k;
        real gv = real(0);
#pragma unroll
        for (int t = 0; t < NT; ++t) gv = Math<real>::fma(wy[t], win[(k + NT - 1 - t) % NT], gv);
        if (i < g.ny) epi(i, c, k, gv, pf[(k / PFD) & 1][k % PFD]);
        if (NPRE > 0 && (k % PFD) == PFD - 1 && k / PFD + 2 < NBLK) {   // block consumed: fetch block + 2 into its slot
#pragma unroll
          for (int kk = 0; kk < PFD; ++kk)
            if (i0 + k + 1 + PFD + kk < g.ny) pre(i0 + k + 1 + PFD + kk, c, pf[(k / PFD) & 1][kk]);
        }
      }
    }
  }
}

// first half: staged X (rows i0 - E ..) -> T = mask_rows(H X) on rows [i0 - hi_y, i0 + RB - lo_y)
template <typename real, int NT, int RB>
__device__ __forceinline__ void band_forward(const Geom<real>& g, const real* X, real* T, int pitch, int i0,
                                             const real (&wx)[NT], const real (&wy)[NT]) {
  const int lox = g.lo_x, hi_y = g.lo_y + NT - 1;
  for (int c = threadIdx.x; c < g.nx; c += BAND_BT) {
    const real* xrow = X + BAND_PAD + c + lox;
    real win[NT];
#pragma unroll
    for (int jj = 0; jj < RB + 2 * (NT - 1); ++jj) {
      real t1 = real(0);
#pragma unroll
      for (int s = 0; s < NT; ++s) t1 = Math<real>::fma(wx[s], xrow[jj * pitch + s], t1);
      win[jj % NT] = t1;
      if (jj >= NT - 1) {
        const int o = jj - (NT - 1), i = i0 - hi_y + o;
        real res = real(0);
        if (i >= 0 && i < g.ny) {
#pragma unroll
          for (int t = 0; t < NT; ++t) res = Math<real>::fma(wy[t], win[(o + t) % NT], res);
        }
        T[o * pitch + BAND_PAD + c] = res;
      }
    }
  }
}

template <typename real>
__device__ __forceinline__ void band_of(const Geom<real>& g, long long w, long long c_lo, int RB, long long& chain,
                                        int& i0) {
  const long long cl = w / g.tiles;
  chain = c_lo + cl;
  i0 = (int)(w - cl * g.tiles) * RB;
}

#define BAND_ITEMS_LOOP                                                        \
  const Geom<real>& g = A.g;                                                   \
  real* X = smem;                                                              \
  real* T = smem + Band<real, NT, RB>::XROWS * A.pitch;                        \
  const int pitch = A.pitch;                                                   \
  const long long items = (c_hi - c_lo) * g.tiles, pall = A.C * g.tiles;       \
  (void)X; (void)T; (void)pall;                                                \
  for (long long w = blockIdx.x; w < items; w += gridDim.x)

// ---------------------------------------------------------------- P0: optb = sigma H^T b
template <typename real, int NT, int RB>
__device__ void band_optb(const UlpdaArgs<real>& A, real* smem, const real (&wx)[NT], const real (&wy)[NT]) {
  const long long c_lo = 0, c_hi = A.b_stride ? A.C : 1;
  const int hi_y = A.g.lo_y + NT - 1;
  BAND_ITEMS_LOOP {
    long long chain; int i0;
    band_of(g, w, c_lo, RB, chain, i0);
    __syncthreads();
    band_stage(T, pitch, A.b + chain * A.b_stride, g.ny, g.nx, i0 - hi_y, Band<real, NT, RB>::TROWS);
    real* ob = A.optb + chain * g.P;
    band_adjoint<real, NT, RB, 0, 4>(g, T, pitch, i0, wx, wy, [&](int, int, real (&)[1]) {},
                                     [&](int i, int c, int, real gv, const real (&)[1]) {
                                       ob[(long long)i * g.nx + c] = A.sigma_l2 * gv;
                                     });
  }
}

// ---------------------------------------------------------------- P1: rhs, u = rhs - Op1 warm
template <typename real, int NT, int RB>
__device__ void band_primal_init(const UlpdaArgs<real>& A, long long c_lo, long long c_hi, real* smem, double* red,
                                 const real (&wx)[NT], const real (&wy)[NT]) {
  constexpr int E = NT - 1;
  BAND_ITEMS_LOOP {
    long long chain; int i0;
    band_of(g, w, c_lo, RB, chain, i0);
    real* zw = A.warm + chain * g.P;
    __syncthreads();
    if (A.warm_on) {
      band_stage(X, pitch, zw, g.ny, g.nx, i0 - E, Band<real, NT, RB>::XROWS);
    } else {
      for (int r = 0; r < Band<real, NT, RB>::XROWS; ++r)
        for (int c = threadIdx.x; c < g.nx; c += BAND_BT) X[r * pitch + BAND_PAD + c] = real(0);
      __syncthreads();
    }
    band_forward<real, NT, RB>(g, X, T, pitch, i0, wx, wy);
    __syncthreads();
    const real* y0 = A.y + chain * 2 * g.P;
    const real* y1 = y0 + g.P;
    const real* ob = A.optb + (A.b_stride ? chain : 0) * g.P;
    double accb = 0, accu = 0;
    // operands of the pointwise part: [0] arg = x - tau (G^T y + z)  (or the prox argument), [1] sigma H^T b
    auto pre = [&](int i, int c, real (&v)[2]) {
      const long long o = (long long)i * g.nx + c;
      if (A.mode == 1) {
        v[0] = A.in[chain * g.P + o];
      } else {
        // G^T y in the accumulation order of pylops (axis 0 then axis 1); v[last] is never read
        real t = real(0);
        if (i < g.ny - 1) t -= y0[o];
        if (i >= 1) t += y0[o - g.nx];
        if (c < g.nx - 1) t -= y1[o];
        if (c >= 1) t += y1[o - 1];
        if (A.zlin != nullptr) t += A.zlin[o];
        v[0] = A.x[chain * g.P + o] - A.tau * t;   // algs.py:439/446: x - tau * ATy
      }
      v[1] = ob[o];
    };
    band_adjoint<real, NT, RB, 2, 4>(g, T, pitch, i0, wx, wy, pre, [&](int i, int c, int k, real gv,
                                                                        const real (&v)[2]) {
      const long long o = (long long)i * g.nx + c;
      const real rhs = v[0] + A.tau * v[1];       // pyproximal L2.prox: y = x + tau * OpTb
      const real zc = X[(k + E) * pitch + BAND_PAD + c];
      const real u = rhs - (zc + A.cA * gv);      // u = b - A x0
      A.u[chain * g.P + o] = u;
      if (!A.warm_on) zw[o] = real(0);
      accb += (double)rhs * (double)rhs;
      accu += (double)u * (double)u;
    });
    const double sb = band_block_sum(accb, red);
    const double su = band_block_sum(accu, red);
    if (threadIdx.x == 0) {
      A.partial[chain * g.tiles + i0 / RB] = sb;
      A.partial[pall + chain * g.tiles + i0 / RB] = su;
    }
  }
}

// ---------------------------------------------------------------- P2: v = Op1 (su * u); w = v
template <typename real, int NT, int RB>
__device__ void band_v0(const UlpdaArgs<real>& A, long long c_lo, long long c_hi, real* smem, double* red,
                        const real (&wx)[NT], const real (&wy)[NT]) {
  constexpr int E = NT - 1;
  BAND_ITEMS_LOOP {
    long long chain; int i0;
    band_of(g, w, c_lo, RB, chain, i0);
    if (!A.st[chain].active) continue;
    const real su = (real)A.st[chain].su;
    __syncthreads();
    band_stage(X, pitch, A.u + chain * g.P, g.ny, g.nx, i0 - E, Band<real, NT, RB>::XROWS);
    band_forward<real, NT, RB>(g, X, T, pitch, i0, wx, wy);
    __syncthreads();
    double acc = 0;
    band_adjoint<real, NT, RB, 0, 4>(g, T, pitch, i0, wx, wy, [&](int, int, real (&)[1]) {},
                                     [&](int i, int c, int k, real gv, const real (&)[1]) {
      const long long o = chain * g.P + (long long)i * g.nx + c;
      const real vv = su * X[(k + E) * pitch + BAND_PAD + c] + A.cA * (su * gv);   // Op1 (su u): the scale commutes
      A.v[o] = vv;
      A.w[o] = vv;
      acc += (double)vv * (double)vv;
    });
    const double s = band_block_sum(acc, red);
    if (threadIdx.x == 0) A.partial[pall + chain * g.tiles + i0 / RB] = s;
  }
}

// ---------------------------------------------------------------- Ka: [deferred x, w update]; u = Op1 v - alfa u
template <typename real, int NT, int RB>
__device__ void band_ka(const UlpdaArgs<real>& A, long long c_lo, long long c_hi, real* smem, double* red,
                        const real (&wx)[NT], const real (&wy)[NT]) {
  constexpr int E = NT - 1;
  BAND_ITEMS_LOOP {
    long long chain; int i0;
    band_of(g, w, c_lo, RB, chain, i0);
    const LsqrState s = A.st[chain];
    if (!s.active) continue;
    const real sv = (real)s.sv, sw = (real)s.sw, su = (real)s.su, t1 = (real)s.t1, t2 = (real)s.t2,
               alfa = (real)s.alfa;
    __syncthreads();
    band_stage(X, pitch, A.v + chain * g.P, g.ny, g.nx, i0 - E, Band<real, NT, RB>::XROWS);
    band_forward<real, NT, RB>(g, X, T, pitch, i0, wx, wy);
    __syncthreads();
    double accu = 0, accw = 0;
    auto pre = [&](int i, int c, real (&v)[3]) {
      const long long o = chain * g.P + (long long)i * g.nx + c;
      v[0] = A.u[o];
      if (s.pending) {
        v[1] = A.w[o];
        v[2] = A.warm[o];
      }
    };
    band_adjoint<real, NT, RB, 3, 4>(g, T, pitch, i0, wx, wy, pre, [&](int i, int c, int k, real gv,
                                                                        const real (&v)[3]) {
      const long long o = chain * g.P + (long long)i * g.nx + c;
      const real vs = sv * X[(k + E) * pitch + BAND_PAD + c];   // the stored v carries the pending scale sv
      if (s.pending) {   // deferred update of the previous LSQR iteration
        real wv = sw * v[1];
        A.warm[o] = v[2] + t1 * wv;    // x = x + t1 * w
        wv = vs + t2 * wv;             // w = v + t2 * w
        A.w[o] = wv;
        accw += (double)wv * (double)wv;
      }
      const real un = (vs + A.cA * (sv * gv)) - alfa * (su * v[0]);   // u = A v - alfa u
      A.u[o] = un;
      accu += (double)un * (double)un;
    });
    const double s1 = band_block_sum(accu, red);
    const double s2 = band_block_sum(accw, red);
    if (threadIdx.x == 0) {
      A.partial[pall + chain * g.tiles + i0 / RB] = s1;
      A.partial[2 * pall + chain * g.tiles + i0 / RB] = s2;
    }
  }
}

// ---------------------------------------------------------------- Kb: v = Op1 u - beta v
template <typename real, int NT, int RB>
__device__ void band_kb(const UlpdaArgs<real>& A, long long c_lo, long long c_hi, real* smem, double* red,
                        const real (&wx)[NT], const real (&wy)[NT]) {
  constexpr int E = NT - 1;
  BAND_ITEMS_LOOP {
    long long chain; int i0;
    band_of(g, w, c_lo, RB, chain, i0);
    const LsqrState s = A.st[chain];
    if (!s.active || !(s.beta > 0)) continue;
    const real su = (real)s.su, sv = (real)s.sv, beta = (real)s.beta;
    __syncthreads();
    band_stage(X, pitch, A.u + chain * g.P, g.ny, g.nx, i0 - E, Band<real, NT, RB>::XROWS);
    band_forward<real, NT, RB>(g, X, T, pitch, i0, wx, wy);
    __syncthreads();
    double acc = 0;
    band_adjoint<real, NT, RB, 1, 4>(g, T, pitch, i0, wx, wy,
                                     [&](int i, int c, real (&v)[1]) { v[0] = A.v[chain * g.P + (long long)i * g.nx + c]; },
                                     [&](int i, int c, int k, real gv, const real (&v)[1]) {
      const long long o = chain * g.P + (long long)i * g.nx + c;
      const real vn = (su * X[(k + E) * pitch + BAND_PAD + c] + A.cA * (su * gv)) - beta * (sv * v[0]);
      A.v[o] = vn;
      acc += (double)vn * (double)vn;
    });
    const double s1 = band_block_sum(acc, red);
    if (threadIdx.x == 0) A.partial[pall + chain * g.tiles + i0 / RB] = s1;
  }
}

// ---------------------------------------------------------------- Pf: last deferred update, noise, xhat, statistics
template <typename real, int RB>
__device__ void band_primal_final(const UlpdaArgs<real>& A, long long c_lo, long long c_hi, long long n) {
  const Geom<real>& g = A.g;
  const long long items = (c_hi - c_lo) * g.tiles;
  const int quads_per_row = (g.nx + 3) / 4;
  const unsigned long long it_global = A.iter_offset + (unsigned long long)n;
  const bool stats = (A.mean != nullptr) && ((long long)it_global >= A.burnin);
  const real inv_cnt = real(1) / (real)(A.stat_count0 + ((long long)it_global - A.burnin) + 1);
  const real cntn = (real)(A.stat_count0 + ((long long)it_global - A.burnin) + 1);
  (void)inv_cnt;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain; int i0;
    band_of(g, w, c_lo, RB, chain, i0);
    const LsqrState s = A.st[chain];
    const real t1 = (real)s.t1, sw = (real)s.sw;
    for (int idx = threadIdx.x; idx < RB * quads_per_row; idx += BAND_BT) {
      const int a = idx / quads_per_row, qd = idx - a * quads_per_row, i = i0 + a, jb = 4 * qd;
      if (i >= g.ny) continue;
      real z[4] = {0, 0, 0, 0};
      if (A.mode == 0 && A.noise == nullptr)
        img_noise4<real>(A.key, A.chain_offset + (unsigned long long)chain, it_global, i, qd, quads_per_row, z);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int j = jb + k;
        if (j >= g.nx) break;
        const long long o = chain * g.P + (long long)i * g.nx + j;
        real zf = A.warm[o];
        if (s.pending) zf += t1 * (sw * A.w[o]);
        A.warm[o] = zf;
        if (A.mode == 1) {
          A.out[o] = zf;
          continue;
        }
        const real xi = (A.noise != nullptr) ? A.noise[(chain * A.N + n) * g.P + (long long)i * g.nx + j] : z[k];
        const real xold = A.x[o];
        const real xn = zf + A.s2t * xi;                 // algs.py:446
        A.x[o] = xn;
        A.xhat[o] = xn + A.theta * (xn - xold);          // algs.py:447
        if (A.samples != nullptr) A.samples[(chain * A.N + n) * g.P + (long long)i * g.nx + j] = xn;
        if (stats) {
          real m = A.mean[o];
          const real d = xn - m;
          m += d / cntn;
          A.mean[o] = m;
          if (A.m2 != nullptr) A.m2[o] += d * (xn - m);
        }
      }
    }
  }
}

// ---------------------------------------------------------------- Pd: y = proxdual(y + mu G xhat)
template <typename real, int RB>
__device__ void band_dual(const UlpdaArgs<real>& A, long long c_lo, long long c_hi, long long n) {
  const Geom<real>& g = A.g;
  const long long items = (c_hi - c_lo) * g.tiles;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain; int i0;
    band_of(g, w, c_lo, RB, chain, i0);
    const real* xh = A.xhat + chain * g.P;
    real* y0 = A.y + chain * 2 * g.P;
    real* y1 = y0 + g.P;
    for (int c = threadIdx.x; c < g.nx; c += BAND_BT) {
      real xv = (i0 < g.ny) ? xh[(long long)i0 * g.nx + c] : real(0);
#pragma unroll 4
      for (int k = 0; k < RB; ++k) {
        const int i = i0 + k;
        if (i >= g.ny) break;
        const long long o = (long long)i * g.nx + c;
        const real xb = (i < g.ny - 1) ? xh[o + g.nx] : real(0);
        const real d0 = (i < g.ny - 1) ? xb - xv : real(0);
        const real d1 = (c < g.nx - 1) ? xh[o + 1] - xv : real(0);
        real v0 = y0[o] + A.mu * d0, v1 = y1[o] + A.mu * d1;   // y + mu * A xhat
        if (A.isotropic) {   // L21 dual prox: v / max(1, |v|_2 / sigma)
          const real nrm = Math<real>::sqrt(v0 * v0 + v1 * v1);
          const real den = Math<real>::max(real(1), nrm / A.g_sigma);
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
        xv = xb;
      }
    }
  }
}

template <typename real, int NT, int RB>
__global__ void __launch_bounds__(BAND_BT, 2) img_ulpda_band_kernel(const __grid_constant__ UlpdaArgs<real> A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  real* smem = reinterpret_cast<real*>(smem_raw);
  __shared__ double red[BAND_BT / 32];
  cg::grid_group grid = cg::this_grid();
  real wx[NT], wy[NT];
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    wx[t] = A.g.wx[t];
    wy[t] = A.g.wy[t];
  }
  // pad columns (and everything else) zero once; the phases only ever write columns [0, nx)
  {
    const int n = (Band<real, NT, RB>::XROWS + Band<real, NT, RB>::TROWS) * A.pitch;
    for (int idx = threadIdx.x; idx < n; idx += BAND_BT) smem[idx] = real(0);
  }
  __syncthreads();
  band_optb<real, NT, RB>(A, smem, wx, wy);
  grid.sync();
  for (long long c_lo = 0; c_lo < A.C; c_lo += A.group) {
    const long long c_hi = (c_lo + A.group < A.C) ? c_lo + A.group : A.C;
    for (long long n = 0; n < A.N; ++n) {
      if (A.mode == 0 && A.gfirst) {
        band_dual<real, RB>(A, c_lo, c_hi, n);
        grid.sync();
      }
      if (blockIdx.x == 0 && threadIdx.x == 0) *A.n_active = 0;
      band_primal_init<real, NT, RB>(A, c_lo, c_hi, smem, red, wx, wy);
      grid.sync();
      phase_s1(A, c_lo, c_hi);
      grid.sync();
      band_v0<real, NT, RB>(A, c_lo, c_hi, smem, red, wx, wy);
      grid.sync();
      phase_s2(A, c_lo, c_hi);
      grid.sync();
      while (*(volatile int*)A.n_active != 0) {
        grid.sync();   // everyone has read n_active before it is reset
        if (blockIdx.x == 0 && threadIdx.x == 0) *A.n_active = 0;
        band_ka<real, NT, RB>(A, c_lo, c_hi, smem, red, wx, wy);
        grid.sync();
        phase_sa(A, c_lo, c_hi);
        grid.sync();
        band_kb<real, NT, RB>(A, c_lo, c_hi, smem, red, wx, wy);
        grid.sync();
        phase_sb(A, c_lo, c_hi);
        grid.sync();
      }
      if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(A.counters + 2, (unsigned long long)(c_hi - c_lo));
      band_primal_final<real, RB>(A, c_lo, c_hi, n);
      grid.sync();
      if (A.mode == 0 && !A.gfirst) {
        band_dual<real, RB>(A, c_lo, c_hi, n);
        grid.sync();
      }
    }
  }
}

}  // namespace
