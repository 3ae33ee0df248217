// MYULA, row-band formulation (included by img_myula.cu; same arguments, TV state machine and scalar phase as
// the tile kernel above, different data movement -- the general path for images that do not fit the on-chip
// cluster kernel, e.g. the reference's 512 x 512 camera image, and for fp64).
//
//   G   : a band of 16 rows + blur halo staged once; g = sigma H^T mask(Hx - b) by the two fused separable
//         passes of img_band.cuh; TV sweep 0 (sol_0 = x) runs in the epilogue out of the staged x
//   T_k : sub-bands of 8 rows: rr, ss = p_k + c_k (p_k - p_{k-1}) staged with one halo row, sol_k on 9 rows into
//         shared memory, then the projected dual update and the objective of the pixel
//   U   : sol_kf recomputed from the staged (rr, ss), the MYULA update, running statistics
// Two image columns per thread (8/16-byte vectors); work items drawn from the BandSched queue.
#include "img_band.cuh"

namespace {

constexpr int MB_RB = 16;   // rows per band = per objective partial
constexpr int MB_HB = 8;    // rows per TV sub-band

template <typename real, int NT>
__host__ __device__ inline size_t myula_band_smem_bytes(int pitch) {
  const int a = Band<real, NT, MB_RB>::XROWS + Band<real, NT, MB_RB>::TROWS, b = 3 * (MB_HB + 2);
  return (size_t)(a > b ? a : b) * pitch * sizeof(real);
}

// rr = p_it + c (p_it - p_{it-1}) (both components) on image rows [r_first, r_first + nrows), all columns;
// zero outside the image (pad columns are never written and stay zero)
template <typename real, int TW>
__device__ __forceinline__ void mband_stage_rr(const MyulaArgs<real>& A, long long chain, int it, int r_first, int nrows,
                                               real* RR, real* SS) {
  constexpr int MB_W = TW;   // image columns per thread of the TV phases (2; 1 for odd widths)
  using P2 = Pack<real, MB_W>;
  const Geom<real>& g = A.g;
  const int pitch = A.pitch;
  const real c = (real)fista_c(it);
  const real* pa = dual_ptr(A, it % 3, 0, chain);
  const real* qa = dual_ptr(A, it % 3, 1, chain);
  const real* pb = dual_ptr(A, (it + 2) % 3, 0, chain);
  const real* qb = dual_ptr(A, (it + 2) % 3, 1, chain);
  constexpr int NB = 5;   // rows whose loads are in flight together
  for (int col = MB_W * threadIdx.x; col < g.nx; col += MB_W * BAND_BT) {
    for (int a0 = 0; a0 < nrows; a0 += NB) {
      P2 p[NB], q2[NB], pp[NB], qq[NB];
      if (it >= 1) {
#pragma unroll
        for (int k = 0; k < NB; ++k) {   // rows outside the image are clamped (their values are discarded below)
          const int o = min(max(r_first + a0 + k, 0), g.ny - 1) * g.nx + col;
          p[k] = ldp<real, MB_W>(pa + o);
          q2[k] = ldp<real, MB_W>(qa + o);
          if (it >= 2) {
            pp[k] = ldp<real, MB_W>(pb + o);
            qq[k] = ldp<real, MB_W>(qb + o);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < NB; ++k) {
        const int a = a0 + k, i = r_first + a;
        if (a >= nrows) break;
        P2 r, s;
#pragma unroll
        for (int q = 0; q < MB_W; ++q) r.a[q] = s.a[q] = real(0);
        if (it >= 1 && i >= 0 && i < g.ny) {
          if (it >= 2) {
#pragma unroll
            for (int q = 0; q < MB_W; ++q) {
              r.a[q] = p[k].a[q] + c * (p[k].a[q] - pp[k].a[q]);
              s.a[q] = q2[k].a[q] + c * (q2[k].a[q] - qq[k].a[q]);
            }
          } else {   // c_1 = 0 and p_0 = 0
            r = p[k];
            s = q2[k];
          }
        }
        stp<real, MB_W>(RR + a * pitch + BAND_PAD + col, r);
        stp<real, MB_W>(SS + a * pitch + BAND_PAD + col, s);
      }
    }
  }
}

// sol = x - gamma div(rr, ss) for columns [col, col + W) of local row a (RR/SS row a = image row, row a-1 above)
template <typename real, int MB_W>
__device__ __forceinline__ Pack<real, MB_W> mband_sol(const real* RR, const real* SS, int pitch, int a, int col,
                                                      const Pack<real, MB_W>& xv, real gamma) {
  using P2 = Pack<real, MB_W>;
  const P2 rc = ldp<real, MB_W>(RR + a * pitch + BAND_PAD + col), ru = ldp<real, MB_W>(RR + (a - 1) * pitch + BAND_PAD + col);
  const P2 sc = ldp<real, MB_W>(SS + a * pitch + BAND_PAD + col);
  const real sl = SS[a * pitch + BAND_PAD + col - 1];
  P2 out;
#pragma unroll
  for (int q = 0; q < MB_W; ++q) {
    const real t = ((-rc.a[q] + ru.a[q]) - sc.a[q]) + (q == 0 ? sl : sc.a[q > 0 ? q - 1 : 0]);   // order of sol_at()
    out.a[q] = xv.a[q] - gamma * (-t);
  }
  return out;
}

// projected dual update of one pixel.  fp64: the reference's divisions (dual_update of the tile kernel).  fp32: the
// shortcuts of the on-chip kernel, far inside the 1e-5 trajectory tolerance: dx / (8 gamma) is a multiply by the
// host-rounded reciprocal and r / max(1, |(r,s)|) is r * min(1, rsqrt(r^2 + s^2)).
template <typename real>
__device__ __forceinline__ void mband_dual_update(real rr, real ss, real dx, real dy, real g8, real inv_g8, real& p,
                                                  real& q) {
  if constexpr (sizeof(real) == 4) {
    const real r = rr - dx * inv_g8, s = ss - dy * inv_g8;
    const real w = Math<real>::min(real(1), Math<real>::rsqrt(Math<real>::max(r * r + s * s, real(1e-30))));
    p = r * w;
    q = s * w;
  } else {
    dual_update(rr, ss, dx, dy, g8, p, q);
  }
}

// ---------------------------------------------------------------- phase G (+ sweep 0)
// GW = image columns per thread of this phase (2 when the blur's tap windows start on even columns, else 1)
template <typename real, int NT, int GW, bool NC>
__device__ void mband_grad_sweep0(const MyulaArgs<real>& A, BandSched& sched, real* smem, double* red) {
  using P2 = Pack<real, GW>;
  constexpr int RB = MB_RB, E = NT - 1;
  const Geom<real>& g = A.g;
  const int pitch = A.pitch, items = (int)A.C * g.tiles;
  real* X = smem;
  real* T = smem + Band<real, NT, RB>::XROWS * pitch;
  const real* wx = A.g.wx;
  const real* wy = A.g.wy;
  const real g8 = real(8) * A.tv_gamma, inv_g8 = (real)(1.0 / (8.0 * (double)A.tv_gamma));
  sched.begin();
  for (int w; (w = sched.next(items)) < items || (sched.end(), false);) {
    long long chain; int i0;
    band_of(g, w, 0, RB, chain, i0);
    const real* bc = A.b + chain * A.b_stride;
    real* gout = A.grad + chain * g.P;
    real* p1 = dual_ptr(A, 1, 0, chain);
    real* q1 = dual_ptr(A, 1, 1, chain);
    band_stage(sched, X, pitch, A.x + chain * g.P, g.ny, g.nx, i0 - E, Band<real, NT, RB>::XROWS);
    band_forward<real, NT, RB, GW>(g, X, T, pitch, i0, wx, wy,
                                     [&](int i, int c) { return ldp<real, GW>(bc + i * g.nx + c); });
    __syncthreads();
    double acc = 0, none = 0;
    BandPf<real, GW, 0, 4> pf;
    band_adjoint<real, NT, RB, GW, 0, 4>(g, T, pitch, i0, wx, wy, pf, [&](int, int, P2 (&)[1]) {},
                                           [&](int i, int c, int k, const P2& gv, const P2 (&)[1]) {
      const int o = i * g.nx + c;
      P2 go;
#pragma unroll
      for (int q = 0; q < GW; ++q) go.a[q] = A.sigma_l2 * gv.a[q];
      stp<real, GW>(gout + o, go);
      if (NC && A.nc.kind == 1) {   // MC-TV: 7-point stencil out of the staged band (halo rows E >= 1, zero pad columns)
        P2 ge;
#pragma unroll
        for (int q = 0; q < GW; ++q) {
          const real* xp = X + (k + E) * pitch + BAND_PAD + c + q;
          ge.a[q] = mctv_at(A.nc.W, i, c + q, g.ny, g.nx, [&](int di, int dj) { return xp[di * pitch + dj]; }) /
                    A.nc.post_div;
        }
        stp<real, GW>(A.gextra_w + chain * g.P + o, ge);
      }
      // TV sweep 0: rr = ss = 0 -> sol = x, obj = gamma sum |grad x|, p_1 = proj(-grad x / (8 gamma))
      const real* xr = X + (k + E) * pitch + BAND_PAD + c;
      const P2 xv = ldp<real, GW>(xr), xd = ldp<real, GW>(xr + pitch);
      const real xright = xr[GW];
      P2 pn, qn;
#pragma unroll
      for (int q = 0; q < GW; ++q) {
        const real dx = (i < g.ny - 1) ? xd.a[q] - xv.a[q] : real(0);
        const real dy = (c + q < g.nx - 1) ? (q + 1 < GW ? xv.a[q + 1 < GW ? q + 1 : 0] : xright) - xv.a[q] : real(0);
        acc += A.tv_objw * (double)Math<real>::sqrt(dx * dx + dy * dy);
        mband_dual_update(real(0), real(0), dx, dy, g8, inv_g8, pn.a[q], qn.a[q]);
      }
      stp<real, GW>(p1 + o, pn);
      stp<real, GW>(q1 + o, qn);
    });
    band_block_sum2(acc, none, red);
    if (threadIdx.x == 0) A.partial[chain * g.tiles + i0 / RB] = acc;
  }
}

// ---------------------------------------------------------------- phase T_it
template <typename real, int TW>
__device__ void mband_tv_sweep(const MyulaArgs<real>& A, BandSched& sched, int it, real* smem, double* red) {
  constexpr int MB_W = TW;
  using P2 = Pack<real, MB_W>;
  constexpr int RB = MB_RB, HB = MB_HB;
  const Geom<real>& g = A.g;
  const int pitch = A.pitch, items = (int)A.C * g.tiles;
  real* RR = smem;                        // local row a <-> image row r0 - 1 + a, a = 0 .. HB + 1
  real* SS = RR + (HB + 2) * pitch;
  real* SOL = SS + (HB + 2) * pitch;      // local row a <-> image row r0 + a, a = 0 .. HB
  const real g8 = real(8) * A.tv_gamma, inv_g8 = (real)(1.0 / (8.0 * (double)A.tv_gamma));
  sched.begin();
  for (int w; (w = sched.next(items)) < items || (sched.end(), false);) {
    long long chain; int i0;
    band_of(g, w, 0, RB, chain, i0);
    if (it > 0 && !A.st[chain].active) continue;   // it == 0: the standalone sweep 0 of a prox-only pipeline (mode 2)
    const real* xc = A.x + chain * g.P;
    real* pn = dual_ptr(A, (it + 1) % 3, 0, chain);
    real* qn = dual_ptr(A, (it + 1) % 3, 1, chain);
    double acc = 0, none = 0;
    for (int r0 = i0; r0 < i0 + RB && r0 < g.ny; r0 += HB) {
      // x on the 9 rows of the sub-band (rows below the image clamped, zeroed when used): requested before the
      // duals are staged, consumed two barriers later by sol and again by the pixel update
      auto load_x = [&](int col, P2 (&xr)[HB + 1]) {
#pragma unroll
        for (int a = 0; a <= HB; ++a) xr[a] = ldp<real, MB_W>(xc + min(r0 + a, g.ny - 1) * g.nx + col);
      };
      const int col0 = MB_W * threadIdx.x;
      P2 xr[HB + 1];
      if (col0 < g.nx) load_x(col0, xr);
      if (r0 != i0) __syncthreads();   // the previous sub-band's readers are done
      mband_stage_rr<real, TW>(A, chain, it, r0 - 1, HB + 2, RR, SS);
      __syncthreads();
      for (int col = col0; col < g.nx; col += MB_W * BAND_BT) {
        if (col != col0) load_x(col, xr);
#pragma unroll
        for (int a = 0; a <= HB; ++a) {
          P2 xv = xr[a];
          if (r0 + a >= g.ny) {   // below the image: x = 0 (the row is never read by an in-image pixel's dx)
#pragma unroll
            for (int q = 0; q < MB_W; ++q) xv.a[q] = real(0);
          }
          stp<real, MB_W>(SOL + a * pitch + BAND_PAD + col, mband_sol<real, MB_W>(RR, SS, pitch, a + 1, col, xv, A.tv_gamma));
        }
      }
      __syncthreads();
      for (int col = col0; col < g.nx; col += MB_W * BAND_BT) {
        if (col != col0 || g.nx > MB_W * BAND_BT) load_x(col, xr);   // more than one column group: xr holds the last one
#pragma unroll
        for (int a = 0; a < HB; ++a) {
          const int i = r0 + a;
          if (i < g.ny) {
            const int o = i * g.nx + col;
            const real* sr = SOL + a * pitch + BAND_PAD + col;
            const P2 sv = ldp<real, MB_W>(sr), sd = ldp<real, MB_W>(sr + pitch), xv = xr[a];
            const real sright = sr[MB_W];
            const P2 rc = ldp<real, MB_W>(RR + (a + 1) * pitch + BAND_PAD + col),
                     sc = ldp<real, MB_W>(SS + (a + 1) * pitch + BAND_PAD + col);
            P2 po, qo;
#pragma unroll
            for (int q = 0; q < MB_W; ++q) {
              const real dx = (i < g.ny - 1) ? sd.a[q] - sv.a[q] : real(0);
              const real dy = (col + q < g.nx - 1) ? (q + 1 < MB_W ? sv.a[q + 1 < MB_W ? q + 1 : 0] : sright) - sv.a[q] : real(0);
              const real df = xv.a[q] - sv.a[q];
              acc += 0.5 * (double)df * (double)df + A.tv_objw * (double)Math<real>::sqrt(dx * dx + dy * dy);
              mband_dual_update(rc.a[q], sc.a[q], dx, dy, g8, inv_g8, po.a[q], qo.a[q]);
            }
            stp<real, MB_W>(pn + o, po);
            stp<real, MB_W>(qn + o, qo);
          }
        }
      }
    }
    band_block_sum2(acc, none, red);
    if (threadIdx.x == 0) A.partial[chain * g.tiles + i0 / RB] = acc;
  }
}

// four consecutive pixels of one row as one 16/32-byte access when the rows allow it, else n <= 4 scalar ones
template <typename real>
__device__ __forceinline__ Pack<real, 4> mb_ld4(const real* p, int n, bool vec) {
  if (vec) return ldp<real, 4>(p);
  Pack<real, 4> v;
#pragma unroll
  for (int k = 0; k < 4; ++k) v.a[k] = (k < n) ? p[k] : real(0);
  return v;
}
template <typename real>
__device__ __forceinline__ void mb_st4(real* p, const Pack<real, 4>& v, int n, bool vec) {
  if (vec) {
    stp<real, 4>(p, v);
    return;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (k < n) p[k] = v.a[k];
}

// ---------------------------------------------------------------- phase U
template <typename real, int TW, bool NC>
__device__ void mband_update(const MyulaArgs<real>& A, BandSched& sched, long long n, real* smem) {
  constexpr int RB = MB_RB, HB = MB_HB;
  const Geom<real>& g = A.g;
  const int pitch = A.pitch, items = (int)A.C * g.tiles;
  real* RR = smem;                        // local row a <-> image row r0 - 1 + a, a = 0 .. HB
  real* SS = RR + (HB + 2) * pitch;
  const int quads_per_row = (g.nx + 3) / 4;
  const unsigned long long it_global = A.iter_offset + (unsigned long long)n;
  const bool stats = (A.mean != nullptr) && ((long long)it_global >= A.burnin);
  const real cntn = (real)(A.stat_count0 + ((long long)it_global - A.burnin) + 1);
  using P4 = Pack<real, 4>;
  bool vec = (g.nx % 4 == 0);
  {
    const void* ptrs[] = {A.x, A.grad, A.gextra, A.noise, A.samples, A.mean, A.m2, A.out};
    for (const void* p : ptrs) vec = vec && (reinterpret_cast<size_t>(p) % (4 * sizeof(real)) == 0);
  }
  sched.begin();
  for (int w; (w = sched.next(items)) < items || (sched.end(), false);) {
    long long chain; int i0;
    band_of(g, w, 0, RB, chain, i0);
    const int kf = A.st[chain].kf;
    for (int r0 = i0; r0 < i0 + RB && r0 < g.ny; r0 += HB) {
      if (r0 != i0) __syncthreads();
      mband_stage_rr<real, TW>(A, chain, kf, r0 - 1, HB + 1, RR, SS);
      __syncthreads();
      // each thread: 4 consecutive pixels of one row (one Philox block of normals), moved as one 16/32-byte access
      // per array when the rows allow it; all loads of the quad are issued before its arithmetic
      for (int idx = threadIdx.x; idx < HB * quads_per_row; idx += BAND_BT) {
        const int a = idx / quads_per_row, qd = idx - a * quads_per_row, i = r0 + a, jb = 4 * qd;
        if (i >= g.ny) continue;
        const int nq = min(4, g.nx - jb);
        const long long o = chain * g.P + (long long)i * g.nx + jb;
        const long long os = (chain * A.N + n) * g.P + (long long)i * g.nx + jb;
        if (NC && A.mode == 2) {   // ME-TV prelude: out = (x - prox) / gamma, nothing else
          const P4 xv = mb_ld4(A.x + o, nq, vec);
          P4 ov;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const real* rr = RR + (a + 1) * pitch + BAND_PAD + jb + k;
            const real* ss = SS + (a + 1) * pitch + BAND_PAD + jb + k;
            const real t = ((-rr[0] + rr[-pitch]) - ss[0]) + ss[-1];
            const real sol = xv.a[k] - A.tv_gamma * (-t);
            ov.a[k] = (xv.a[k] - sol) / A.me_den;
          }
          mb_st4(A.out + o, ov, nq, vec);
          continue;
        }
        const P4 xv = mb_ld4(A.x + o, nq, vec), gr0 = mb_ld4(A.grad + o, nq, vec);
        P4 ge, xi, mn, m2v;
        if (A.gextra != nullptr) ge = mb_ld4(A.gextra + o, nq, vec);
        if (A.noise != nullptr) xi = mb_ld4(A.noise + os, nq, vec);
        if (stats) {
          mn = mb_ld4(A.mean + o, nq, vec);
          if (A.m2 != nullptr) m2v = mb_ld4(A.m2 + o, nq, vec);
        }
        if (A.noise == nullptr)
          img_noise4<real>(A.key, A.chain_offset + (unsigned long long)chain, it_global, i, qd, quads_per_row, xi.a);
        P4 xn;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const real* rr = RR + (a + 1) * pitch + BAND_PAD + jb + k;
          const real* ss = SS + (a + 1) * pitch + BAND_PAD + jb + k;
          const real t = ((-rr[0] + rr[-pitch]) - ss[0]) + ss[-1];   // order of sol_at()
          const real sol = xv.a[k] - A.tv_gamma * (-t);
          real gr = gr0.a[k];
          if (A.gextra != nullptr) gr = gr + A.gscale * ge.a[k];   // algs.py:291
          // algs.py:569: (1 - tau/gamma) x - tau grad f(x) + tau/gamma prox_g(x) + sqrt(2 tau) xi
          xn.a[k] = A.a_x * xv.a[k] - A.tau * gr + A.tog * sol + A.s2t * xi.a[k];
        }
        mb_st4(A.x + o, xn, nq, vec);
        if (A.samples != nullptr) mb_st4(A.samples + os, xn, nq, vec);
        if (stats) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const real d = xn.a[k] - mn.a[k];
            mn.a[k] += d / cntn;
            if (A.m2 != nullptr) m2v.a[k] += d * (xn.a[k] - mn.a[k]);
          }
          mb_st4(A.mean + o, mn, nq, vec);
          if (A.m2 != nullptr) mb_st4(A.m2 + o, m2v, nq, vec);
        }
      }
    }
  }
}

// NC = the launch carries a non-convex data term (MC-TV stencil in phase G and / or the ME-TV prelude B): a separate
// instantiation, so that the plain kernel keeps its register budget (the stencil inside the unrolled epilogue cost
// the shared instantiation 1.3 KB of spills and 19 % of its throughput, r02m)
template <typename real, int NT, int GW, int TW, bool NC>
__device__ __forceinline__ void myula_band_iteration(const MyulaArgs<real>& A, BandSched& sched, long long n, real* smem,
                                                     double* red, cg::grid_group& grid) {
  if (blockIdx.x == 0 && threadIdx.x == 0) *A.n_active = 0;
  if (!NC || A.mode == 0) mband_grad_sweep0<real, NT, GW, NC>(A, sched, smem, red);
  else mband_tv_sweep<real, TW>(A, sched, 0, smem, red);   // prox-only pipeline: sweep 0 without the data term
  grid.sync();
  phase_tv_scalar(A, 0);
  grid.sync();
  for (int it = 1; it < A.tv_passes; ++it) {
    if (*(volatile int*)A.n_active == 0) break;
    grid.sync();   // everyone has read n_active before it is reset
    if (blockIdx.x == 0 && threadIdx.x == 0) *A.n_active = 0;
    mband_tv_sweep<real, TW>(A, sched, it, smem, red);
    grid.sync();
    phase_tv_scalar(A, it);
    grid.sync();
  }
  mband_update<real, TW, NC>(A, sched, n, smem);
  grid.sync();
}

// A = the sampler, B = the ME-TV prelude or disabled (see img_myula_kernel)
template <typename real, int NT, int GW, int TW, bool NC>
__global__ void __launch_bounds__(BAND_BT, 2) img_myula_band_kernel(const __grid_constant__ MyulaArgs<real> A,
                                                                    const __grid_constant__ MyulaArgs<real> B) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  real* smem = reinterpret_cast<real*>(smem_raw);
  __shared__ double red[2 * BAND_NW];
  __shared__ int s_next[2];
  cg::grid_group grid = cg::this_grid();
  BandSched sched{A.queue, s_next, 0, 0, 0};
  if (blockIdx.x == 0 && threadIdx.x < 4) A.queue[threadIdx.x] = gridDim.x;
  // pad columns (and everything else) zero once; the phases only ever write columns [0, nx)
  for (int idx = threadIdx.x; idx < (int)(myula_band_smem_bytes<real, NT>(A.pitch) / sizeof(real)); idx += BAND_BT)
    smem[idx] = real(0);
  band_tma_init(sched);
  grid.sync();
  for (long long n = 0; n < A.N; ++n) {
    if (NC && B.mode == 2) myula_band_iteration<real, NT, GW, TW, true>(B, sched, n, smem, red, grid);
    myula_band_iteration<real, NT, GW, TW, NC>(A, sched, n, smem, red, grid);
  }
}

}  // namespace
