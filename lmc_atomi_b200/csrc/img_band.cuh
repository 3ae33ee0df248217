// Row-band machinery shared by the band kernels of the image chains (img_ulpda_band.inl, img_myula_band.inl).
//
// A work item is a band of RB full image rows of one chain.  The band plus its blur halo rows is staged once
// in shared memory (coalesced rows, zero pad columns = the zero boundary of Convolve2D); the fused operator
// H^T mask(H x [- b]) is two separable passes with compile-time tap counts, each a horizontal filter out of
// shared memory feeding a vertical sliding window in registers; the pointwise work of a phase runs in the
// epilogue of the second pass.  A thread owns W adjacent image columns (W = 2: 8/16-byte shared/global
// vectors, index arithmetic paid once per pair).  Work items are drawn from a global queue (BandSched).
#pragma once
#include <type_traits>

#include "img_common.cuh"

namespace lmcimg {

constexpr int BAND_PAD = 8;     // zero columns each side of a staged row (>= max taps - 1)
constexpr int BAND_BT = 256;    // threads per CTA
constexpr int BAND_NW = BAND_BT / 32;

template <typename real, int NT, int RB>
struct Band {
  static constexpr int E = NT - 1;
  static constexpr int XROWS = RB + 2 * E;    // staged input rows
  static constexpr int TROWS = RB + E;        // intermediate (H x) rows
  __host__ __device__ static size_t smem_bytes(int pitch) { return (size_t)(XROWS + TROWS) * pitch * sizeof(real); }
};

__host__ __device__ inline int band_pitch(int nx) { return ((nx + 2 * BAND_PAD + 3) / 4) * 4; }

// W adjacent elements moved as one vector
template <typename real, int W>
struct alignas(sizeof(real) * W) Pack {
  real a[W];
};
template <typename real, int W>
__device__ __forceinline__ Pack<real, W> ldp(const real* p) {
  return *reinterpret_cast<const Pack<real, W>*>(p);
}
template <typename real, int W>
__device__ __forceinline__ void stp(real* p, const Pack<real, W>& v) {
  *reinterpret_cast<Pack<real, W>*>(p) = v;
}

// block-wide fp64 sums of two values behind ONE barrier (fixed order: deterministic); results valid in
// thread 0.  `red` (2 x BAND_NW doubles) is next written only after the barrier at the top of the next item.
__device__ __forceinline__ void band_block_sum2(double& a, double& b, double* red) {
  a = lmc_warp_sum(a);
  b = lmc_warp_sum(b);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) {
    red[wid] = a;
    red[BAND_NW + wid] = b;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    a = 0;
    b = 0;
#pragma unroll
    for (int i = 0; i < BAND_NW; ++i) {
      a += red[i];
      b += red[BAND_NW + i];
    }
  }
}

// Dynamic work distribution: the CTAs of a phase draw items from a global counter instead of striding, so
// a CTA on a slower path to memory (far L2 partition, busier SM) simply takes fewer items and the grid
// barrier at the end of the phase is reached together.  Thread 0 draws the next item at the START of the
// current one (the atomic's latency is covered by the item) and publishes it through shared memory at the
// barrier that opens the next round.  Results do not depend on which CTA ran an item (one partial per item).
struct BandSched {
  int* q;        // 4 rotating global counters: phase p uses q[p & 3] and re-arms q[(p + 2) & 3]
  int* s_next;   // 2 shared ints (double-buffered by round)
  int phase, round, nxt;
  unsigned tma_par;   // phase parity of the staging mbarrier (band_stage), tracked per thread
  __device__ __forceinline__ void begin() {
    if (blockIdx.x == 0 && threadIdx.x == 0) q[(phase + 2) & 3] = gridDim.x;
    nxt = blockIdx.x;
  }
  __device__ __forceinline__ int next(int items) {   // also the barrier that frees the staging buffers
    int* slot = s_next + (round & 1);
    if (threadIdx.x == 0) *slot = nxt;
    __syncthreads();
    ++round;
    const int w = *slot;
    if (w < items && threadIdx.x == 0) nxt = atomicAdd(q + (phase & 3), 1);
    return w;
  }
  __device__ __forceinline__ void end() { ++phase; }
};


// ---------------------------------------------------------------- staging through the TMA engine
// band_stage moves whole image rows: with 16-byte aligned rows each is ONE 1-D bulk copy (cp.async.bulk global ->
// shared, SASS UBLKCP) issued by a lane of warp 0 and tracked by an mbarrier (expect_tx = bytes of the in-image rows);
// the other warps issue nothing and go straight to the wait.  Rows outside the image are zero-filled by the CTA.
#ifndef LMC_BAND_TMA
#define LMC_BAND_TMA 1
#endif
__device__ __forceinline__ unsigned long long& band_mbar() {
  __shared__ __align__(8) unsigned long long bar;
  return bar;
}
__device__ __forceinline__ void band_tma_init(BandSched& sched) {
  sched.tma_par = 0;
  if (threadIdx.x == 0) {
    const unsigned bar = (unsigned)__cvta_generic_to_shared(&band_mbar());
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
}

// rows [ifirst, ifirst + rows) of src -> S (zero outside the image): bulk copies (above) when the rows are 16-byte
// aligned, else element copies.  Ends with the data visible to the whole CTA.  The caller's preceding barrier
// (BandSched::next or __syncthreads) has retired every access to S of the previous item.
template <typename real>
__device__ __forceinline__ void band_stage(BandSched& sched, real* S, int pitch, const real* __restrict__ src, int ny,
                                           int nx, int ifirst, int rows) {
  constexpr int EPV = 16 / sizeof(real);
  const bool vec = (nx % EPV == 0) && ((reinterpret_cast<size_t>(src) & 15) == 0);
#if LMC_BAND_TMA
  if (vec && rows <= 32) {
    const int a = ifirst > 0 ? ifirst : 0, b = (ifirst + rows < ny) ? ifirst + rows : ny;   // in-image rows [a, b)
    const unsigned bar = (unsigned)__cvta_generic_to_shared(&band_mbar());
    const unsigned row_bytes = (unsigned)nx * (unsigned)sizeof(real);
    if (threadIdx.x < 32) {
      if (threadIdx.x == 0) {
        // generic-proxy writes to S (zero fill, previous phases) are ordered before the async-proxy writes below
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
                     "r"((unsigned)(b > a ? b - a : 0) * row_bytes)
                     : "memory");
      }
      __syncwarp();
      const int r = (int)threadIdx.x, i = ifirst + r;
      if (r < rows && i >= a && i < b) {
        const unsigned d = (unsigned)__cvta_generic_to_shared(S + r * pitch + BAND_PAD);
        const real* s = src + (long long)i * nx;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d),
                     "l"(s), "r"(row_bytes), "r"(bar)
                     : "memory");
      }
    }
    if (a > ifirst || b < ifirst + rows) {   // top / bottom bands only (uniform over the CTA)
      for (int r = 0; r < rows; ++r) {
        const int i = ifirst + r;
        if (i >= 0 && i < ny) continue;
        for (int c = threadIdx.x * EPV; c < nx; c += BAND_BT * EPV)
          *reinterpret_cast<uint4*>(S + r * pitch + BAND_PAD + c) = make_uint4(0u, 0u, 0u, 0u);
      }
      __syncthreads();
    }
    unsigned done = 0;
    while (!done) {
      asm volatile(
          "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(bar), "r"(sched.tma_par)
          : "memory");
    }
    sched.tma_par ^= 1u;
    return;
  }
#endif
  const int cpr = nx / EPV;   // 16-byte chunks per row
  if (vec && cpr <= BAND_BT && BAND_BT % cpr == 0) {
    // a thread keeps its chunk column and walks down the rows: no per-copy index arithmetic
    const int rstep = BAND_BT / cpr, r0 = threadIdx.x / cpr, c = (threadIdx.x - r0 * cpr) * EPV;
    real* dst = S + r0 * pitch + BAND_PAD + c;
    const real* s = src + (long long)(ifirst + r0) * nx + c;
    const int dstep = rstep * pitch, sstep = rstep * nx;
    for (int r = r0; r < rows; r += rstep, dst += dstep, s += sstep) {
      const int i = ifirst + r;
      if (i >= 0 && i < ny) {
        const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(s) : "memory");
      } else {
        *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
  } else {
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
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
}

// The k x k blurs of the driver (prox_lmc_deconv.py:55-69): first tap of the forward window and the parities of the
// first shared-memory column the forward / adjoint passes read for an even image column.  Two columns per
// thread (W = 2) need vector loads from even columns: an odd window start is read from one column earlier.
template <int NT>
struct BandWin {
  static constexpr int LO = (NT == 7) ? -3 : -2;   // pylops pads an even kernel by one trailing zero row / column
  static constexpr int FWD_OFF = (LO & 1), ADJ_OFF = ((LO + NT - 1) & 1);
};

// the NT + W - 1 consecutive elements starting at p; p - OFF is W-element aligned
template <typename real, int NT, int W, int OFF>
__device__ __forceinline__ void band_taps(const real* p, real (&e)[NT + W - 1]) {
  if constexpr (W == 1) {
#pragma unroll
    for (int s = 0; s < NT; ++s) e[s] = p[s];
  } else {
    constexpr int NE = NT + W - 1, NV = (NE + OFF + W - 1) / W;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const Pack<real, W> t = ldp<real, W>(p - OFF + W * v);
#pragma unroll
      for (int q = 0; q < W; ++q)
        if (W * v + q >= OFF && W * v + q - OFF < NE) e[W * v + q - OFF] = t.a[q];
    }
  }
}

template <typename real, int W, int NPRE, int PFD>
using BandPf = Pack<real, W>[2][PFD][NPRE > 0 ? NPRE : 1];   // two blocks of PFD rows in flight

// request the epilogue operands of the first two row blocks of columns [c, c + W)
template <typename real, int RB, int W, int NPRE, int PFD, class Pre>
__device__ __forceinline__ void band_prefetch(const Geom<real>& g, int i0, int c, BandPf<real, W, NPRE, PFD>& pf,
                                              Pre&& pre) {
  constexpr int NBLK = RB / PFD;
  if (NPRE > 0 && c < g.nx) {
#pragma unroll
    for (int bl = 0; bl < 2 && bl < NBLK; ++bl)
#pragma unroll
      for (int kk = 0; kk < PFD; ++kk) pre(min(i0 + bl * PFD + kk, g.ny - 1), c, pf[bl][kk]);
  }
}

// second half of the operator: T (rows i0 - hi_y .. of a row-masked image) -> H^T T on the band.
// pre(i, c, v) loads the NPRE global operands of pixels (i, c .. c+W-1) of the phase's pointwise update; they
// are fetched PFD rows ahead (rows past the image are clamped, not skipped: no branch around the loads);
// the first two blocks of the thread's first column group arrive in `pf` (band_prefetch, issued by the
// caller before staging).  epi(i, c, k, value, v) consumes them (k = row within the band), in-image rows only.
template <typename real, int NT, int RB, int W, int NPRE, int PFD, class Pre, class Epi>
__device__ __forceinline__ void band_adjoint(const Geom<real>& g, const real* T, int pitch, int i0, const real* wx,
                                             const real* wy, BandPf<real, W, NPRE, PFD>& pf, Pre&& pre,
                                             Epi&& epi) {
  static_assert(RB % PFD == 0, "prefetch depth must divide the band height");
  constexpr int NBLK = RB / PFD;
  const int lox = g.lo_x;
  for (int c = W * threadIdx.x; c < g.nx; c += W * BAND_BT) {
    if (c != W * (int)threadIdx.x) band_prefetch<real, RB, W, NPRE, PFD>(g, i0, c, pf, pre);
    const real* trow = T + BAND_PAD + c - lox - (NT - 1);
    real win[W][NT];
#pragma unroll
    for (int jj = 0; jj < RB + NT - 1; ++jj) {
      real e[NT + W - 1];
      band_taps<real, NT, W, (W > 1 ? BandWin<NT>::ADJ_OFF : 0)>(trow + jj * pitch, e);
#pragma unroll
      for (int q = 0; q < W; ++q) {
        real u = real(0);
#pragma unroll
        for (int s = 0; s < NT; ++s) u = Math<real>::fma(wx[s], e[NT - 1 - s + q], u);
        win[q][jj % NT] = u;
      }
      if (jj >= NT - 1) {
        const int k = jj - (NT - 1), i = i0 + k;
        Pack<real, W> gv;
#pragma unroll
        for (int q = 0; q < W; ++q) {
          real r = real(0);
#pragma unroll
          for (int t = 0; t < NT; ++t) r = Math<real>::fma(wy[t], win[q][(k + NT - 1 - t) % NT], r);
          gv.a[q] = r;
        }
        if (i < g.ny) epi(i, c, k, gv, pf[(k / PFD) & 1][k % PFD]);
        if (NPRE > 0 && (k % PFD) == PFD - 1 && k / PFD + 2 < NBLK) {   // block consumed: fetch block + 2 into its slot
#pragma unroll
          for (int kk = 0; kk < PFD; ++kk) pre(min(i0 + k + 1 + PFD + kk, g.ny - 1), c, pf[(k / PFD) & 1][kk]);
        }
      }
    }
  }
}

// rows of the forward pass that need no loads are computed and selected instead of branched around (ULPDA: +1.6 %)
#ifndef LMC_BAND_FWD_SELECT
#define LMC_BAND_FWD_SELECT 1
#endif
struct BandNoSub {};

// first half: staged X (rows i0 - E ..) -> T = mask_rows(H X [- sub]) on rows [i0 - hi_y, i0 + RB - lo_y).
// `sub(i, c)` (optional) returns the Pack subtracted at in-image pixels (i, c .. c+W-1): the data b of the residual.
template <typename real, int NT, int RB, int W, class Sub = BandNoSub>
__device__ __forceinline__ void band_forward(const Geom<real>& g, const real* X, real* T, int pitch, int i0,
                                             const real* wx, const real* wy, Sub sub = Sub()) {
  const int lox = g.lo_x, hi_y = g.lo_y + NT - 1;
  for (int c = W * threadIdx.x; c < g.nx; c += W * BAND_BT) {
    const real* xrow = X + BAND_PAD + c + lox;
    real* trow = T + BAND_PAD + c;
    real win[W][NT];
#pragma unroll
    for (int jj = 0; jj < RB + 2 * (NT - 1); ++jj) {
      real e[NT + W - 1];
      band_taps<real, NT, W, (W > 1 ? BandWin<NT>::FWD_OFF : 0)>(xrow + jj * pitch, e);
#pragma unroll
      for (int q = 0; q < W; ++q) {
        real t1 = real(0);
#pragma unroll
        for (int s = 0; s < NT; ++s) t1 = Math<real>::fma(wx[s], e[s + q], t1);
        win[q][jj % NT] = t1;
      }
      if (jj >= NT - 1) {
        const int o = jj - (NT - 1), i = i0 - hi_y + o;
        Pack<real, W> res;
#pragma unroll
        for (int q = 0; q < W; ++q) res.a[q] = real(0);
#if LMC_BAND_FWD_SELECT
        if constexpr (std::is_same<Sub, BandNoSub>::value) {   // no loads in the row: compute it and select, no branch
          const bool in = i >= 0 && i < g.ny;
#pragma unroll
          for (int q = 0; q < W; ++q) {
            real r = real(0);
#pragma unroll
            for (int t = 0; t < NT; ++t) r = Math<real>::fma(wy[t], win[q][(o + t) % NT], r);
            res.a[q] = in ? r : real(0);
          }
        } else
#endif
        if (i >= 0 && i < g.ny) {
#pragma unroll
          for (int q = 0; q < W; ++q) {
            real r = real(0);
#pragma unroll
            for (int t = 0; t < NT; ++t) r = Math<real>::fma(wy[t], win[q][(o + t) % NT], r);
            res.a[q] = r;
          }
          if constexpr (!std::is_same<Sub, BandNoSub>::value) {
            const Pack<real, W> bv = sub(i, c);
#pragma unroll
            for (int q = 0; q < W; ++q) res.a[q] -= bv.a[q];
          }
        }
        stp<real, W>(trow + o * pitch, res);
      }
    }
  }
}

template <typename real>
__device__ __forceinline__ void band_of(const Geom<real>& g, int w, long long c_lo, int RB, long long& chain, int& i0) {
  const int cl = w / g.tiles;
  chain = c_lo + cl;
  i0 = (w - cl * g.tiles) * RB;
}

// L2 prefetch of rows [r0, r0 + rows) of one image (clamped to the image): one bulk-prefetch instruction
// (cp.async.bulk.prefetch.L2, SASS UBLKPF) issued by a single thread for the rows the CTA's NEXT work item will stage,
// so that its staging loads find them in L2 instead of paying the HBM latency in front of the forward pass.
#ifndef LMC_BAND_L2_PREFETCH
#define LMC_BAND_L2_PREFETCH 1
#endif
template <typename real>
__device__ __forceinline__ void band_l2_prefetch(const real* base, int ny, int nx, int r0, int rows) {
#if LMC_BAND_L2_PREFETCH
  const int a = r0 > 0 ? r0 : 0, b = (r0 + rows < ny) ? r0 + rows : ny;
  if (b <= a) return;
  size_t p0 = (size_t)(base + (size_t)a * nx), p1 = (size_t)(base + (size_t)b * nx);
  p0 = (p0 + 15) & ~(size_t)15;
  p1 &= ~(size_t)15;
  if (p1 > p0)
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p0), "r"((unsigned)(p1 - p0)) : "memory");
#endif
}

}  // namespace lmcimg
