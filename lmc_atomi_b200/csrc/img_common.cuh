// Tile machinery shared by the image-chain kernels (MYULA / ULPDA deconvolution).
//
// Data layout in HBM: every per-chain array is a dense row-major (ny, nx) image,
// chains stacked with stride P = ny*nx ("(C, P)" arrays).  A work item is one
// TH x TW tile of one chain; persistent CTAs stride over the items, stage the tile
// plus its halo in shared memory (zero fill outside the image = the zero boundary
// of pylops Convolve2D), and run the separable blur / difference stencils out of
// shared memory.  Reductions (norms, objectives) are accumulated in fp64 in both
// arithmetic modes, written as one partial per (chain, tile) and summed in a fixed
// order by the scalar phases, so every run is deterministic.
#pragma once
#include <cooperative_groups.h>

#include "lmc_common.cuh"

namespace cg = cooperative_groups;

namespace lmcimg {

constexpr int TH = 32;              // tile rows
constexpr int TW = 32;              // tile cols
constexpr int HMAX = 6;             // max halo = max blur taps - 1  (7-tap box -> 6)
constexpr int RS = TH + 2 * HMAX;   // side of the staged region (44)
constexpr int NTHREADS = 256;
constexpr int MAXTAP = LMC_IMG_MAX_TAPS;

template <typename real>
struct Geom {
  int ny, nx, tiles_y, tiles_x, tiles;  // tiles per chain
  long long P;                          // ny * nx
  int ntap_y, ntap_x, lo_y, lo_x;       // forward blur: out[i] = sum_t w[t] * in[i + lo + t]
  int e;                                // halo of the fused H^T H stencil: max(ntap) - 1
  real wy[MAXTAP], wx[MAXTAP];
};

// ---- block-wide fp64 sum (all threads must call); result valid in thread 0
__device__ __forceinline__ double block_sum(double v, double* red /* >= 8 doubles */) {
  v = lmc_warp_sum(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  __syncthreads();  // protect `red` from the previous use
  if (lane == 0) red[wid] = v;
  __syncthreads();
  double s = 0;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < NTHREADS / 32; ++i) s += red[i];
  }
  return s;
}

// ---- stage a (rows x cols) window of an image whose top-left is (r0, c0); zero outside
template <typename real>
__device__ __forceinline__ void load_region(real* S, int pitch, const real* img,
                                            int ny, int nx, int r0, int c0, int rows, int cols) {
  for (int idx = threadIdx.x; idx < rows * cols; idx += NTHREADS) {
    int a = idx / cols, b = idx - a * cols;
    int i = r0 + a, j = c0 + b;
    real v = real(0);
    if (i >= 0 && i < ny && j >= 0 && j < nx) v = img[(long long)i * nx + j];
    S[a * pitch + b] = v;
  }
}

// ---- one separable pass.  out[a][b] = sum_t w[t] * in[a + SA*(lo+t)][b + SB*(lo+t)]
// for a in [a0,a1), b in [b0,b1).  (SA,SB) = (0,1) row pass, (1,0) column pass,
// negative = adjoint direction.
template <typename real, int SA, int SB>
__device__ __forceinline__ void blur_pass(const real* __restrict__ in, real* __restrict__ out, int pitch,
                                          int a0, int a1, int b0, int b1, const real* __restrict__ w,
                                          int ntap, int lo) {
  const int cols = b1 - b0, n = (a1 - a0) * cols;
  for (int idx = threadIdx.x; idx < n; idx += NTHREADS) {
    int a = a0 + idx / cols, b = b0 + idx % cols;
    real acc = real(0);
    for (int t = 0; t < ntap; ++t) {
      int o = lo + t;
      acc = Math<real>::fma(w[t], in[(a + SA * o) * pitch + (b + SB * o)], acc);
    }
    out[a * pitch + b] = acc;
  }
}

// Fused normal-equations stencil on one tile: G = H^T mask(H X - B) where X is the staged
// region (tile extended by e on each side, local origin = (i0-e, j0-e)), B an optional
// global image (data term), mask = zero outside the image.  Uses two scratch buffers.
// On return T2 holds G on the tile rows/cols [e, e+TH) x [e, e+TW).
template <typename real>
__device__ __forceinline__ void normal_stencil(const Geom<real>& g, const real* X, real* T1, real* T2,
                                               const real* bimg, int i0, int j0) {
  const int e = g.e;
  const int hy = g.lo_y + g.ntap_y - 1, hx = g.lo_x + g.ntap_x - 1;  // upper end of the forward window
  // rows/cols of the residual needed by the adjoint on this tile
  const int ra0 = e - hy, ra1 = e + TH - g.lo_y;   // [ra0, ra1)
  const int rb0 = e - hx, rb1 = e + TW - g.lo_x;
  // 1. forward row pass on every staged row that the column pass will read
  const int xa0 = ra0 + g.lo_y, xa1 = ra1 + hy;    // = [e-hy+lo, e+TH-lo+hy) within [0, TH+2e)
  blur_pass<real, 0, 1>(X, T1, RS, xa0, xa1, rb0, rb1, g.wx, g.ntap_x, g.lo_x);
  __syncthreads();
  // 2. forward column pass, minus data, masked to the image
  {
    const int cols = rb1 - rb0, n = (ra1 - ra0) * cols;
    for (int idx = threadIdx.x; idx < n; idx += NTHREADS) {
      int a = ra0 + idx / cols, b = rb0 + idx % cols;
      int i = i0 - e + a, j = j0 - e + b;
      real acc = real(0);
      if (i >= 0 && i < g.ny && j >= 0 && j < g.nx) {
        for (int t = 0; t < g.ntap_y; ++t) acc = Math<real>::fma(g.wy[t], T1[(a + g.lo_y + t) * RS + b], acc);
        if (bimg != nullptr) acc -= bimg[(long long)i * g.nx + j];
      }
      T2[a * RS + b] = acc;
    }
  }
  __syncthreads();
  // 3. adjoint row pass (tile columns only)
  blur_pass<real, 0, -1>(T2, T1, RS, ra0, ra1, e, e + TW, g.wx, g.ntap_x, g.lo_x);
  __syncthreads();
  // 4. adjoint column pass -> tile
  blur_pass<real, -1, 0>(T1, T2, RS, e, e + TH, e, e + TW, g.wy, g.ntap_y, g.lo_y);
  __syncthreads();
}

// ---- Philox noise for the image chains: 4 normals for the 4-pixel group `quad` of row `i`
template <typename real>
__device__ __forceinline__ void img_noise4(uint2 key, unsigned long long chain, unsigned long long iter,
                                           int i, int quad, int quads_per_row, real z[4]) {
  uint32_t c0 = (uint32_t)((unsigned long long)i * quads_per_row + quad);
  uint4 r = lmc_philox4x32_10(
      make_uint4(c0, (uint32_t)iter, (uint32_t)chain, ((uint32_t)(chain >> 32) & 0xffffu) |
                                                           (((uint32_t)(iter >> 32) & 0xffu) << 16) | (2u << 24)),
      key);
  Math<real>::normals(r.x, r.y, z[0], z[1]);
  Math<real>::normals(r.z, r.w, z[2], z[3]);
}

}  // namespace lmcimg
