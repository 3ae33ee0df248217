// MYULA image chains (algs.py:477-587) with the L2 data term and the isotropic TV prior
// (prox_lmc_deconv.py:101-103,122,465-473) as ONE persistent cooperative kernel.
//
// Per sampler iteration (all chains of the batch advance together):
//   phase G  : g = sigma_L2 * H^T (H x - b)   [fused 4-pass separable stencil, halo 2(k-1)]
//              + TV dual sweep 0 (sol_0 = x)  [same staged tile of x]
//   phase S  : per chain: obj_k from the per-tile partials (fixed order), relative-change
//              test of pyproximal.TV.prox, early exit per chain
//   phase T_k: k = 1..niter-1 for chains still active: rr = p_k + c_k (p_k - p_{k-1}),
//              sol_k = x - gamma div(rr), obj_k partials, speculative p_{k+1}
//   phase U  : prox = sol_kf recomputed from (p_kf, p_kf-1); x <- (1-tau/gam) x - tau g
//              + tau/gam prox + sqrt(2 tau) xi; running mean / M2 (Welford); optional sample
// Phases are separated by grid-wide barriers; a chain's tiles are consecutive work items,
// so halos are served by L2.  The dual variables are stored as the projected iterates
// p_k only (3 rotating buffers): 7 array transfers per executed sweep instead of the 9 of
// a literal (rr, ss, pold, qold) implementation.
#include "img_myula_tile.cuh"

namespace {

// A = the sampler; B = the ME-TV prelude (mode 2: every iteration first evaluates (x - TV_1.prox(x, gamma)) / gamma
// into A.gextra with the same TV machinery, algs.py:280-282) or disabled (mode -1)
template <typename real>
__global__ void __launch_bounds__(NTHREADS) img_myula_kernel(const __grid_constant__ MyulaArgs<real> A,
                                                             const __grid_constant__ MyulaArgs<real> B) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  real* smem = reinterpret_cast<real*>(smem_raw);
  __shared__ double red[NTHREADS / 32];
  cg::grid_group grid = cg::this_grid();
  for (long long n = 0; n < A.N; ++n) {
    if (B.mode == 2) myula_tile_iteration(B, n, smem, red, grid);
    myula_tile_iteration(A, n, smem, red, grid);
  }
}

}  // namespace

#include "img_myula_band.inl"
#include "img_host.inl"
