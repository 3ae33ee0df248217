// ULPDA image chains (algs.py:295-474) with proxf = pyproximal.L2 (LSQR-solved prox, warm
// start; prox_lmc_deconv.py:101-103), proxg = L21 / L1 and A = pylops.Gradient
// (prox_lmc_deconv.py:98,116,119,455-464) as ONE persistent cooperative kernel.
//
// Per sampler iteration (gfirst = False order; gfirst = True rotates the dual phase first):
//   P1  ATy = G^T y; rhs = x - tau (ATy + z) + tau sigma H^T b; u = rhs - Op1 warm
//       (Op1 = I + tau sigma H^T H as ONE fused 4-pass separable stencil); ||rhs||^2, ||u||^2
//   S1  bnorm, beta                                 (scalar phases: one warp per chain,
//   P2  v = Op1 (u / beta); w = v; ||v||^2           fixed-order sums of per-tile partials)
//   S2  alfa, LSQR recurrences initialised
//   LSQR loop (scipy.sparse.linalg.lsqr, damp = 0), per chain until its own stopping test:
//     Ka  [deferred x += t1 w; w = v + t2 w; ||w||^2]  u = Op1 v - alfa u; ||u||^2
//     Sa  beta, anorm
//     Kb  v = Op1 u - beta v; ||v||^2
//     Sb  alfa, plane rotations, t1, t2, norms, istop tests (atol, btol, conlim, iter_lim)
//   Pf  x_lsqr += t1 w (last deferred update); warm = x_lsqr; x = x_lsqr + sqrt(2 tau) xi;
//       xhat = x + theta (x - xold); running mean / M2; optional sample
//   Pd  y = proxdual_{mu}(y + mu G xhat): projection onto the l2 (L21) / l-inf (L1) ball
// The normalisations u/beta, v/alfa are never materialised: each stored vector carries a
// pending scale that the next reader applies (saves one sweep per normalisation).
#include "img_ulpda_tile.cuh"
#include "img_band.cuh"   // band geometry for the host side (the band kernels live in img_ulpda_band{5,6,7}.cu)

namespace {

// A = the sampler; B = the TV pipeline of an ME-TV term (mode 3) or disabled (mode -1)
template <typename real>
__global__ void __launch_bounds__(NTHREADS) img_ulpda_kernel(const __grid_constant__ UlpdaArgs<real> A,
                                                             const __grid_constant__ MyulaArgs<real> B) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  real* smem = reinterpret_cast<real*>(smem_raw);
  __shared__ double red[NTHREADS / 32];
  cg::grid_group grid = cg::this_grid();
  phase_optb(A, smem);
  grid.sync();
  for (long long n = 0; n < A.N; ++n) {
    if (A.mode == 0 && A.gfirst) {
      phase_dual(A, n, smem);
      grid.sync();
    }
    if (A.nc.kind != 0) ncvx_prestep(A, B, 0, A.C, n, smem, red, grid);
    if (blockIdx.x == 0 && threadIdx.x == 0) *A.n_active = 0;
    phase_primal_init(A, smem, red);
    grid.sync();
    phase_s1(A, 0, A.C);
    grid.sync();
    phase_v0(A, smem, red);
    grid.sync();
    phase_s2(A, 0, A.C);
    grid.sync();
    while (*(volatile int*)A.n_active != 0) {
      grid.sync();   // everyone has read n_active before it is reset
      if (blockIdx.x == 0 && threadIdx.x == 0) *A.n_active = 0;
      phase_ka(A, smem, red);
      grid.sync();
      phase_sa(A, 0, A.C);
      grid.sync();
      phase_kb(A, smem, red);
      grid.sync();
      phase_sb(A, 0, A.C);
      grid.sync();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(A.counters + 2, (unsigned long long)A.C);
    phase_primal_final(A, n);
    grid.sync();
    if (A.mode == 0 && !A.gfirst) {
      phase_dual(A, n, smem);
      grid.sync();
    }
  }
}

}  // namespace

#include "img_host_ulpda.inl"
