// MAP branch of the reference driver (prox_lmc_deconv.py:138-445): pyproximal.optimization.primaldual.
// AdaptivePrimalDual(proxf, proxg, A = Gradient, tau, mu) -- the deterministic twin of ULPDA with the step-size
// adaptation of Goldstein et al.  One iteration is
//     x    = proxf.prox(x - tau A^T y, tau)                 [lmc_img_pd_arg + lmc_img_l2_prox]
//     y    = proxg.proxdual(y + mu (2 A x - A x_old), mu)   [lmc_img_pdhg_dual, below]
//     p, d = ||(x_old - x)/tau - (A^T y_old - A^T y)||, ||(y_old - y)/mu - (A x_old - A x)||   (l2 norms)
// followed by the host-side rule on (p, d).  The two kernels below fuse the stencils with the residual sums; the
// sums are fp64 per-block partials added in a fixed order (deterministic), as everywhere on the image path.
#include "img_common.cuh"

using namespace lmcimg;

namespace {

constexpr int PD_BT = 256;

template <typename real>
__device__ __forceinline__ void pd_project(int iso, real sigma, real& v0, real& v1) {
  if (iso) {   // L21.proxdual: v / max(1, |v| / sigma)  (img_proxdual_kernel of img_ops.cu)
    const real den = Math<real>::max(real(1), Math<real>::sqrt(v0 * v0 + v1 * v1) / sigma);
    v0 /= den;
    v1 /= den;
  } else {     // L1.proxdual: clip
    v0 = Math<real>::max(-sigma, Math<real>::min(sigma, v0));
    v1 = Math<real>::max(-sigma, Math<real>::min(sigma, v1));
  }
}

__device__ __forceinline__ void pd_block_partial(double acc, double* partial) {
  __shared__ double red[PD_BT / 32];
  acc = lmc_warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0;
#pragma unroll
    for (int i = 0; i < PD_BT / 32; ++i) s += red[i];
    partial[blockIdx.x] = s;
  }
}

// dual step: Ax <- G x (in place over Ax_old), y <- proxdual(y + mu (2 Ax - Ax_old)), partial sums of
// ((y_old - y)/mu - (Ax_old - Ax))^2 over both components
template <typename real>
__global__ void __launch_bounds__(PD_BT) img_pdhg_dual_kernel(int ny, int nx, real mu, real sigma, int iso,
                                                              const real* __restrict__ x, real* __restrict__ Ax,
                                                              real* __restrict__ y, double* __restrict__ partial) {
  const long long P = (long long)ny * nx;
  double acc = 0;
  for (long long o = (long long)blockIdx.x * PD_BT + threadIdx.x; o < P; o += (long long)gridDim.x * PD_BT) {
    const int i = (int)(o / nx), j = (int)(o - (long long)i * nx);
    const real xc = x[o];
    const real a0 = (i < ny - 1) ? x[o + nx] - xc : real(0);     // pylops.Gradient, forward, edge=False
    const real a1 = (j < nx - 1) ? x[o + 1] - xc : real(0);
    const real o0 = Ax[o], o1 = Ax[P + o], y0 = y[o], y1 = y[P + o];
    real v0 = y0 + mu * (real(2) * a0 - o0), v1 = y1 + mu * (real(2) * a1 - o1);
    pd_project(iso, sigma, v0, v1);
    const real r0 = (y0 - v0) / mu - (o0 - a0), r1 = (y1 - v1) / mu - (o1 - a1);
    acc += (double)r0 * (double)r0 + (double)r1 * (double)r1;
    Ax[o] = a0;
    Ax[P + o] = a1;
    y[o] = v0;
    y[P + o] = v1;
  }
  pd_block_partial(acc, partial);
}

// primal residual: ATy <- G^T y (in place over ATy_old), partial sums of ((x_old - x)/tau - (ATy_old - ATy))^2
template <typename real>
__global__ void __launch_bounds__(PD_BT) img_pdhg_primal_kernel(int ny, int nx, real tau, const real* __restrict__ xold,
                                                                const real* __restrict__ x, const real* __restrict__ y,
                                                                real* __restrict__ ATy, double* __restrict__ partial) {
  const long long P = (long long)ny * nx;
  const real* y0 = y;
  const real* y1 = y + P;
  double acc = 0;
  for (long long o = (long long)blockIdx.x * PD_BT + threadIdx.x; o < P; o += (long long)gridDim.x * PD_BT) {
    const int i = (int)(o / nx), j = (int)(o - (long long)i * nx);
    real t = real(0);    // accumulation order of pylops.Gradient.rmatvec (img_pd_arg_kernel)
    if (i < ny - 1) t -= y0[o];
    if (i >= 1) t += y0[o - nx];
    if (j < nx - 1) t -= y1[o];
    if (j >= 1) t += y1[o - 1];
    const real r = (xold[o] - x[o]) / tau - (ATy[o] - t);
    acc += (double)r * (double)r;
    ATy[o] = t;
  }
  pd_block_partial(acc, partial);
}

__global__ void img_pdhg_finish_kernel(const double* __restrict__ partial, int n, double* __restrict__ out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double s = 0;
    for (int i = 0; i < n; ++i) s += partial[i];
    *out = sqrt(s);
  }
}

inline int pd_grid(long long P) {
  long long g = (P + PD_BT - 1) / PD_BT;
  return (int)(g < 1 ? 1 : (g > 148 * 8 ? 148 * 8 : g));
}

}  // namespace

extern "C" size_t lmc_img_pdhg_workspace_bytes(void) { return (size_t)(148 * 8) * sizeof(double); }

extern "C" int lmc_img_pdhg_dual(int dtype, int32_t ny, int32_t nx, double mu, double g_sigma, int isotropic,
                                 const void* d_x, void* d_Ax, void* d_y, double* d_resid, void* d_workspace,
                                 size_t workspace_bytes, void* stream) {
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_pdhg_dual: bad dtype");
  LMC_REQUIRE(ny >= 1 && nx >= 1 && mu > 0 && g_sigma > 0, "lmc_img_pdhg_dual: bad arguments");
  LMC_REQUIRE(d_x && d_Ax && d_y && d_resid, "lmc_img_pdhg_dual: null buffer");
  LMC_REQUIRE(d_workspace != nullptr && workspace_bytes >= lmc_img_pdhg_workspace_bytes(),
              "lmc_img_pdhg_dual: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = pd_grid((long long)ny * nx);
  double* part = (double*)d_workspace;
  if (dtype == LMC_F64)
    img_pdhg_dual_kernel<double><<<grid, PD_BT, 0, st>>>(ny, nx, mu, g_sigma, isotropic, (const double*)d_x,
                                                        (double*)d_Ax, (double*)d_y, part);
  else
    img_pdhg_dual_kernel<float><<<grid, PD_BT, 0, st>>>(ny, nx, (float)mu, (float)g_sigma, isotropic,
                                                       (const float*)d_x, (float*)d_Ax, (float*)d_y, part);
  img_pdhg_finish_kernel<<<1, 32, 0, st>>>(part, grid, d_resid);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int lmc_img_pdhg_primal_resid(int dtype, int32_t ny, int32_t nx, double tau, const void* d_xold,
                                         const void* d_x, const void* d_y, void* d_ATy, double* d_resid,
                                         void* d_workspace, size_t workspace_bytes, void* stream) {
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_pdhg_primal_resid: bad dtype");
  LMC_REQUIRE(ny >= 1 && nx >= 1 && tau > 0, "lmc_img_pdhg_primal_resid: bad arguments");
  LMC_REQUIRE(d_xold && d_x && d_y && d_ATy && d_resid, "lmc_img_pdhg_primal_resid: null buffer");
  LMC_REQUIRE(d_workspace != nullptr && workspace_bytes >= lmc_img_pdhg_workspace_bytes(),
              "lmc_img_pdhg_primal_resid: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = pd_grid((long long)ny * nx);
  double* part = (double*)d_workspace;
  if (dtype == LMC_F64)
    img_pdhg_primal_kernel<double><<<grid, PD_BT, 0, st>>>(ny, nx, tau, (const double*)d_xold, (const double*)d_x,
                                                          (const double*)d_y, (double*)d_ATy, part);
  else
    img_pdhg_primal_kernel<float><<<grid, PD_BT, 0, st>>>(ny, nx, (float)tau, (const float*)d_xold, (const float*)d_x,
                                                         (const float*)d_y, (float*)d_ATy, part);
  img_pdhg_finish_kernel<<<1, 32, 0, st>>>(part, grid, d_resid);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}
