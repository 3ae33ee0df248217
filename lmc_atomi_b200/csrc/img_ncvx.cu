// Pieces of the reference's non-convex data term algs.L2_ncvx_tv (algs.py:22-291):
// f(x) = sigma/2 ||Hx - b||^2 - lamda * env_gamma(TV)(x).  The samplers need the gradient of the Moreau
// envelope, which is either a 7-point stencil (MC-TV, Op2 = Gradient) or a TV prox followed by a pointwise
// combination (ME-TV); both run between the single-iteration launches of the sampler kernels, which take the
// result through lmc_img_io::d_fgrad_extra (MYULA) / d_prox_arg (ULPDA).  HBM-bound pointwise/stencil work:
// one thread per pixel, neighbours served by L1/L2.
#include "img_common.cuh"
#include "img_ncvx.cuh"

using namespace lmcimg;

namespace {

// out = base * x + scale * (G^T(w(Gx) Gx) / post_div); G = forward differences, last row/col zero
template <typename real>
__global__ void img_mctv_grad_kernel(int ny, int nx, MctvW<real> W, real base, real scale, real post_div, long long C,
                                     const real* __restrict__ xin, real* __restrict__ out) {
  const long long P = (long long)ny * nx, total = C * P;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const long long chain = t / P, o = t - chain * P;
    const int i = (int)(o / nx), j = (int)(o - (long long)i * nx);
    const real* x = xin + chain * P;
    const real xc = x[o];
    const real acc = mctv_at(W, i, j, ny, nx, [&](int di, int dj) { return x[o + di * nx + dj]; });
    const real m = scale * (acc / post_div);
    out[t] = (base != real(0)) ? base * xc + m : m;
  }
}

template <typename real>
__global__ void img_metv_combine_kernel(long long n, real base, real scale, real den, const real* __restrict__ x,
                                        const real* __restrict__ p, real* __restrict__ out) {
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
    const real xv = x[t], d = xv - p[t];
    const real f = (den != real(0)) ? d / den : scale * d;
    out[t] = (base != real(0)) ? base * xv + f : f;
  }
}

// out = x - tau * (G^T y + z)   (algs.py:436-446); y is (C,2,P), z (P,) shared or null
template <typename real>
__global__ void img_pd_arg_kernel(int ny, int nx, real tau, long long C, const real* __restrict__ xin,
                                  const real* __restrict__ yin, const real* __restrict__ z, real* __restrict__ out) {
  const long long P = (long long)ny * nx, total = C * P;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const long long chain = t / P, o = t - chain * P;
    const int i = (int)(o / nx), j = (int)(o - (long long)i * nx);
    const real* y0 = yin + chain * 2 * P;
    const real* y1 = y0 + P;
    real acc = real(0);
    if (i < ny - 1) acc -= y0[o];
    if (i >= 1) acc += y0[o - nx];
    if (j < nx - 1) acc -= y1[o];
    if (j >= 1) acc += y1[o - 1];
    if (z != nullptr) acc += z[o];
    out[t] = xin[t] - tau * acc;
  }
}

inline int ncvx_grid(long long n) {
  long long g = (n + 255) / 256;
  return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g));
}

}  // namespace

extern "C" int lmc_img_mctv_grad(int dtype, int32_t ny, int32_t nx, double gamma, int isotropic, double base,
                                 double scale, double post_div, int64_t C, const void* d_x, void* d_out,
                                 void* stream) {
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_mctv_grad: bad dtype");
  LMC_REQUIRE(ny >= 1 && nx >= 1 && gamma > 0 && post_div != 0, "lmc_img_mctv_grad: bad arguments");
  if (C <= 0) return 0;
  LMC_REQUIRE(d_x && d_out && d_x != d_out, "lmc_img_mctv_grad: needs distinct in/out buffers");
  cudaStream_t st = (cudaStream_t)stream;
  const long long n = (long long)C * ny * nx;
  if (dtype == LMC_F64) {
    MctvW<double> W{1.0 / gamma, gamma, isotropic};
    img_mctv_grad_kernel<double><<<ncvx_grid(n), 256, 0, st>>>(ny, nx, W, base, scale, post_div, C, (const double*)d_x,
                                                              (double*)d_out);
  } else {
    MctvW<float> W{(float)(1.0 / gamma), (float)gamma, isotropic};
    img_mctv_grad_kernel<float><<<ncvx_grid(n), 256, 0, st>>>(ny, nx, W, (float)base, (float)scale, (float)post_div, C,
                                                             (const float*)d_x, (float*)d_out);
  }
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int lmc_img_metv_combine(int dtype, int64_t n, double base, double scale, double den, const void* d_x,
                                    const void* d_p, void* d_out, void* stream) {
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_metv_combine: bad dtype");
  if (n <= 0) return 0;
  LMC_REQUIRE(d_x && d_p && d_out, "lmc_img_metv_combine: null buffer");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == LMC_F64)
    img_metv_combine_kernel<double><<<ncvx_grid(n), 256, 0, st>>>(n, base, scale, den, (const double*)d_x,
                                                                 (const double*)d_p, (double*)d_out);
  else
    img_metv_combine_kernel<float><<<ncvx_grid(n), 256, 0, st>>>(n, (float)base, (float)scale, (float)den,
                                                                (const float*)d_x, (const float*)d_p, (float*)d_out);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int lmc_img_pd_arg(int dtype, int32_t ny, int32_t nx, double tau, int64_t C, const void* d_x,
                              const void* d_y, const void* d_z, void* d_out, void* stream) {
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_pd_arg: bad dtype");
  LMC_REQUIRE(ny >= 1 && nx >= 1, "lmc_img_pd_arg: bad size");
  if (C <= 0) return 0;
  LMC_REQUIRE(d_x && d_y && d_out, "lmc_img_pd_arg: null buffer");
  cudaStream_t st = (cudaStream_t)stream;
  const long long n = (long long)C * ny * nx;
  if (dtype == LMC_F64)
    img_pd_arg_kernel<double><<<ncvx_grid(n), 256, 0, st>>>(ny, nx, tau, C, (const double*)d_x, (const double*)d_y,
                                                           (const double*)d_z, (double*)d_out);
  else
    img_pd_arg_kernel<float><<<ncvx_grid(n), 256, 0, st>>>(ny, nx, (float)tau, C, (const float*)d_x,
                                                          (const float*)d_y, (const float*)d_z, (float*)d_out);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}
