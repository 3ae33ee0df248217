// Stand-alone batched operators of the image path (unit entry points; also what the
// host-side operator stand-ins call): Convolve2D matvec/rmatvec, Gradient matvec/rmatvec,
// L21/L1 dual prox, and the Philox normals the samplers draw (test hook).
#include "img_common.cuh"

using namespace lmcimg;

#include "img_host_common.inl"

namespace {

template <typename real>
__global__ void __launch_bounds__(NTHREADS) img_blur_kernel(const __grid_constant__ Geom<real> g, int adjoint,
                                                            long long C, const real* in, real* out) {
  __shared__ real X[RS * RS];
  __shared__ real T1[RS * RS];
  const long long items = C * g.tiles;
  const int e = g.e, hy = g.lo_y + g.ntap_y - 1;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    long long chain = w / g.tiles;
    int tile = (int)(w - chain * g.tiles);
    int i0 = (tile / g.tiles_x) * TH, j0 = (tile % g.tiles_x) * TW;
    __syncthreads();
    load_region(X, RS, in + chain * g.P, g.ny, g.nx, i0 - e, j0 - e, TH + 2 * e, TW + 2 * e);
    __syncthreads();
    if (!adjoint) blur_pass<real, 0, 1>(X, T1, RS, e + g.lo_y, e + TH + hy, e, e + TW, g.wx, g.ntap_x, g.lo_x);
    else blur_pass<real, 0, -1>(X, T1, RS, e - hy, e + TH - g.lo_y, e, e + TW, g.wx, g.ntap_x, g.lo_x);
    __syncthreads();
    if (!adjoint) blur_pass<real, 1, 0>(T1, X, RS, e, e + TH, e, e + TW, g.wy, g.ntap_y, g.lo_y);
    else blur_pass<real, -1, 0>(T1, X, RS, e, e + TH, e, e + TW, g.wy, g.ntap_y, g.lo_y);
    __syncthreads();
    for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
      int a = idx / TW, b = idx % TW, i = i0 + a, j = j0 + b;
      if (i < g.ny && j < g.nx) out[chain * g.P + (long long)i * g.nx + j] = X[(e + a) * RS + e + b];
    }
  }
}

// forward: in (C,P) -> out (C,2,P); adjoint: in (C,2,P) -> out (C,P)   
template <typename real>
__global__ void img_grad_kernel(int ny, int nx, int adjoint, long long C, const real* in, real* out) {
  const long long P = (long long)ny * nx, total = C * P;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    long long chain = t / P, o = t - chain * P;
    int i = (int)(o / nx), j = (int)(o - (long long)i * nx);
    if (!adjoint) {
      const real* x = in + chain * P;
      real xv = x[o];
      out[chain * 2 * P + o] = (i < ny - 1) ? x[o + nx] - xv : real(0);
      out[chain * 2 * P + P + o] = (j < nx - 1) ? x[o + 1] - xv : real(0);
    } else {
      const real* v0 = in + chain * 2 * P;
      const real* v1 = v0 + P;
      real acc = real(0);
      if (i < ny - 1) acc -= v0[o];
      if (i >= 1) acc += v0[o - nx];
      if (j < nx - 1) acc -= v1[o];
      if (j >= 1) acc += v1[o - 1];
      out[chain * P + o] = acc;
    }
  }
}

template <typename real>
__global__ void img_proxdual_kernel(int isotropic, real sigma, long long n, real* v) {
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n;
       t += (long long)gridDim.x * blockDim.x) {
    real v0 = v[t], v1 = v[n + t];
    if (isotropic) {
      real den = Math<real>::max(real(1), Math<real>::sqrt(v0 * v0 + v1 * v1) / sigma);
      v0 /= den;
      v1 /= den;
    } else {
      v0 = Math<real>::max(-sigma, Math<real>::min(sigma, v0));
      v1 = Math<real>::max(-sigma, Math<real>::min(sigma, v1));
    }
    v[t] = v0;
    v[n + t] = v1;
  }
}

template <typename real>
__global__ void img_noise_kernel(int ny, int nx, long long C, long long N, uint2 key, unsigned long long chain_offset,
                                 unsigned long long iter_offset, real* out) {
  const int qpr = (nx + 3) / 4;
  const long long per = (long long)ny * qpr, total = C * N * per;
  const long long P = (long long)ny * nx;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    long long cn = t / per, r = t - cn * per;
    long long chain = cn / N, n = cn - chain * N;
    int i = (int)(r / qpr), q = (int)(r - (long long)i * qpr);
    real z[4];
    img_noise4<real>(key, chain_offset + (unsigned long long)chain, iter_offset + (unsigned long long)n, i, q, qpr, z);
    for (int k = 0; k < 4; ++k)
      if (4 * q + k < nx) out[cn * P + (long long)i * nx + 4 * q + k] = z[k];
  }
}

inline int grid_for(long long n, int block) {
  long long g = (n + block - 1) / block;
  return (int)(g < 1 ? 1 : (g > 148 * 32 ? 148 * 32 : g));
}

}  // namespace

extern "C" int lmc_img_blur(const lmc_img_problem* pb, int dtype, int adjoint, int64_t C, const void* d_in,
                            void* d_out, void* stream) {
  if (check_problem(pb, "lmc_img_blur") != 0) return -1;
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_blur: bad dtype");
  if (C <= 0) return 0;
  LMC_REQUIRE(d_in && d_out && d_in != d_out, "lmc_img_blur: needs distinct in/out buffers");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == LMC_F64) {
    Geom<double> g;
    fill_geom(pb, g);
    long long items = C * g.tiles;
    img_blur_kernel<double><<<(int)(items < 148 * 8 ? items : 148 * 8), NTHREADS, 0, st>>>(
        g, adjoint, C, (const double*)d_in, (double*)d_out);
  } else {
    Geom<float> g;
    fill_geom(pb, g);
    long long items = C * g.tiles;
    img_blur_kernel<float><<<(int)(items < 148 * 8 ? items : 148 * 8), NTHREADS, 0, st>>>(
        g, adjoint, C, (const float*)d_in, (float*)d_out);
  }
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int lmc_img_grad(int dtype, int32_t ny, int32_t nx, int adjoint, int64_t C, const void* d_in, void* d_out,
                            void* stream) {
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_grad: bad dtype");
  LMC_REQUIRE(ny >= 1 && nx >= 1, "lmc_img_grad: bad size");
  if (C <= 0) return 0;
  LMC_REQUIRE(d_in && d_out, "lmc_img_grad: null buffer");
  cudaStream_t st = (cudaStream_t)stream;
  long long n = (long long)C * ny * nx;
  if (dtype == LMC_F64)
    img_grad_kernel<double><<<grid_for(n, 256), 256, 0, st>>>(ny, nx, adjoint, C, (const double*)d_in, (double*)d_out);
  else
    img_grad_kernel<float><<<grid_for(n, 256), 256, 0, st>>>(ny, nx, adjoint, C, (const float*)d_in, (float*)d_out);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int lmc_img_proxdual(int dtype, int isotropic, double sigma, int64_t n_pix, void* d_v, void* stream) {
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_proxdual: bad dtype");
  LMC_REQUIRE(sigma > 0, "lmc_img_proxdual: sigma must be positive");
  if (n_pix <= 0) return 0;
  LMC_REQUIRE(d_v != nullptr, "lmc_img_proxdual: null buffer");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == LMC_F64)
    img_proxdual_kernel<double><<<grid_for(n_pix, 256), 256, 0, st>>>(isotropic, sigma, n_pix, (double*)d_v);
  else
    img_proxdual_kernel<float><<<grid_for(n_pix, 256), 256, 0, st>>>(isotropic, (float)sigma, n_pix, (float*)d_v);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int lmc_img_noise(int dtype, int32_t ny, int32_t nx, int64_t C, int64_t N, uint64_t seed,
                             uint64_t chain_offset, uint64_t iter_offset, void* d_out, void* stream) {
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_noise: bad dtype");
  LMC_REQUIRE(ny >= 1 && nx >= 1, "lmc_img_noise: bad size");
  if (C <= 0 || N <= 0) return 0;
  LMC_REQUIRE(d_out != nullptr, "lmc_img_noise: null buffer");
  cudaStream_t st = (cudaStream_t)stream;
  long long n = (long long)C * N * ny * ((nx + 3) / 4);
  if (dtype == LMC_F64)
    img_noise_kernel<double><<<grid_for(n, 256), 256, 0, st>>>(ny, nx, C, N, img_key(seed), chain_offset, iter_offset,
                                                              (double*)d_out);
  else
    img_noise_kernel<float><<<grid_for(n, 256), 256, 0, st>>>(ny, nx, C, N, img_key(seed), chain_offset, iter_offset,
                                                             (float*)d_out);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}
