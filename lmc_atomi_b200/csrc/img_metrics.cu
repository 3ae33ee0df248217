// Per-sample quality metrics on the device: what the reference's per-iteration host callback
// computes (prox_lmc_deconv.py:35-36,128-133: cost f(x) + g(Kx), ||x - x*||, SNR, PSNR, MSE) reduced to four
// sums per image so that a run can return N x 4 doubles instead of copying N x P samples to the host:
//   [0] ||H x - b||^2   [1] sum_pixels |G x|  (isotropic: l2 norm of the two differences; else l1)
//   [2] ||x - x*||^2    [3] ||x*||^2
// One work item = (image, slice of its tiles); fp64 accumulation, fixed-order reductions: deterministic.
#include <mutex>

#include "img_common.cuh"

using namespace lmcimg;

#include "img_host_common.inl"

namespace {

constexpr int MET_SLICES = 16;   // partial sums per image
constexpr int MET_NQ = 4;

template <typename real>
__global__ void __launch_bounds__(NTHREADS) img_metrics_kernel(const __grid_constant__ Geom<real> g, int isotropic,
                                                               long long M, long long imgs_per_chain,
                                                               const real* __restrict__ x, const real* __restrict__ b,
                                                               long long b_stride, const real* __restrict__ xtrue,
                                                               double* __restrict__ partial) {
  __shared__ real X[RS * RS];
  __shared__ real T1[RS * RS];
  __shared__ double red[NTHREADS / 32];
  const int e = g.e, hy = g.lo_y + g.ntap_y - 1;
  const long long items = M * MET_SLICES;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    const long long m = w / MET_SLICES;
    const int sl = (int)(w - m * MET_SLICES);
    const real* xi = x + m * g.P;
    const real* bi = b + (m / imgs_per_chain) * b_stride;
    double acc[MET_NQ] = {0, 0, 0, 0};
    for (int tile = sl; tile < g.tiles; tile += MET_SLICES) {
      const int i0 = (tile / g.tiles_x) * TH, j0 = (tile % g.tiles_x) * TW;
      __syncthreads();
      load_region(X, RS, xi, g.ny, g.nx, i0 - e, j0 - e, TH + 2 * e, TW + 2 * e);
      __syncthreads();
      blur_pass<real, 0, 1>(X, T1, RS, e + g.lo_y, e + TH + hy, e, e + TW, g.wx, g.ntap_x, g.lo_x);
      __syncthreads();
      for (int idx = threadIdx.x; idx < TH * TW; idx += NTHREADS) {
        const int a = idx / TW, c = idx % TW, i = i0 + a, j = j0 + c;
        if (i >= g.ny || j >= g.nx) continue;
        real hx = real(0);
        for (int t = 0; t < g.ntap_y; ++t) hx = Math<real>::fma(g.wy[t], T1[(e + a + g.lo_y + t) * RS + e + c], hx);
        const long long o = (long long)i * g.nx + j;
        const real xv = xi[o];
        const real r = hx - bi[o];
        const real d0 = (i < g.ny - 1) ? xi[o + g.nx] - xv : real(0);
        const real d1 = (j < g.nx - 1) ? xi[o + 1] - xv : real(0);
        acc[0] += (double)r * (double)r;
        acc[1] += isotropic ? (double)Math<real>::sqrt(d0 * d0 + d1 * d1)
                            : (double)Math<real>::abs(d0) + (double)Math<real>::abs(d1);
        if (xtrue != nullptr) {
          const real xt = xtrue[o], dd = xv - xt;
          acc[2] += (double)dd * (double)dd;
          acc[3] += (double)xt * (double)xt;
        }
      }
    }
#pragma unroll
    for (int q = 0; q < MET_NQ; ++q) {
      const double s = block_sum(acc[q], red);
      if (threadIdx.x == 0) partial[(m * MET_SLICES + sl) * MET_NQ + q] = s;
    }
  }
}

__global__ void img_metrics_reduce_kernel(long long M, const double* __restrict__ partial, double* __restrict__ out) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= M * MET_NQ) return;
  const long long m = t / MET_NQ;
  const int q = (int)(t - m * MET_NQ);
  double s = 0;
  for (int sl = 0; sl < MET_SLICES; ++sl) s += partial[(m * MET_SLICES + sl) * MET_NQ + q];
  out[t] = s;
}

template <typename real>
int run_metrics(const lmc_img_problem* pb, int isotropic, long long M, long long per_chain, const void* d_x,
                const void* d_xtrue, double* d_out, void* ws, cudaStream_t st) {
  Geom<real> g;
  fill_geom(pb, g);
  const long long items = M * MET_SLICES;
  const int grid = (int)(items < 148 * 8 ? items : 148 * 8);
  img_metrics_kernel<real><<<grid, NTHREADS, 0, st>>>(g, isotropic, M, per_chain, (const real*)d_x,
                                                     (const real*)pb->d_b, pb->b_chain_stride, (const real*)d_xtrue,
                                                     (double*)ws);
  LMC_CUDA_OK(cudaGetLastError());
  const long long n = M * MET_NQ;
  img_metrics_reduce_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>(M, (const double*)ws, d_out);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

extern "C" size_t lmc_img_metrics_workspace_bytes(int64_t M) {
  return M < 1 ? 0 : (size_t)M * MET_SLICES * MET_NQ * sizeof(double);
}

extern "C" int lmc_img_sample_metrics(const lmc_img_problem* pb, int dtype, int isotropic, int64_t M,
                                      int64_t images_per_chain, const void* d_x, const void* d_xtrue, double* d_out,
                                      void* d_workspace, size_t workspace_bytes, void* stream) {
  if (check_problem(pb, "lmc_img_sample_metrics") != 0) return -1;
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_sample_metrics: bad dtype");
  if (M <= 0) return 0;
  LMC_REQUIRE(images_per_chain >= 1, "lmc_img_sample_metrics: images_per_chain must be >= 1");
  LMC_REQUIRE(d_x && d_out && pb->d_b, "lmc_img_sample_metrics: d_x, d_out and pb->d_b are required");
  LMC_REQUIRE(pb->b_chain_stride == 0 || pb->b_chain_stride == (int64_t)pb->ny * pb->nx,
              "lmc_img_sample_metrics: b_chain_stride must be 0 or ny*nx");
  LMC_REQUIRE(d_workspace && workspace_bytes >= lmc_img_metrics_workspace_bytes(M),
              "lmc_img_sample_metrics: workspace too small (%zu < %zu)", workspace_bytes,
              lmc_img_metrics_workspace_bytes(M));
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == LMC_F64
             ? run_metrics<double>(pb, isotropic, M, images_per_chain, d_x, d_xtrue, d_out, d_workspace, st)
             : run_metrics<float>(pb, isotropic, M, images_per_chain, d_x, d_xtrue, d_out, d_workspace, st);
}

// ------------------------------------------------------------------ no-allocation variant
// Same kernels, same arguments and results as lmc_img_sample_metrics, but the per-slice partial sums live in a
// file-scope device array instead of a caller-supplied workspace: for callers that cannot allocate (CUDA graph
// capture, serving loops).  Every GPU that loads the library holds LMC_METRICS_STATIC_MAX_IMAGES * 16 * 4 doubles
// (512 KiB) until exit.  `metrics_scratch` is a host-side stand-in here, NOT a device address: the device address
// is looked up once per device with cudaGetSymbolAddress, with that device current, and cached.
// Every (image, slice, quantity) partial is written by img_metrics_kernel before img_metrics_reduce_kernel reads it
// in the same call, so nothing of the previous call's contents has to be cleared.  All calls on a device share the
// array: two calls that overlap on the GPU (different streams, no dependency) race.
static __device__ __align__(256) double metrics_scratch[(size_t)LMC_METRICS_STATIC_MAX_IMAGES * MET_SLICES * MET_NQ];

namespace {
constexpr int MET_MAX_DEVICES = 64;
std::mutex metrics_scratch_mu;
double* metrics_scratch_ptr[MET_MAX_DEVICES] = {};
}  // namespace

extern "C" int lmc_img_sample_metrics_static(const lmc_img_problem* pb, int dtype, int isotropic, int64_t M,
                                             int64_t images_per_chain, const void* d_x, const void* d_xtrue,
                                             double* d_out, void* stream) {
  if (check_problem(pb, "lmc_img_sample_metrics_static") != 0) return -1;
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_sample_metrics_static: bad dtype");
  if (M <= 0) return 0;
  LMC_REQUIRE(M <= LMC_METRICS_STATIC_MAX_IMAGES,
              "lmc_img_sample_metrics_static: M = %lld images exceeds the static scratch (max %d); split the call or "
              "use lmc_img_sample_metrics with a workspace", (long long)M, LMC_METRICS_STATIC_MAX_IMAGES);
  LMC_REQUIRE(images_per_chain >= 1, "lmc_img_sample_metrics_static: images_per_chain must be >= 1");
  LMC_REQUIRE(d_x && d_out && pb->d_b, "lmc_img_sample_metrics_static: d_x, d_out and pb->d_b are required");
  LMC_REQUIRE(pb->b_chain_stride == 0 || pb->b_chain_stride == (int64_t)pb->ny * pb->nx,
              "lmc_img_sample_metrics_static: b_chain_stride must be 0 or ny*nx");
  int dev = -1;
  LMC_CUDA_OK(cudaGetDevice(&dev));
  LMC_REQUIRE(dev >= 0 && dev < MET_MAX_DEVICES, "lmc_img_sample_metrics_static: device ordinal %d out of range", dev);
  double* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(metrics_scratch_mu);
    if (metrics_scratch_ptr[dev] == nullptr) {   // once per device, with that device current
      void* p = nullptr;
      LMC_CUDA_OK(cudaGetSymbolAddress(&p, metrics_scratch));
      LMC_REQUIRE(p != nullptr, "lmc_img_sample_metrics_static: cudaGetSymbolAddress returned a null address");
      metrics_scratch_ptr[dev] = (double*)p;
    }
    scratch = metrics_scratch_ptr[dev];
  }
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == LMC_F64
             ? run_metrics<double>(pb, isotropic, M, images_per_chain, d_x, d_xtrue, d_out, scratch, st)
             : run_metrics<float>(pb, isotropic, M, images_per_chain, d_x, d_xtrue, d_out, scratch, st);
}
