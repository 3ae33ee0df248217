// 2-Wasserstein evaluation on the device: the quantity the reference's mixture drivers plot,
//   ot.emd2(a, b, ot.dist(Z1, Z[:k+1]))**.5      (lmc.py:386-416, lmc_laplace.py:426-451, prox_lmc.py)
// between K true samples and a chain prefix, both with uniform weights.  POT's network simplex is a sequential
// CPU algorithm; here the transport problem is solved by log-domain Sinkhorn iterations with epsilon scaling,
// the cost |x_i - y_j|^2 recomputed on the fly (no n x m matrix in memory), all arithmetic in fp64:
//   f_i <- -eps log sum_j b_j exp((g_j - C_ij)/eps),   g_j <- -eps log sum_i a_i exp((f_i - C_ij)/eps)
// (optionally over-relaxed: new = old + omega (update - old), omega in [1,2).  Measured: omega = 1.5 needs 1.2-2.3x
// fewer iterations on cold starts but failed to converge on warm-started prefix curves; the host default is 1.)
// and the reported cost is the primal value <C, P_eps> of the plan P_ij = a_i b_j exp((f_i + g_j - C_ij)/eps).
// It is an approximation of the exact EMD whose gap shrinks with eps (tests/test_w2_gpu.py pins it against the
// exact linear programme of oracle/w2_ref.py: |W2 - exact| < 1e-3 at the default eps).
// One warp per group of ROWS rows; lanes stride over the other point set, which stays in L1/L2.
#include "lmc_common.cuh"

namespace {

constexpr int W2_BT = 256;
constexpr int W2_NW = W2_BT / 32;
constexpr int W2_MAXD = 4;
constexpr int W2_UNROLL = 4;
constexpr double W2_NEG = -1.0e300;

struct Lse {
  double m, s;
  __device__ __forceinline__ void push(double v) {
    if (v > m) {
      s = s * exp(m - v) + 1.0;
      m = v;
    } else {
      // v - m <= 0 is formed in fp64 (the 1/eps factor magnifies rounding of the cost); the exponential itself only
      // needs the ~2e-6 relative accuracy of ex2.approx: its error reaches f as eps * 2e-6, far below tol * eps
      s += (double)__expf((float)(v - m));
    }
  }
  __device__ __forceinline__ void merge(double m2, double s2) {
    const double M = fmax(m, m2);
    s = s * exp(m - M) + s2 * exp(m2 - M);
    m = M;
  }
};

// out_i += omega (-eps (LSE_j((pot_j - |p_i - q_j|^2)/eps) + logw) - out_i),  i < np, j < nq;  err = max_i |change|
template <int D, int ROWS>
__global__ void __launch_bounds__(W2_BT) w2_half_step_kernel(int np, int nq, const double* __restrict__ p,
                                                             const double* __restrict__ q,
                                                             const double* __restrict__ pot, double* __restrict__ out,
                                                             double eps, double logw, double omega,
                                                             unsigned long long* __restrict__ err) {
  const int lane = threadIdx.x & 31;
  const int grp = blockIdx.x * W2_NW + (threadIdx.x >> 5);
  const int i0 = grp * ROWS;
  if (i0 >= np) return;
  double pi[ROWS][D];
  Lse acc[ROWS];
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
    const int i = min(i0 + r, np - 1);
#pragma unroll
    for (int k = 0; k < D; ++k) pi[r][k] = p[(size_t)i * D + k];
    acc[r].m = W2_NEG;
    acc[r].s = 0.0;
  }
  const double ieps = 1.0 / eps;
  // W2_UNROLL columns per trip, their loads issued together: the capture of the one-column loop showed 11 long-
  // scoreboard stall cycles per issued instruction (every trip waited for its own q_j, g_j)
  int j = lane;
  for (; j + 32 * (W2_UNROLL - 1) < nq; j += 32 * W2_UNROLL) {
    double qj[W2_UNROLL][D], gj[W2_UNROLL];
#pragma unroll
    for (int u = 0; u < W2_UNROLL; ++u) {
#pragma unroll
      for (int k = 0; k < D; ++k) qj[u][k] = q[(size_t)(j + 32 * u) * D + k];
      gj[u] = pot[j + 32 * u];
    }
#pragma unroll
    for (int u = 0; u < W2_UNROLL; ++u) {
#pragma unroll
      for (int r = 0; r < ROWS; ++r) {
        double c = 0.0;
#pragma unroll
        for (int k = 0; k < D; ++k) {
          const double d = pi[r][k] - qj[u][k];
          c = fma(d, d, c);
        }
        acc[r].push((gj[u] - c) * ieps);
      }
    }
  }
  for (; j < nq; j += 32) {
    double qj[D];
#pragma unroll
    for (int k = 0; k < D; ++k) qj[k] = q[(size_t)j * D + k];
    const double gj = pot[j];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      double c = 0.0;
#pragma unroll
      for (int k = 0; k < D; ++k) {
        const double d = pi[r][k] - qj[k];
        c = fma(d, d, c);
      }
      acc[r].push((gj - c) * ieps);
    }
  }
  double worst = 0.0;
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double m2 = __shfl_xor_sync(0xffffffffu, acc[r].m, o);
      const double s2 = __shfl_xor_sync(0xffffffffu, acc[r].s, o);
      acc[r].merge(m2, s2);
    }
    if (lane == 0 && i0 + r < np) {
      const double old = out[i0 + r];
      const double v = fma(omega, -eps * (acc[r].m + log(acc[r].s) + logw) - old, old);   // over-relaxed update
      worst = fmax(worst, fabs(v - old));
      out[i0 + r] = v;
    }
  }
  if (lane == 0 && err != nullptr) atomicMax(err, (unsigned long long)__double_as_longlong(worst));   // worst >= 0
}

// partial[2 b] = sum_ij C_ij P_ij,  partial[2 b + 1] = sum_ij P_ij over the rows of block b,
// with P_ij = exp((f_i + g_j - C_ij)/eps + logab); summed in block order by w2_finish_kernel (deterministic).
template <int D>
__global__ void __launch_bounds__(W2_BT) w2_cost_kernel(int np, int nq, const double* __restrict__ p,
                                                        const double* __restrict__ q, const double* __restrict__ f,
                                                        const double* __restrict__ g, double eps, double logab,
                                                        double* __restrict__ partial) {
  __shared__ double red[2][W2_NW];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double c_sum = 0.0, m_sum = 0.0;
  const double ieps = 1.0 / eps;
  for (int i = blockIdx.x * W2_NW + warp; i < np; i += gridDim.x * W2_NW) {
    double pi[D];
#pragma unroll
    for (int k = 0; k < D; ++k) pi[k] = p[(size_t)i * D + k];
    const double fi = f[i];
    for (int j = lane; j < nq; j += 32) {
      double c = 0.0;
#pragma unroll
      for (int k = 0; k < D; ++k) {
        const double d = pi[k] - q[(size_t)j * D + k];
        c = fma(d, d, c);
      }
      const double w = exp((fi + g[j] - c) * ieps + logab);
      c_sum = fma(c, w, c_sum);
      m_sum += w;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    c_sum += __shfl_xor_sync(0xffffffffu, c_sum, o);
    m_sum += __shfl_xor_sync(0xffffffffu, m_sum, o);
  }
  if (lane == 0) {
    red[0][warp] = c_sum;
    red[1][warp] = m_sum;
  }
  __syncthreads();
  if (threadIdx.x < 2) {
    double s = 0.0;
    for (int w = 0; w < W2_NW; ++w) s += red[threadIdx.x][w];
    partial[2 * blockIdx.x + threadIdx.x] = s;
  }
}

__global__ void w2_finish_kernel(int blocks, const double* __restrict__ partial, double* __restrict__ scal) {
  if (threadIdx.x < 2) {
    double s = 0.0;
    for (int b = 0; b < blocks; ++b) s += partial[2 * b + threadIdx.x];
    scal[1 + threadIdx.x] = s;
  }
}

// scal[3] = squared diagonal of the joint bounding box of the two point sets (an upper bound of max C_ij)
template <int D>
__global__ void __launch_bounds__(W2_BT) w2_extent_kernel(int n, int m, const double* __restrict__ x,
                                                          const double* __restrict__ y, double* __restrict__ scal) {
  __shared__ double lo[W2_NW][D], hi[W2_NW][D];
  double l[D], h[D];
#pragma unroll
  for (int k = 0; k < D; ++k) {
    l[k] = 1.0e300;
    h[k] = -1.0e300;
  }
  for (int i = threadIdx.x; i < n + m; i += W2_BT) {
    const double* pt = i < n ? x + (size_t)i * D : y + (size_t)(i - n) * D;
#pragma unroll
    for (int k = 0; k < D; ++k) {
      l[k] = fmin(l[k], pt[k]);
      h[k] = fmax(h[k], pt[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < D; ++k)
    for (int o = 16; o > 0; o >>= 1) {
      l[k] = fmin(l[k], __shfl_xor_sync(0xffffffffu, l[k], o));
      h[k] = fmax(h[k], __shfl_xor_sync(0xffffffffu, h[k], o));
    }
  if ((threadIdx.x & 31) == 0)
    for (int k = 0; k < D; ++k) {
      lo[threadIdx.x >> 5][k] = l[k];
      hi[threadIdx.x >> 5][k] = h[k];
    }
  __syncthreads();
  if (threadIdx.x == 0) {
    double diag = 0.0;
    for (int k = 0; k < D; ++k) {
      double a = lo[0][k], b = hi[0][k];
      for (int w = 1; w < W2_NW; ++w) {
        a = fmin(a, lo[w][k]);
        b = fmax(b, hi[w][k]);
      }
      diag += (b - a) * (b - a);
    }
    scal[3] = diag;
  }
}

constexpr int W2_COST_BLOCKS = 296;
constexpr int W2_NSCAL = 8;

template <int D>
int half_step(int np, int nq, const double* p, const double* q, const double* pot, double* out, double eps,
              double logw, double omega, unsigned long long* err, cudaStream_t st) {
  // one row per warp: measured on B200 at 5000^2 / 10000^2 pairs, rows per warp 1 -> 149 / 428 us per iteration,
  // 2 -> 172 / 422 us, 4 -> 242 / 489 us (fewer warps in flight outweigh the reuse of q_j, g_j)
  w2_half_step_kernel<D, 1><<<(np + W2_NW - 1) / W2_NW, W2_BT, 0, st>>>(np, nq, p, q, pot, out, eps, logw, omega, err);
  LMC_CUDA_OK(cudaGetLastError());
  return 0;
}

template <int D>
int run_sinkhorn(const double* x, int n, const double* y, int m, double eps_final, double eps_start,
                 double eps_scale, double tol, double omega, int max_iters, int check_every, int warm,
                 double* ws,
                 lmc_w2_result* out, cudaStream_t st) {
  double* f = ws;
  double* g = ws + n;
  double* scal = g + m;
  double* partial = scal + W2_NSCAL;
  unsigned long long* err = (unsigned long long*)scal;
  const double la = -log((double)n), lb = -log((double)m);
  if (!warm) LMC_CUDA_OK(cudaMemsetAsync(f, 0, sizeof(double) * (size_t)n, st));
  LMC_CUDA_OK(cudaMemsetAsync(g, 0, sizeof(double) * (size_t)m, st));
  double eps = eps_start;
  if (!(eps > 0.0)) {
    w2_extent_kernel<D><<<1, W2_BT, 0, st>>>(n, m, x, y, scal);
    LMC_CUDA_OK(cudaGetLastError());
    LMC_CUDA_OK(cudaMemcpyAsync(&eps, scal + 3, sizeof(double), cudaMemcpyDeviceToHost, st));
    LMC_CUDA_OK(cudaStreamSynchronize(st));
    if (!(eps > eps_final)) eps = eps_final;          // coincident points: nothing to anneal
    eps = fmax(eps * eps_scale, eps_final);
  }
  long long iters = 0;
  double last_err = 0.0;
  int converged = 1;
  for (;;) {
    int done = 0;
    for (int it = 1; it <= max_iters && !done; ++it) {
      const bool check = (it % check_every) == 0 || it == max_iters;
      // g_j <- -eps log sum_i a_i exp((f_i - C_ij)/eps), then f_i <- -eps log sum_j b_j exp((g_j - C_ij)/eps)
      if (half_step<D>(m, n, y, x, f, g, eps, la, omega, nullptr, st) != 0) return -2;
      if (check) LMC_CUDA_OK(cudaMemsetAsync(err, 0, sizeof(unsigned long long), st));
      if (half_step<D>(n, m, x, y, g, f, eps, lb, omega, check ? err : nullptr, st) != 0) return -2;
      ++iters;
      if (check) {
        LMC_CUDA_OK(cudaMemcpyAsync(&last_err, err, sizeof(double), cudaMemcpyDeviceToHost, st));
        LMC_CUDA_OK(cudaStreamSynchronize(st));
        done = last_err <= tol * eps * omega;
      }
    }
    if (eps <= eps_final) {
      converged = done;
      break;
    }
    eps = fmax(eps * eps_scale, eps_final);
  }
  // one plain (omega = 1) f update makes the row marginals exact; the primal value of that plan is the estimate
  if (half_step<D>(n, m, x, y, g, f, eps, lb, 1.0, nullptr, st) != 0) return -2;
  w2_cost_kernel<D><<<W2_COST_BLOCKS, W2_BT, 0, st>>>(n, m, x, y, f, g, eps, la + lb, partial);
  LMC_CUDA_OK(cudaGetLastError());
  w2_finish_kernel<<<1, 32, 0, st>>>(W2_COST_BLOCKS, partial, scal);
  LMC_CUDA_OK(cudaGetLastError());
  double cm[2];
  LMC_CUDA_OK(cudaMemcpyAsync(cm, scal + 1, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
  LMC_CUDA_OK(cudaStreamSynchronize(st));
  out->cost = cm[0];
  out->w2 = sqrt(fmax(cm[0], 0.0));
  out->mass = cm[1];
  out->eps = eps;
  out->err = last_err;
  out->iters = iters;
  out->converged = converged;
  return 0;
}

}  // namespace

extern "C" size_t lmc_w2_workspace_bytes(int64_t n, int64_t m) {
  return (n < 1 || m < 1) ? 0 : sizeof(double) * (size_t)(n + m + W2_NSCAL + 2 * W2_COST_BLOCKS);
}

extern "C" int lmc_w2_sinkhorn(const double* d_x, int64_t n, const double* d_y, int64_t m, int32_t dim,
                               double eps_final, double eps_start, double eps_scale, double tol, double omega,
                               int32_t max_iters, int32_t check_every, int32_t warm_start, void* d_workspace, size_t workspace_bytes,
                               lmc_w2_result* out, void* stream) {
  LMC_REQUIRE(out != nullptr, "lmc_w2_sinkhorn: out is required");
  LMC_REQUIRE(n >= 1 && m >= 1 && n < (1ll << 30) && m < (1ll << 30), "lmc_w2_sinkhorn: bad sizes %lld, %lld",
              (long long)n, (long long)m);
  LMC_REQUIRE(dim >= 1 && dim <= W2_MAXD, "lmc_w2_sinkhorn: dim must be in 1..%d, got %d", W2_MAXD, dim);
  LMC_REQUIRE(d_x && d_y, "lmc_w2_sinkhorn: d_x and d_y are required");
  LMC_REQUIRE(eps_final > 0.0 && eps_scale > 0.0 && eps_scale < 1.0 && tol > 0.0,
              "lmc_w2_sinkhorn: need eps_final > 0, 0 < eps_scale < 1, tol > 0");
  LMC_REQUIRE(omega >= 1.0 && omega < 2.0, "lmc_w2_sinkhorn: the relaxation factor omega must be in [1, 2)");
  LMC_REQUIRE(max_iters >= 1 && check_every >= 1, "lmc_w2_sinkhorn: max_iters and check_every must be >= 1");
  LMC_REQUIRE(!warm_start || eps_start > 0.0, "lmc_w2_sinkhorn: a warm start needs an explicit eps_start");
  LMC_REQUIRE(d_workspace && workspace_bytes >= lmc_w2_workspace_bytes(n, m),
              "lmc_w2_sinkhorn: workspace too small (%zu < %zu)", workspace_bytes, lmc_w2_workspace_bytes(n, m));
  cudaStream_t st = (cudaStream_t)stream;
  double* ws = (double*)d_workspace;
  switch (dim) {
    case 1: return run_sinkhorn<1>(d_x, (int)n, d_y, (int)m, eps_final, eps_start, eps_scale, tol, omega, max_iters, check_every, warm_start, ws, out, st);
    case 2: return run_sinkhorn<2>(d_x, (int)n, d_y, (int)m, eps_final, eps_start, eps_scale, tol, omega, max_iters, check_every, warm_start, ws, out, st);
    case 3: return run_sinkhorn<3>(d_x, (int)n, d_y, (int)m, eps_final, eps_start, eps_scale, tol, omega, max_iters, check_every, warm_start, ws, out, st);
    default: return run_sinkhorn<4>(d_x, (int)n, d_y, (int)m, eps_final, eps_start, eps_scale, tol, omega, max_iters, check_every, warm_start, ws, out, st);
  }
}
