// MYULA image chains (algs.py:477-587; L2 data term + isotropic TV prior,
// prox_lmc_deconv.py:101-103,122,465-473) with the WHOLE CHAIN STATE ON CHIP.
//
// One thread-block cluster of 8 CTAs owns one chain for all N sampler iterations:
// CTA `rank` holds image rows [rank*RPC, (rank+1)*RPC) (RPC = 32 -> images up to
// 256 x 256, fp32).  Per CTA, in shared memory (row pitch 272 floats, 8 zero pad
// columns each side = the zero boundary of pylops Convolve2D):
//   XS   : x on the CTA's rows plus 6 halo rows above and below (blur halo)
//   pool : two (rr, ss) buffers of the TV dual field, RPC + 2 halo rows each; the
//          second pair doubles as the two scratch images of the blur passes
// and in registers (thread = one column x a strip of RPC/2 rows): the pointwise
// arrays p, q (current projected dual iterate) and `base` (everything of the MYULA
// update that does not depend on the TV prox).  Halo rows are PUSHED into the
// neighbour CTA's shared memory (DSMEM) by the thread that produces them; the
// per-chain TV objective is reduced across the cluster through 8 doubles per CTA.
// One cluster barrier per TV sweep + one per sampler iteration; chains exit the
// TV loop individually (no batch-wide lockstep).  HBM traffic per iteration: the
// observation b (L2 resident when shared), the running mean / samples if requested.
//
// Arithmetic is the same recipe as the streaming kernel img_myula.cu (pyproximal
// TV.prox as restated in SURVEY 8a I4) with two fp32 shortcuts that stay far inside
// the 1e-5 trajectory tolerance: dx/(8 gamma) is a multiply by the host-rounded
// reciprocal, and r / max(1,|(r,s)|) is r * min(1, rsqrt(r^2+s^2)).
#include <cooperative_groups.h>

#include "img_common.cuh"
#include "img_host_common.inl"

namespace cg = cooperative_groups;
using namespace lmcimg;

namespace {

constexpr int OC_CL = 8;               // CTAs per cluster
constexpr int OC_W = 256;              // columns handled (nx <= OC_W)
constexpr int OC_PADL = 8;             // zero columns left of column 0
constexpr int OC_PITCH = OC_W + 16;    // floats per shared-memory row
constexpr int OC_HB = 6;               // x halo rows per side (max blur taps - 1)
constexpr int OC_NT = 512;             // threads per CTA: 256 columns x 2 row strips

template <int RPC>
struct OcLayout {
  static constexpr int STRIP = RPC / 2;
  static constexpr int XS_ROWS = RPC + 2 * OC_HB;
  static constexpr int DROWS = RPC + 2;             // halo row above, own rows, halo row below
  static constexpr int T_ROWS = RPC + 2 * OC_HB;    // blur scratch images
  // pool rows.  Buffer Y (odd TV sweeps write it) aliases the blur scratch; buffer X
  // (written by sweep 0, which runs before the blur of the same iteration) does not.
  static constexpr int Y_RR = 0, Y_SS = DROWS;
  static constexpr int T1 = 0, T2 = T_ROWS;
  static constexpr int X_RR = (2 * DROWS > 2 * T_ROWS) ? 2 * DROWS : 2 * T_ROWS;
  static constexpr int X_SS = X_RR + DROWS;
  static constexpr int POOL_ROWS = X_SS + DROWS;
  static constexpr size_t SMEM_BYTES = (size_t)(XS_ROWS + POOL_ROWS) * OC_PITCH * sizeof(float);
  static_assert(STRIP % 4 == 0, "strips are processed in groups of 4 rows (one Philox block per lane)");
  static_assert(RPC >= OC_HB, "halo rows must come from the adjacent CTA only");
};

struct OcArgs {
  Geom<float> g;
  float* x;
  const float* b;
  long long b_stride;
  const float* noise;   // [C][N][P] or null
  float* samples;       // [C][N][P] or null
  float* mean;
  float* m2;
  unsigned long long* counters;
  long long C, N, burnin, stat_count0;
  unsigned long long chain_offset, iter_offset;
  uint2 key;
  float sigma_l2, tau, a_x, tog, s2t, tv_gamma, inv_g8;
  int tv_niter;
  double tv_rtol;
};

__device__ __forceinline__ float oc_fista_c(int it) {
  // same recurrence as fista_c() of img_myula.cu: t_k = t_{k-1} + 1/2 (sic)
  if (it < 1) return 0.0f;
  double told = 1.0 + 0.5 * (it - 1), t = (1.0 + sqrt(4.0 * told * told)) / 2.0;
  return (float)((told - 1.0) / t);
}

// sol = x - gamma * div(rr, ss), div = -grad^T, accumulated in the order of sol_at() (img_myula.cu)
__device__ __forceinline__ float oc_sol(float xv, float rr, float rr_up, float ss, float ss_left, float gamma) {
  float t = ((-rr + rr_up) - ss) + ss_left;
  return xv - gamma * (-t);
}

// transpose of a 4x4 block held as 4 registers in each of 4 consecutive lanes
__device__ __forceinline__ void oc_transpose4(float v[4], int m) {
#pragma unroll
  for (int s = 2; s >= 1; s >>= 1) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if ((k & s) == 0) {
        const bool up = (m & s) != 0;
        float send = up ? v[k] : v[k | s];
        float recv = __shfl_xor_sync(0xffffffffu, send, s);
        if (up) v[k] = recv; else v[k | s] = recv;
      }
    }
  }
}

// Per-thread view of the CTA's shared memory and of the neighbours' windows.
template <int RPC>
struct OcCtx {
  using L = OcLayout<RPC>;
  float* xs;
  float* pool;
  float* xs_prev;
  float* xs_next;
  float* pool_prev;
  float* pool_next;
  int rank, c, h, r0, a0, ny, nx;
  bool colin, has_right;
  __device__ __forceinline__ float& XS(int a, int cc) const { return xs[(a + OC_HB) * OC_PITCH + OC_PADL + cc]; }
  // pool row `row` (buffers store local row a at row base + a + 1; scratch images at base + a + OC_HB)
  __device__ __forceinline__ float& PL(int row, int cc) const { return pool[row * OC_PITCH + OC_PADL + cc]; }
};

// One TV dual sweep on the thread's strip.  Reads (rr_it, ss_it) from buffer `cur`, leaves
// p_{it+1} in (P, Q) and writes rr_{it+1} = p_{it+1} + cn (p_{it+1} - p_it) into buffer `nxt`
// (own rows, plus the halo rows of the neighbour CTAs).  Returns the thread's part of the
// objective 1/2 ||x - sol||^2 + gamma sum |grad sol|.  FIRST: rr_0 = ss_0 = 0, sol_0 = x.
template <int RPC, bool FIRST>
__device__ __forceinline__ double oc_tv_sweep(const OcCtx<RPC>& K, int cur_rr, int cur_ss, int nxt_rr, int nxt_ss,
                                              float cn, float gamma, float inv_g8,
                                              float (&P)[RPC / 2], float (&Q)[RPC / 2]) {
  constexpr int STRIP = RPC / 2;
  const int c = K.c, a0 = K.a0;
  double acc = 0.0;
  float accf = 0.0f;
  float x_c = K.XS(a0, c), rr_c = 0.0f, ss_c = 0.0f, rr_ur = 0.0f, sol_c;
  if (FIRST) {
    sol_c = x_c;
  } else {
    rr_c = K.PL(cur_rr + a0 + 1, c);
    ss_c = K.PL(cur_ss + a0 + 1, c);
    sol_c = oc_sol(x_c, rr_c, K.PL(cur_rr + a0, c), ss_c, K.PL(cur_ss + a0 + 1, c - 1), gamma);
    rr_ur = K.PL(cur_rr + a0, c + 1);
  }
#pragma unroll
  for (int k = 0; k < STRIP; ++k) {
    const int a = a0 + k, i = K.r0 + a;
    const float x_b = K.XS(a + 1, c), x_r = K.XS(a, c + 1);
    float sol_b, sol_r, rr_b = 0.0f, ss_b = 0.0f;
    if (FIRST) {
      sol_b = x_b;
      sol_r = x_r;
    } else {
      rr_b = K.PL(cur_rr + a + 2, c);
      ss_b = K.PL(cur_ss + a + 2, c);
      sol_b = oc_sol(x_b, rr_b, rr_c, ss_b, K.PL(cur_ss + a + 2, c - 1), gamma);
      const float rr_r = K.PL(cur_rr + a + 1, c + 1);
      sol_r = oc_sol(x_r, rr_r, rr_ur, K.PL(cur_ss + a + 1, c + 1), ss_c, gamma);
      rr_ur = rr_r;
    }
    const bool in = K.colin && i < K.ny;
    const float dx = (i < K.ny - 1) ? sol_b - sol_c : 0.0f;
    const float dy = K.has_right ? sol_r - sol_c : 0.0f;
    const float nrm = lmc_sqrtf(fmaf(dx, dx, dy * dy));
    if (in) {
      if (!FIRST) {
        const double df = (double)(x_c - sol_c);
        acc = fma(0.5 * df, df, acc);
      }
      accf += nrm;
    }
    if ((k & 3) == 3) {          // |grad sol| summed in fp32 over 4 pixels, then in fp64
      acc = fma((double)gamma, (double)accf, acc);
      accf = 0.0f;
    }
    // projected dual update: r = rr - dx/(8 gamma); p = r / max(1, |(r,s)|)
    const float r = fmaf(-dx, inv_g8, rr_c), s = fmaf(-dy, inv_g8, ss_c);
    const float n2 = fmaf(r, r, s * s);
    const float sc = n2 > 1.0f ? lmc_rsqrtf(n2) : 1.0f;
    float p = r * sc, q = s * sc;
    float rrn, ssn;
    if (FIRST) {
      rrn = p;   // c_1 = 0 and p_0 = 0
      ssn = q;
    } else {
      rrn = fmaf(cn, p - P[k], p);
      ssn = fmaf(cn, q - Q[k], q);
    }
    if (!in) p = q = rrn = ssn = 0.0f;
    P[k] = p;
    Q[k] = q;
    K.PL(nxt_rr + a + 1, c) = rrn;
    K.PL(nxt_ss + a + 1, c) = ssn;
    if (k == 0 && K.h == 0) {          // first own row -> below-halo of the previous CTA
      if (K.rank > 0) {
        K.pool_prev[(nxt_rr + RPC + 1) * OC_PITCH + OC_PADL + c] = rrn;
        K.pool_prev[(nxt_ss + RPC + 1) * OC_PITCH + OC_PADL + c] = ssn;
      } else {
        K.PL(nxt_rr + 0, c) = 0.0f;    // row -1 of the image
      }
    }
    if (k == STRIP - 1 && K.h == 1) {  // last own row -> above-halo of the next CTA
      if (K.rank < OC_CL - 1) {
        K.pool_next[(nxt_rr + 0) * OC_PITCH + OC_PADL + c] = rrn;
      } else {
        K.PL(nxt_rr + RPC + 1, c) = 0.0f;
        K.PL(nxt_ss + RPC + 1, c) = 0.0f;
      }
    }
    x_c = x_b;
    sol_c = sol_b;
    rr_c = rr_b;
    ss_c = ss_b;
  }
  return acc;
}

// g = H^T mask(H x - b) by four separable passes through the scratch images, fused with the
// noise draw:  base = (1 - tau/gamma) x - tau sigma_L2 g + sqrt(2 tau) xi      (algs.py:569)
template <int RPC>
__device__ __forceinline__ void oc_blur_base(const OcCtx<RPC>& K, const OcArgs& A, long long chain, long long n,
                                             unsigned long long it_global, float (&base)[RPC / 2]) {
  using L = OcLayout<RPC>;
  constexpr int STRIP = RPC / 2;
  const Geom<float>& g = A.g;
  const int c = K.c, h = K.h, a0 = K.a0, r0 = K.r0, ny = K.ny, nx = K.nx;
  const int ey = g.ntap_y - 1, hi_y = g.lo_y + g.ntap_y - 1;
  // A: forward row pass  XS -> T1 on rows [-ey, RPC+ey)
  for (int a = -ey + h; a < RPC + ey; a += 2) {
    float s = 0.0f;
    for (int t = 0; t < g.ntap_x; ++t) s = fmaf(g.wx[t], K.XS(a, c + g.lo_x + t), s);
    K.PL(L::T1 + a + OC_HB, c) = s;
  }
  __syncthreads();
  // B: forward column pass - b, masked to the image  T1 -> T2 on rows [-hi_y, RPC-lo_y)
  {
    const float* bc = A.b + chain * A.b_stride;
    for (int a = -hi_y + h; a < RPC - g.lo_y; a += 2) {
      const int i = r0 + a;
      float s = 0.0f;
      if (K.colin && i >= 0 && i < ny) {
        const float bv = __ldg(bc + (long long)i * nx + c);
        for (int t = 0; t < g.ntap_y; ++t) s = fmaf(g.wy[t], K.PL(L::T1 + a + g.lo_y + t + OC_HB, c), s);
        s -= bv;
      }
      K.PL(L::T2 + a + OC_HB, c) = s;
    }
  }
  __syncthreads();
  // C: adjoint row pass  T2 -> T1 (same rows)
  for (int a = -hi_y + h; a < RPC - g.lo_y; a += 2) {
    float s = 0.0f;
    for (int t = 0; t < g.ntap_x; ++t) s = fmaf(g.wx[t], K.PL(L::T2 + a + OC_HB, c - g.lo_x - t), s);
    K.PL(L::T1 + a + OC_HB, c) = s;
  }
  __syncthreads();
  // D: adjoint column pass on the own strip
  const int quads_per_row = (nx + 3) / 4;
#pragma unroll
  for (int k4 = 0; k4 < STRIP; k4 += 4) {
    float z[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (A.noise == nullptr) {
      const int m = c & 3;
      img_noise4<float>(A.key, A.chain_offset + (unsigned long long)chain, it_global, r0 + a0 + k4 + m, c >> 2,
                        quads_per_row, z);
      oc_transpose4(z, m);
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const int k = k4 + kk, a = a0 + k, i = r0 + a;
      float s = 0.0f;
      for (int t = 0; t < g.ntap_y; ++t) s = fmaf(g.wy[t], K.PL(L::T1 + a - g.lo_y - t + OC_HB, c), s);
      float xi = z[kk];
      if (A.noise != nullptr && K.colin && i < ny)
        xi = __ldg(A.noise + (chain * A.N + n) * g.P + (long long)i * nx + c);
      base[k] = A.a_x * K.XS(a, c) - A.tau * (A.sigma_l2 * s) + A.s2t * xi;
    }
  }
}

template <int RPC>
__global__ void __launch_bounds__(OC_NT, 1) img_myula_onchip_kernel(const __grid_constant__ OcArgs A) {
  using L = OcLayout<RPC>;
  constexpr int STRIP = L::STRIP;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ double red[OC_NT / 32];
  __shared__ double cpart[2][OC_CL];

  cg::cluster_group cluster = cg::this_cluster();
  const long long cid = blockIdx.x / OC_CL, ncl = gridDim.x / OC_CL;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const Geom<float>& g = A.g;
  OcCtx<RPC> K;
  K.xs = reinterpret_cast<float*>(smem_raw);
  K.pool = K.xs + L::XS_ROWS * OC_PITCH;
  K.rank = (int)cluster.block_rank();
  K.c = tid & (OC_W - 1);
  K.h = tid >> 8;
  K.ny = g.ny;
  K.nx = g.nx;
  K.r0 = K.rank * RPC;
  K.a0 = K.h * STRIP;
  K.colin = K.c < g.nx;
  K.has_right = K.c < g.nx - 1;
  const int rank = K.rank, c = K.c, h = K.h, r0 = K.r0, a0 = K.a0, ny = g.ny, nx = g.nx;
  const float gamma = A.tv_gamma, inv_g8 = A.inv_g8;

  for (int idx = tid; idx < (L::XS_ROWS + L::POOL_ROWS) * OC_PITCH; idx += OC_NT) K.xs[idx] = 0.0f;
  // neighbour windows (DSMEM)
  K.xs_prev = rank > 0 ? cluster.map_shared_rank(K.xs, rank - 1) : nullptr;
  K.xs_next = rank < OC_CL - 1 ? cluster.map_shared_rank(K.xs, rank + 1) : nullptr;
  K.pool_prev = rank > 0 ? cluster.map_shared_rank(K.pool, rank - 1) : nullptr;
  K.pool_next = rank < OC_CL - 1 ? cluster.map_shared_rank(K.pool, rank + 1) : nullptr;
  cluster.sync();   // every CTA of the cluster is resident and zeroed before the first push

  unsigned long long sweeps_done = 0;

  for (long long chain = cid; chain < A.C; chain += ncl) {
    // ---- load x (own rows + blur halo rows straight from HBM; zero outside the image)
    {
      const float* xc = A.x + chain * g.P;
      for (int a = -OC_HB + h; a < RPC + OC_HB; a += 2) {
        const int i = r0 + a;
        float v = 0.0f;
        if (i >= 0 && i < ny && K.colin) v = xc[(long long)i * nx + c];
        K.XS(a, c) = v;
      }
    }
    __syncthreads();

    for (long long n = 0; n < A.N; ++n) {
      float P[STRIP], Q[STRIP], base[STRIP];
      double prev_obj = 0.0;
      int kf = 0;
      const unsigned long long it_global = A.iter_offset + (unsigned long long)n;

      // ============================================================ TV sweeps
      // even sweeps write buffer X, odd sweeps buffer Y (= the blur scratch, free by then)
      for (int it = 0;; ++it) {
        const int cur_rr = (it & 1) ? L::X_RR : L::Y_RR, cur_ss = (it & 1) ? L::X_SS : L::Y_SS;
        const int nxt_rr = (it & 1) ? L::Y_RR : L::X_RR, nxt_ss = (it & 1) ? L::Y_SS : L::X_SS;
        double acc;
        if (it == 0) {
          acc = oc_tv_sweep<RPC, true>(K, cur_rr, cur_ss, nxt_rr, nxt_ss, 0.0f, gamma, inv_g8, P, Q);
        } else {
          acc = oc_tv_sweep<RPC, false>(K, cur_rr, cur_ss, nxt_rr, nxt_ss, oc_fista_c(it + 1), gamma, inv_g8, P, Q);
        }
        acc = lmc_warp_sum(acc);
        if (lane == 0) red[wid] = acc;
        // the data-term gradient does not depend on the TV loop: it runs between sweep 0 and its barrier
        if (it == 0) oc_blur_base<RPC>(K, A, chain, n, it_global, base);
        __syncthreads();    // red[] complete (and, on sweep 0, the scratch reads are done)
        if (tid < OC_CL) {
          double s = 0.0;
#pragma unroll
          for (int w = 0; w < OC_NT / 32; ++w) s += red[w];
          double* dst = cluster.map_shared_rank(&cpart[it & 1][0], tid);
          dst[rank] = s;
        }
        cluster.sync();     // halo rows of buffer nxt and all 8 partials have landed
        double obj = 0.0;
#pragma unroll
        for (int r = 0; r < OC_CL; ++r) obj += cpart[it & 1][r];
        const double prev = it == 0 ? 0.0 : prev_obj;
        const double rel = obj > 1e-10 ? fabs(obj - prev) / obj : 2.0;
        prev_obj = obj;
        kf = it;
        ++sweeps_done;
        const bool stop = rel < A.tv_rtol || it == A.tv_niter - 1;
        if (stop || it + 1 >= A.tv_niter) break;
      }

      // ============================================================ update
      // prox = sol_kf = x - gamma div(rr_kf);  x <- base + (tau/gamma) prox; halo rows pushed to the neighbours
      {
        const int cur_rr = (kf & 1) ? L::X_RR : L::Y_RR, cur_ss = (kf & 1) ? L::X_SS : L::Y_SS;
        float rr_up = kf > 0 ? K.PL(cur_rr + a0, c) : 0.0f;
#pragma unroll
        for (int k = 0; k < STRIP; ++k) {
          const int a = a0 + k, i = r0 + a;
          const float xv = K.XS(a, c);
          float sol = xv;
          if (kf > 0) {
            const float rr = K.PL(cur_rr + a + 1, c);
            sol = oc_sol(xv, rr, rr_up, K.PL(cur_ss + a + 1, c), K.PL(cur_ss + a + 1, c - 1), gamma);
            rr_up = rr;
          }
          const float xn = fmaf(A.tog, sol, base[k]);
          base[k] = xn;
          if (K.colin && i < ny) {
            K.XS(a, c) = xn;
            if (a < OC_HB && rank > 0) K.xs_prev[(RPC + a + OC_HB) * OC_PITCH + OC_PADL + c] = xn;
            if (a >= RPC - OC_HB && rank < OC_CL - 1) K.xs_next[(a - RPC + OC_HB) * OC_PITCH + OC_PADL + c] = xn;
          }
        }
        const bool stats = (A.mean != nullptr) && ((long long)it_global >= A.burnin);
        if (K.colin && (A.samples != nullptr || stats)) {
          const long long o0 = (long long)(r0 + a0) * nx + c;
          if (A.samples != nullptr) {
            float* sp = A.samples + (chain * A.N + n) * g.P + o0;
#pragma unroll
            for (int k = 0; k < STRIP; ++k)
              if (r0 + a0 + k < ny) sp[(long long)k * nx] = base[k];
          }
          if (stats) {
            const float cntn = (float)(A.stat_count0 + ((long long)it_global - A.burnin) + 1);
            float* mp = A.mean + chain * g.P + o0;
            float mv[STRIP];
#pragma unroll
            for (int k = 0; k < STRIP; ++k) mv[k] = (r0 + a0 + k < ny) ? mp[(long long)k * nx] : 0.0f;
            if (A.m2 != nullptr) {
              float* m2p = A.m2 + chain * g.P + o0;
              float m2v[STRIP];
#pragma unroll
              for (int k = 0; k < STRIP; ++k) m2v[k] = (r0 + a0 + k < ny) ? m2p[(long long)k * nx] : 0.0f;
#pragma unroll
              for (int k = 0; k < STRIP; ++k) {
                const float d = base[k] - mv[k];
                const float mnew = mv[k] + d / cntn;
                if (r0 + a0 + k < ny) m2p[(long long)k * nx] = m2v[k] + d * (base[k] - mnew);
              }
            }
#pragma unroll
            for (int k = 0; k < STRIP; ++k) {
              const float d = base[k] - mv[k];
              if (r0 + a0 + k < ny) mp[(long long)k * nx] = mv[k] + d / cntn;
            }
          }
        }
      }
      cluster.sync();   // new x (own rows + halo rows) visible; dual buffers free for the next iteration
    }

    // ---- store the chain state
    {
      float* xc = A.x + chain * g.P;
#pragma unroll
      for (int k = 0; k < STRIP; ++k) {
        const int a = a0 + k, i = r0 + a;
        if (K.colin && i < ny) xc[(long long)i * nx + c] = K.XS(a, c);
      }
    }
    __syncthreads();   // XS is reloaded by the next chain
  }
  if (rank == 0 && tid == 0 && sweeps_done) atomicAdd(A.counters, sweeps_done);
  cluster.sync();      // no CTA exits while a neighbour may still push into its shared memory
}

template <int RPC>
int launch_onchip(const OcArgs& A, cudaStream_t st) {
  using L = OcLayout<RPC>;
  auto kern = img_myula_onchip_kernel<RPC>;
  LMC_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::SMEM_BYTES));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = OC_CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(OC_NT);
  cfg.dynamicSmemBytes = L::SMEM_BYTES;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cfg.gridDim = dim3(OC_CL);
  int max_clusters = 0;
  LMC_CUDA_OK(cudaOccupancyMaxActiveClusters(&max_clusters, kern, &cfg));
  LMC_REQUIRE(max_clusters >= 1, "on-chip MYULA kernel: no cluster of %d CTAs with %zu B shared memory fits", OC_CL,
              L::SMEM_BYTES);
  long long ncl = A.C < max_clusters ? A.C : max_clusters;
  cfg.gridDim = dim3((unsigned)(ncl * OC_CL));
  LMC_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, A));
  return 0;
}

}  // namespace

extern "C" int lmc_img_myula_onchip_supported(const lmc_img_problem* pb, int dtype) {
  if (pb == nullptr || dtype != LMC_F32) return 0;
  if (pb->ny < 1 || pb->nx < 1 || pb->ny > OC_CL * 32 || pb->nx > OC_W) return 0;
  if (pb->ntap_y < 1 || pb->ntap_x < 1 || pb->ntap_y > OC_HB + 1 || pb->ntap_x > OC_HB + 1) return 0;
  return 1;
}

extern "C" int lmc_img_myula_onchip_clusters(void) {
  auto kern = img_myula_onchip_kernel<32>;
  using L = OcLayout<32>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::SMEM_BYTES) != cudaSuccess) {
    lmc_set_error("cudaFuncSetAttribute failed: %s", cudaGetErrorString(cudaGetLastError()));
    return -2;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = OC_CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(OC_NT);
  cfg.gridDim = dim3(OC_CL);
  cfg.dynamicSmemBytes = L::SMEM_BYTES;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) {
    lmc_set_error("cudaOccupancyMaxActiveClusters failed: %s", cudaGetErrorString(cudaGetLastError()));
    return -2;
  }
  return n;
}

extern "C" int lmc_img_myula_run_onchip(const lmc_img_problem* pb, const lmc_img_myula_params* pr, int dtype,
                                        int64_t C, int64_t N, uint64_t seed, uint64_t chain_offset,
                                        const lmc_img_io* io, void* stream) {
  if (check_problem(pb, "lmc_img_myula_run_onchip") != 0) return -1;
  LMC_REQUIRE(pr != nullptr && io != nullptr, "lmc_img_myula_run_onchip: null descriptor");
  LMC_REQUIRE(lmc_img_myula_onchip_supported(pb, dtype),
              "lmc_img_myula_run_onchip: needs LMC_F32, ny <= %d, nx <= %d, blur taps <= %d (use lmc_img_myula_run)",
              OC_CL * 32, OC_W, OC_HB + 1);
  LMC_REQUIRE(C >= 0 && N >= 0, "lmc_img_myula_run_onchip: negative C or N");
  LMC_REQUIRE(pr->tau > 0 && pr->gamma > 0, "lmc_img_myula_run_onchip: tau and gamma must be positive");
  LMC_REQUIRE(pr->tv_niter >= 0 && pr->tv_sigma * pr->epsg * pr->gamma > 0,
              "lmc_img_myula_run_onchip: TV needs niter >= 0 and sigma*epsg*gamma > 0");
  if (C == 0 || N == 0) return 0;
  LMC_REQUIRE(io->d_x != nullptr && pb->d_b != nullptr, "lmc_img_myula_run_onchip: d_x and d_b are required");
  LMC_REQUIRE(pb->b_chain_stride == 0 || pb->b_chain_stride == (int64_t)pb->ny * pb->nx,
              "lmc_img_myula_run_onchip: b_chain_stride must be 0 or ny*nx");
  LMC_REQUIRE(io->d_counters != nullptr, "lmc_img_myula_run_onchip: d_counters is required");
  OcArgs A;
  memset(&A, 0, sizeof(A));
  fill_geom(pb, A.g);
  A.x = (float*)io->d_x;
  A.b = (const float*)pb->d_b;
  A.b_stride = pb->b_chain_stride;
  A.noise = (const float*)io->d_noise;
  A.samples = (float*)io->d_samples;
  A.mean = (float*)io->d_mean;
  A.m2 = (float*)io->d_m2;
  A.counters = io->d_counters;
  A.C = C;
  A.N = N;
  A.burnin = io->burnin;
  A.stat_count0 = io->stat_count0;
  A.chain_offset = chain_offset;
  A.iter_offset = io->iter_offset;
  A.key = img_key(seed);
  A.sigma_l2 = (float)pb->sigma_l2;
  A.tau = (float)pr->tau;
  A.a_x = (float)(1 - pr->tau / pr->gamma);
  A.tog = (float)(pr->tau / pr->gamma);
  A.s2t = (float)sqrt(2 * pr->tau);
  const double tvg = pr->tv_sigma * (pr->epsg * pr->gamma);
  A.tv_gamma = (float)tvg;
  A.inv_g8 = (float)(1.0 / (8.0 * (double)A.tv_gamma));
  A.tv_niter = pr->tv_niter;
  A.tv_rtol = pr->tv_rtol;
  return launch_onchip<32>(A, (cudaStream_t)stream);
}
