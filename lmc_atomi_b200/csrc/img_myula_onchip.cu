// MYULA image chains (algs.py:477-587; L2 data term + isotropic TV prior,
// prox_lmc_deconv.py:101-103,122,465-473) with the WHOLE CHAIN STATE ON CHIP.
//
// One thread-block cluster of 8 CTAs owns one chain for all N sampler iterations:
// CTA `rank` holds image rows [rank*RPC, (rank+1)*RPC) (RPC = 32 -> images up to
// 256 x 256, fp32).  Per CTA, in shared memory (row pitch 272 floats, 8 zero pad
// columns each side = the zero boundary of pylops Convolve2D):
//   XS   : x on the CTA's rows plus 6 halo rows above and below (blur halo)
//   pool : two (rr, ss) buffers of the TV dual field, RPC + 2 halo rows each; the
//          second pair doubles as the scratch image of the blur passes
//   BS   : `base` = everything of the MYULA update that does not depend on the TV prox
// and in registers (thread = one column x a strip of RPC/OC_SPLIT rows): the pointwise
// arrays p, q (current projected dual iterate).  Halo rows are PUSHED into the
// neighbour CTA's shared memory (DSMEM) by the thread that produces them; the
// per-chain TV objective is reduced across the cluster through 8 doubles per CTA.
// One cluster barrier per TV sweep + one per sampler iteration; chains exit the
// TV loop individually (no batch-wide lockstep).  HBM traffic per iteration: the
// observation b (L2 resident when shared), the running mean / samples if requested.
//
// The blur H^T(Hx - b) is two fused separable passes with compile-time tap counts
// (5, 6, 7 = the reference's H5/H6/H7): a horizontal 1-D filter out of shared memory
// feeding a vertical sliding window held in registers.
//
// Arithmetic is the same recipe as the streaming kernel img_myula.cu (pyproximal
// TV.prox as restated in SURVEY 8a I4) with fp32 shortcuts that stay far inside
// the 1e-5 trajectory tolerance: dx/(8 gamma) is a multiply by the host-rounded
// reciprocal, r / max(1,|(r,s)|) is r * min(1, rsqrt(r^2+s^2)), and the objective
// is summed in fp32 over groups of 4 pixels before it enters the fp64 accumulator.
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
#ifndef OC_SPLIT
#define OC_SPLIT 2
#endif
constexpr int OC_NT = OC_W * OC_SPLIT; // threads per CTA: 256 columns x OC_SPLIT row strips

template <int RPC>
struct OcLayout {
  static constexpr int STRIP = RPC / OC_SPLIT;
  static constexpr int XS_ROWS = RPC + 2 * OC_HB;
  static constexpr int DROWS = RPC + 2;             // halo row above, own rows, halo row below
  static constexpr int T_ROWS = RPC + 2 * OC_HB;    // blur scratch image (residual), local row a at T2 + a + OC_HB
  // pool rows.  Buffer Y (odd TV sweeps write it) aliases the blur scratch; buffer X
  // (written by sweep 0, which runs before the blur of the same iteration) does not.
  static constexpr int Y_RR = 0, Y_SS = DROWS;
  static constexpr int T2 = 0;
  static constexpr int X_RR = (2 * DROWS > T_ROWS) ? 2 * DROWS : T_ROWS;
  static constexpr int X_SS = X_RR + DROWS;
  static constexpr int POOL_ROWS = X_SS + DROWS;
  static constexpr int PADDED_FLOATS = (XS_ROWS + POOL_ROWS) * OC_PITCH;
  static constexpr int BS_FLOATS = RPC * OC_W;      // pointwise: no pad columns
  static constexpr size_t SMEM_BYTES = (size_t)(PADDED_FLOATS + BS_FLOATS) * sizeof(float);
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
  int tv_passes;    // niter + LMC_TV_EXTRA_PASS
  double tv_rtol;
  double tv_objw;   // gamma * sigma_TV: weight of sum |grad sol| in the objective of the exit test
};

// c_it = (t_{it-1} - 1) / t_it with t_k = t_{k-1} + 1/2, t_0 = 1 (fista_c() of img_myula.cu,
// the "sic" recurrence of SURVEY I4) in closed form: (it - 1) / (it + 2).
__device__ __forceinline__ float oc_fista_c(int it) {
  return it < 1 ? 0.0f : (float)(it - 1) / (float)(it + 2);
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

// split cluster barrier: everything stored before the arrive (shared memory of any CTA of the
// cluster included) is visible to the cluster's threads after the matching wait
__device__ __forceinline__ void oc_cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void oc_cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// Per-thread view of the CTA's shared memory and of the neighbours' windows.
template <int RPC>
struct OcCtx {
  float* xs;
  float* pool;
  float* bs;
  float* xs_prev;
  float* xs_next;
  float* pool_prev;
  float* pool_next;
  int rank, c, h, r0, a0, ny, nx;
  bool colin, has_right;
  __device__ __forceinline__ float& XS(int a, int cc) const { return xs[(a + OC_HB) * OC_PITCH + OC_PADL + cc]; }
  // pool row `row` (dual buffers store local row a at row base + a + 1; the scratch image at base + a + OC_HB)
  __device__ __forceinline__ float& PL(int row, int cc) const { return pool[row * OC_PITCH + OC_PADL + cc]; }
  __device__ __forceinline__ float& BS(int a) const { return bs[a * OC_W + c]; }
};

// One TV dual sweep on the thread's strip.  Reads (rr_it, ss_it) from buffer `cur`, leaves
// p_{it+1} in (P, Q) and writes rr_{it+1} = p_{it+1} + cn (p_{it+1} - p_it) into buffer `nxt`
// (own rows, plus the halo rows of the neighbour CTAs).  Returns the thread's part of the
// objective 1/2 ||x - sol||^2 + gamma sum |grad sol|.  FIRST: rr_0 = ss_0 = 0, sol_0 = x.
// FULL: the image fills the cluster's 8*RPC x 256 footprint (no in-image masks).
template <int RPC, bool FIRST, bool FULL>
__device__ __forceinline__ double oc_tv_sweep(const OcCtx<RPC>& K, int cur_rr, int cur_ss, int nxt_rr, int nxt_ss,
                                              float cn, float gamma, double objw, float inv_g8,
                                              float (&P)[RPC / OC_SPLIT], float (&Q)[RPC / OC_SPLIT]) {
  constexpr int STRIP = RPC / OC_SPLIT;
  const int c = K.c, a0 = K.a0;
  double acc = 0.0;
  float accn = 0.0f, accd = 0.0f;
  float x_c = K.XS(a0, c), rr_c = 0.0f, ss_c = 0.0f, rr_ur = 0.0f, sol_c;
  if (FIRST) {
    sol_c = x_c;
  } else {
    rr_c = K.PL(cur_rr + a0 + 1, c);
    ss_c = K.PL(cur_ss + a0 + 1, c);
    sol_c = oc_sol(x_c, rr_c, K.PL(cur_rr + a0, c), ss_c, K.PL(cur_ss + a0 + 1, c - 1), gamma);
    rr_ur = K.PL(cur_rr + a0, c + 1);
  }
  const bool last_cta = K.rank == OC_CL - 1;
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
    bool in, has_below;
    if (FULL) {
      in = true;
      has_below = !(k == STRIP - 1 && K.h == OC_SPLIT - 1 && last_cta);
    } else {
      in = K.colin && i < K.ny;
      has_below = i < K.ny - 1;
    }
    const float dx = has_below ? sol_b - sol_c : 0.0f;
    const float dy = K.has_right ? sol_r - sol_c : 0.0f;
    const float nrm = lmc_sqrtf(fmaf(dx, dx, dy * dy));
    if (in) {
      if (!FIRST) {
        const float df = x_c - sol_c;
        accd = fmaf(df, df, accd);
      }
      accn += nrm;
    }
    if ((k & 3) == 3) {          // fp32 over 4 pixels, then fp64
      acc = fma(objw, (double)accn, acc);
      if (!FIRST) acc = fma(0.5, (double)accd, acc);
      accn = accd = 0.0f;
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
    if (k == STRIP - 1 && K.h == OC_SPLIT - 1) {  // last own row -> above-halo of the next CTA
      if (!last_cta) {
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

// g = H^T mask(H x - b) in two fused separable passes (NT taps per axis, compile time), each a
// horizontal filter out of shared memory feeding a vertical sliding window in registers:
//   AB: x -> residual  (scratch image T2, rows [-hi_y, RPC - lo_y) of the CTA)
//   CD: residual -> g on the thread's own strip, fused with the noise draw:
//       base = (1 - tau/gamma) x - tau sigma_L2 g + sqrt(2 tau) xi      (algs.py:569)  -> BS
// the observation rows a thread's residual rows need (issued early: the loads fly during sweep 0)
template <int RPC, int NT, bool FULL>
__device__ __forceinline__ void oc_load_b(const OcCtx<RPC>& K, const OcArgs& A, long long chain,
                                          float (&bv)[(RPC + NT - 1 + OC_SPLIT - 1) / OC_SPLIT]) {
  constexpr int TOT = RPC + NT - 1, HALF = (TOT + OC_SPLIT - 1) / OC_SPLIT;
  const int hi_y = A.g.lo_y + NT - 1;
  const float* bc = A.b + chain * A.b_stride;
  const int j0 = K.h * HALF;
#pragma unroll
  for (int o = 0; o < HALF; ++o) {
    const int i = K.r0 - hi_y + j0 + o;
    const bool in = (FULL || K.colin) && i >= 0 && i < K.ny && j0 + o < TOT;
    bv[o] = in ? __ldg(bc + (long long)i * K.nx + K.c) : 0.0f;
  }
}

template <int RPC, int NT, bool FULL>
__device__ __forceinline__ void oc_blur_base(const OcCtx<RPC>& K, const OcArgs& A, long long chain, long long n,
                                             unsigned long long it_global,
                                             const float (&bv)[(RPC + NT - 1 + OC_SPLIT - 1) / OC_SPLIT]) {
  using L = OcLayout<RPC>;
  constexpr int STRIP = RPC / OC_SPLIT;
  constexpr int TOT = RPC + NT - 1, HALF = (TOT + OC_SPLIT - 1) / OC_SPLIT;
  const Geom<float>& g = A.g;
  const int c = K.c, a0 = K.a0, r0 = K.r0, ny = K.ny, nx = K.nx;
  const int lo_y = g.lo_y, lo_x = g.lo_x, hi_y = lo_y + NT - 1;
  float wx[NT], wy[NT];
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    wx[t] = g.wx[t];
    wy[t] = g.wy[t];
  }
  // ---- AB
  {
    const int j0 = K.h * HALF;
    const int aout0 = -hi_y + j0;          // first residual row of this thread
    const int rin0 = aout0 + lo_y;         // first x row it reads
    const float* xrow = &K.XS(rin0, c + lo_x);
    float win[NT];
#pragma unroll
    for (int jj = 0; jj < HALF + NT - 1; ++jj) {
      float t1 = 0.0f;
      if (rin0 + jj < RPC + OC_HB) {
#pragma unroll
        for (int s = 0; s < NT; ++s) t1 = fmaf(wx[s], xrow[jj * OC_PITCH + s], t1);
      }
      win[jj % NT] = t1;
      if (jj >= NT - 1) {
        const int o = jj - (NT - 1), a = aout0 + o, i = r0 + a;
        if (j0 + o < TOT) {
          float res = 0.0f;
          const bool in = FULL ? (i >= 0 && i < ny) : (K.colin && i >= 0 && i < ny);
          if (in) {
#pragma unroll
            for (int t = 0; t < NT; ++t) res = fmaf(wy[t], win[(o + t) % NT], res);
            res -= bv[o];
          }
          K.PL(L::T2 + a + OC_HB, c) = res;
        }
      }
    }
  }
  __syncthreads();
  // ---- CD
  {
    const int rin0 = a0 - hi_y;
    const float* rrow = &K.PL(L::T2 + rin0 + OC_HB, c - lo_x);
    const int quads_per_row = (nx + 3) / 4;
    const float* nz = A.noise != nullptr ? A.noise + (chain * A.N + n) * g.P + (long long)r0 * nx + c : nullptr;
    float win[NT];
    float z[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
    for (int jj = 0; jj < STRIP + NT - 1; ++jj) {
      float u = 0.0f;
#pragma unroll
      for (int s = 0; s < NT; ++s) u = fmaf(wx[s], rrow[jj * OC_PITCH - s], u);
      win[jj % NT] = u;
      if (jj >= NT - 1) {
        const int k = jj - (NT - 1), a = a0 + k, i = r0 + a;
        if ((k & 3) == 0 && nz == nullptr) {
          const int m = c & 3;
          img_noise4<float>(A.key, A.chain_offset + (unsigned long long)chain, it_global, i + m, c >> 2,
                            quads_per_row, z);
          oc_transpose4(z, m);
        }
        float gv = 0.0f;
#pragma unroll
        for (int t = 0; t < NT; ++t) gv = fmaf(wy[t], win[(k + NT - 1 - t) % NT], gv);
        float xi = z[k & 3];
        if (nz != nullptr && K.colin && i < ny) xi = __ldg(nz + (long long)a * nx);
        K.BS(a) = A.a_x * K.XS(a, c) - A.tau * (A.sigma_l2 * gv) + A.s2t * xi;
      }
    }
  }
}

template <int RPC, int NT, bool FULL>
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
  K.bs = K.xs + L::PADDED_FLOATS;
  K.rank = (int)cluster.block_rank();
  K.c = tid & (OC_W - 1);
  K.h = tid >> 8;
  K.ny = g.ny;
  K.nx = g.nx;
  K.r0 = K.rank * RPC;
  K.a0 = K.h * STRIP;
  K.colin = FULL ? true : K.c < g.nx;
  K.has_right = K.c < g.nx - 1;
  const int rank = K.rank, c = K.c, h = K.h, r0 = K.r0, a0 = K.a0, ny = g.ny, nx = g.nx;
  const float gamma = A.tv_gamma, inv_g8 = A.inv_g8;

  for (int idx = tid; idx < L::PADDED_FLOATS; idx += OC_NT) K.xs[idx] = 0.0f;
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
      for (int a = -OC_HB + h; a < RPC + OC_HB; a += OC_SPLIT) {
        const int i = r0 + a;
        float v = 0.0f;
        if (i >= 0 && i < ny && K.colin) v = xc[(long long)i * nx + c];
        K.XS(a, c) = v;
      }
    }
    __syncthreads();

    for (long long n = 0; n < A.N; ++n) {
      float P[STRIP], Q[STRIP];
      float bv[(RPC + NT - 1 + OC_SPLIT - 1) / OC_SPLIT];
      oc_load_b<RPC, NT, FULL>(K, A, chain, bv);
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
          acc = oc_tv_sweep<RPC, true, FULL>(K, cur_rr, cur_ss, nxt_rr, nxt_ss, 0.0f, gamma, A.tv_objw, inv_g8, P, Q);
        } else {
          acc = oc_tv_sweep<RPC, false, FULL>(K, cur_rr, cur_ss, nxt_rr, nxt_ss, oc_fista_c(it + 1), gamma, A.tv_objw,
                                              inv_g8, P, Q);
        }
        acc = lmc_warp_sum(acc);
        if (lane == 0) red[wid] = acc;
        // the data-term gradient does not depend on the TV loop: it runs between sweep 0 and its barrier
        if (it == 0) oc_blur_base<RPC, NT, FULL>(K, A, chain, n, it_global, bv);
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
        // rel = |obj - prev| / obj < rtol  (2 when obj <= 1e-10: never stops)
        const bool small = obj > 1e-10 && fabs(obj - prev) < A.tv_rtol * obj;
        prev_obj = obj;
        kf = it;
        ++sweeps_done;
        if (small || it + 1 >= A.tv_passes) break;
      }

      // ============================================================ update
      // prox = sol_kf = x - gamma div(rr_kf);  x <- base + (tau/gamma) prox; halo rows pushed to the neighbours
      {
        const bool stats = (A.mean != nullptr) && ((long long)it_global >= A.burnin);
        const long long o0 = (long long)(r0 + a0) * nx + c;
        float* const mp = A.mean + chain * g.P + o0;
        float mv[STRIP];
        if (stats && K.colin) {
#pragma unroll
          for (int k = 0; k < STRIP; ++k) mv[k] = (FULL || r0 + a0 + k < ny) ? mp[(long long)k * nx] : 0.0f;
        }
        const int cur_rr = (kf & 1) ? L::X_RR : L::Y_RR, cur_ss = (kf & 1) ? L::X_SS : L::Y_SS;
        float xn[STRIP];
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
          xn[k] = fmaf(A.tog, sol, K.BS(a));
          if (FULL || (K.colin && i < ny)) {
            K.XS(a, c) = xn[k];
            if (a < OC_HB && rank > 0) K.xs_prev[(RPC + a + OC_HB) * OC_PITCH + OC_PADL + c] = xn[k];
            if (a >= RPC - OC_HB && rank < OC_CL - 1) K.xs_next[(a - RPC + OC_HB) * OC_PITCH + OC_PADL + c] = xn[k];
          }
        }
        // x (own rows + pushed halo rows) is complete: arrive now, let the statistics traffic to HBM
        // overlap the barrier latency, wait afterwards
        oc_cluster_arrive();
        if (K.colin && A.samples != nullptr) {
          float* sp = A.samples + (chain * A.N + n) * g.P + o0;
#pragma unroll
          for (int k = 0; k < STRIP; ++k)
            if (FULL || r0 + a0 + k < ny) sp[(long long)k * nx] = xn[k];
        }
        if (stats && K.colin) {
          const float inv_cnt = 1.0f / (float)(A.stat_count0 + ((long long)it_global - A.burnin) + 1);
          if (A.m2 != nullptr) {
            float* m2p = A.m2 + chain * g.P + o0;
            float m2v[STRIP];
#pragma unroll
            for (int k = 0; k < STRIP; ++k) m2v[k] = (FULL || r0 + a0 + k < ny) ? m2p[(long long)k * nx] : 0.0f;
#pragma unroll
            for (int k = 0; k < STRIP; ++k) {
              const float d = xn[k] - mv[k];
              const float mnew = fmaf(d, inv_cnt, mv[k]);
              if (FULL || r0 + a0 + k < ny) m2p[(long long)k * nx] = fmaf(d, xn[k] - mnew, m2v[k]);
            }
          }
#pragma unroll
          for (int k = 0; k < STRIP; ++k) {
            const float d = xn[k] - mv[k];
            if (FULL || r0 + a0 + k < ny) mp[(long long)k * nx] = fmaf(d, inv_cnt, mv[k]);
          }
        }
      }
      oc_cluster_wait();   // new x (own rows + halo rows) visible; dual buffers free for the next iteration
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

cudaLaunchConfig_t oc_config(cudaLaunchAttribute* attr, size_t smem, cudaStream_t st) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = OC_CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(OC_NT);
  cfg.gridDim = dim3(OC_CL);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cfg;
}

template <int RPC, int NT, bool FULL>
int oc_max_clusters(int* out) {
  using L = OcLayout<RPC>;
  auto kern = img_myula_onchip_kernel<RPC, NT, FULL>;
  LMC_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::SMEM_BYTES));
  cudaLaunchAttribute attr[1];
  cudaLaunchConfig_t cfg = oc_config(attr, L::SMEM_BYTES, nullptr);
  LMC_CUDA_OK(cudaOccupancyMaxActiveClusters(out, kern, &cfg));
  return 0;
}

template <int RPC, int NT, bool FULL>
int launch_onchip(const OcArgs& A, cudaStream_t st) {
  using L = OcLayout<RPC>;
  auto kern = img_myula_onchip_kernel<RPC, NT, FULL>;
  int max_clusters = 0;
  if (oc_max_clusters<RPC, NT, FULL>(&max_clusters) != 0) return -2;
  LMC_REQUIRE(max_clusters >= 1, "on-chip MYULA kernel: no cluster of %d CTAs with %zu B shared memory fits", OC_CL,
              L::SMEM_BYTES);
  cudaLaunchAttribute attr[1];
  cudaLaunchConfig_t cfg = oc_config(attr, L::SMEM_BYTES, st);
  const long long ncl = A.C < max_clusters ? A.C : max_clusters;
  cfg.gridDim = dim3((unsigned)(ncl * OC_CL));
  LMC_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, A));
  return 0;
}

template <int NT>
int launch_onchip_nt(const OcArgs& A, cudaStream_t st) {
  const bool full = A.g.ny == OC_CL * 32 && A.g.nx == OC_W;
  return full ? launch_onchip<32, NT, true>(A, st) : launch_onchip<32, NT, false>(A, st);
}

}  // namespace

extern "C" int lmc_img_myula_onchip_supported(const lmc_img_problem* pb, int dtype) {
  if (pb == nullptr || dtype != LMC_F32) return 0;
  if (pb->ny < 1 || pb->nx < 1 || pb->ny > OC_CL * 32 || pb->nx > OC_W) return 0;
  if (pb->ntap_y != pb->ntap_x || pb->ntap_y < 5 || pb->ntap_y > 7) return 0;   // H5 / H6 / H7
  return 1;
}

extern "C" int lmc_img_myula_onchip_clusters(void) {
  int n = 0;
  if (oc_max_clusters<32, 5, true>(&n) != 0) return -2;
  return n;
}

extern "C" int lmc_img_myula_run_onchip(const lmc_img_problem* pb, const lmc_img_myula_params* pr, int dtype,
                                        int64_t C, int64_t N, uint64_t seed, uint64_t chain_offset,
                                        const lmc_img_io* io, void* stream) {
  if (check_problem(pb, "lmc_img_myula_run_onchip") != 0) return -1;
  LMC_REQUIRE(pr != nullptr && io != nullptr, "lmc_img_myula_run_onchip: null descriptor");
  LMC_REQUIRE(lmc_img_myula_onchip_supported(pb, dtype),
              "lmc_img_myula_run_onchip: needs LMC_F32, ny <= %d, nx <= %d, 5/6/7 blur taps on both axes "
              "(use lmc_img_myula_run)", OC_CL * 32, OC_W);
  LMC_REQUIRE(C >= 0 && N >= 0, "lmc_img_myula_run_onchip: negative C or N");
  LMC_REQUIRE(pr->tau > 0 && pr->gamma > 0, "lmc_img_myula_run_onchip: tau and gamma must be positive");
  LMC_REQUIRE(io->ncvx == nullptr || io->ncvx->kind == 0,
              "lmc_img_myula_run_onchip: non-convex data terms run on the band / streaming kernels");
  LMC_REQUIRE(pr->tv_niter >= 0 && pr->tv_sigma * pr->epsg * pr->gamma > 0,
              "lmc_img_myula_run_onchip: TV needs niter >= 0 and sigma*epsg*gamma > 0");
  if (C == 0 || N == 0) return 0;
  LMC_REQUIRE(io->d_x != nullptr && pb->d_b != nullptr, "lmc_img_myula_run_onchip: d_x and d_b are required");
  LMC_REQUIRE(pb->b_chain_stride == 0 || pb->b_chain_stride == (int64_t)pb->ny * pb->nx,
              "lmc_img_myula_run_onchip: b_chain_stride must be 0 or ny*nx");
  LMC_REQUIRE(io->d_counters != nullptr, "lmc_img_myula_run_onchip: d_counters is required");
  LMC_REQUIRE(io->d_fgrad_extra == nullptr, "lmc_img_myula_run_onchip: d_fgrad_extra is not supported (use lmc_img_myula_run)");
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
  // samples folded in by this call before global iteration `it` = it - max(burnin, iter_offset); the kernels evaluate
  // count = stat_base + (it - burnin) + 1, so stat_base absorbs the part of iter_offset beyond the burn-in
  A.stat_count0 = io->stat_count0 - ((long long)io->iter_offset > io->burnin ? (long long)io->iter_offset - io->burnin : 0);
  A.chain_offset = chain_offset;
  A.iter_offset = io->iter_offset;
  A.key = img_key(seed);
  A.sigma_l2 = (float)pb->sigma_l2;
  A.tau = (float)pr->tau;
  A.a_x = (float)(1 - pr->tau / pr->gamma);
  A.tog = (float)(pr->tau / pr->gamma);
  A.s2t = (float)sqrt(2 * pr->tau);
  A.tv_gamma = (float)(pr->tv_sigma * (pr->epsg * pr->gamma));
  A.inv_g8 = (float)(1.0 / (8.0 * (double)A.tv_gamma));
  A.tv_passes = pr->tv_niter + LMC_TV_EXTRA_PASS;
  A.tv_objw = (double)A.tv_gamma * pr->tv_sigma;
  A.tv_rtol = pr->tv_rtol;
  cudaStream_t st = (cudaStream_t)stream;
  switch (pb->ntap_y) {
    case 5: return launch_onchip_nt<5>(A, st);
    case 6: return launch_onchip_nt<6>(A, st);
    default: return launch_onchip_nt<7>(A, st);
  }
}
