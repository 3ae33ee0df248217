// Host side of img_ulpda.cu: workspace layout, argument marshalling, C entry points.
#include "img_host_common.inl"

namespace {

template <typename real>
size_t ulpda_ws_bytes(int ny, int nx, long long C) {
  const size_t P = (size_t)ny * nx;
  const size_t tiles = (size_t)((ny + TH - 1) / TH) * ((nx + TW - 1) / TW);
  return 256 + align_up(C * sizeof(LsqrState)) + align_up((size_t)3 * C * tiles * sizeof(double)) +
         4 * align_up((size_t)C * P * sizeof(real)) /* optb, u, v, w */;
}

template <typename real>
int run_ulpda(const lmc_img_problem* pb, const lmc_img_ulpda_params* pr, long long C, long long N, uint64_t seed,
              uint64_t chain_offset, const lmc_img_io* io, int mode, const void* prox_in, void* prox_out,
              cudaStream_t st) {
  UlpdaArgs<real> A;
  memset(&A, 0, sizeof(A));
  fill_geom(pb, A.g);
  const size_t need = ulpda_ws_bytes<real>(pb->ny, pb->nx, C);
  LMC_REQUIRE(io->d_workspace != nullptr && io->workspace_bytes >= need, "image workspace too small (%zu < %zu)",
              io->workspace_bytes, need);
  char* ws = (char*)io->d_workspace;
  A.n_active = (int*)ws;
  unsigned long long* local_counters = (unsigned long long*)(ws + 64);
  ws += 256;
  A.st = (LsqrState*)ws;
  ws += align_up(C * sizeof(LsqrState));
  A.partial = (double*)ws;
  ws += align_up((size_t)3 * C * A.g.tiles * sizeof(double));
  const size_t arr = align_up((size_t)C * A.g.P * sizeof(real));
  A.optb = (real*)ws; ws += arr;
  A.u = (real*)ws; ws += arr;
  A.v = (real*)ws; ws += arr;
  A.w = (real*)ws; ws += arr;
  A.counters = io->d_counters ? io->d_counters : local_counters;
  A.C = C;
  A.N = N;
  A.mode = mode;
  A.key = img_key(seed);
  A.chain_offset = chain_offset;
  A.iter_offset = io->iter_offset;
  A.burnin = io->burnin;
  A.stat_count0 = io->stat_count0;
  A.b = (const real*)pb->d_b;
  A.b_stride = pb->b_chain_stride;
  A.warm = (real*)io->d_warm;
  A.sigma_l2 = (real)pb->sigma_l2;
  A.tau = (real)pr->tau;
  A.cA = (real)(pr->tau * pb->sigma_l2);
  A.iter_lim = pr->lsqr_iter_lim;
  A.warm_on = pr->warm;
  A.atol = pr->lsqr_atol;
  A.btol = pr->lsqr_btol;
  A.ctol = pr->lsqr_conlim > 0 ? 1.0 / pr->lsqr_conlim : 0.0;
  if (mode == 0) {
    A.x = (real*)io->d_x;
    A.xhat = (real*)io->d_xhat;
    A.y = (real*)io->d_y;
    A.zlin = (const real*)io->d_z;
    A.noise = (const real*)io->d_noise;
    A.samples = (real*)io->d_samples;
    A.ysamples = (real*)io->d_ysamples;
    A.mean = (real*)io->d_mean;
    A.m2 = (real*)io->d_m2;
    A.mu = (real)pr->mu;
    A.theta = (real)pr->theta;
    A.s2t = (real)sqrt(2 * pr->tau);
    A.g_sigma = (real)pr->g_sigma;
    A.isotropic = pr->g_isotropic;
    A.gfirst = pr->gfirst;
  } else {
    A.in = (const real*)prox_in;
    A.out = (real*)prox_out;
  }
  LMC_CUDA_OK(cudaMemsetAsync(io->d_workspace, 0, 256, st));
  return launch_persistent(img_ulpda_kernel<real>, A, C * A.g.tiles, (size_t)3 * RS * RS * sizeof(real), st);
}

}  // namespace

extern "C" size_t lmc_img_ulpda_workspace_bytes(int dtype, int32_t ny, int32_t nx, int64_t C) {
  if (ny < 1 || nx < 1 || C < 1) return 256;
  return dtype == LMC_F64 ? ulpda_ws_bytes<double>(ny, nx, C) : ulpda_ws_bytes<float>(ny, nx, C);
}

extern "C" int lmc_img_ulpda_run(const lmc_img_problem* pb, const lmc_img_ulpda_params* pr, int dtype, int64_t C,
                                 int64_t N, uint64_t seed, uint64_t chain_offset, const lmc_img_io* io,
                                 void* stream) {
  if (check_problem(pb, "lmc_img_ulpda_run") != 0) return -1;
  LMC_REQUIRE(pr != nullptr && io != nullptr, "lmc_img_ulpda_run: null descriptor");
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_ulpda_run: dtype must be LMC_F32 or LMC_F64");
  LMC_REQUIRE(C >= 0 && N >= 0, "lmc_img_ulpda_run: negative C or N");
  LMC_REQUIRE(pr->tau > 0 && pr->mu > 0 && pr->g_sigma > 0, "lmc_img_ulpda_run: tau, mu, g_sigma must be positive");
  if (C == 0 || N == 0) return 0;
  LMC_REQUIRE(io->d_x && io->d_xhat && io->d_y && io->d_warm && pb->d_b,
              "lmc_img_ulpda_run: d_x, d_xhat, d_y, d_warm and d_b are required");
  LMC_REQUIRE(pb->b_chain_stride == 0 || pb->b_chain_stride == (int64_t)pb->ny * pb->nx,
              "lmc_img_ulpda_run: b_chain_stride must be 0 or ny*nx");
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == LMC_F64 ? run_ulpda<double>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st)
                          : run_ulpda<float>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st);
}

extern "C" int lmc_img_l2_prox(const lmc_img_problem* pb, int dtype, double tau, int32_t iter_lim, double atol,
                               double btol, double conlim, int64_t C, const void* d_in, void* d_warm, void* d_out,
                               unsigned long long* d_counters, void* d_workspace, size_t workspace_bytes,
                               void* stream) {
  if (check_problem(pb, "lmc_img_l2_prox") != 0) return -1;
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_l2_prox: bad dtype");
  LMC_REQUIRE(C >= 0 && tau > 0, "lmc_img_l2_prox: bad arguments");
  if (C == 0) return 0;
  LMC_REQUIRE(d_in && d_warm && d_out && pb->d_b, "lmc_img_l2_prox: null buffer");
  lmc_img_ulpda_params pr;
  memset(&pr, 0, sizeof(pr));
  pr.tau = tau;
  pr.lsqr_iter_lim = iter_lim;
  pr.warm = 1;
  pr.lsqr_atol = atol;
  pr.lsqr_btol = btol;
  pr.lsqr_conlim = conlim;
  lmc_img_io io;
  memset(&io, 0, sizeof(io));
  io.d_warm = d_warm;
  io.d_workspace = d_workspace;
  io.workspace_bytes = workspace_bytes;
  io.d_counters = d_counters;
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == LMC_F64 ? run_ulpda<double>(pb, &pr, C, 1, 0, 0, &io, 1, d_in, d_out, st)
                          : run_ulpda<float>(pb, &pr, C, 1, 0, 0, &io, 1, d_in, d_out, st);
}
