// Host side of img_ulpda.cu: workspace layout, argument marshalling, C entry points.
#include "img_host_common.inl"

// The band kernels are instantiated per tap count in their own translation units (img_ulpda_band{5,6,7}.cu, compiled
// in parallel): dtype and w (image columns per thread) are resolved there.  A / B are UlpdaArgs / MyulaArgs<float|double>.
int lmc_ulpda_band_launch_nt5(int dtype, void* A, const void* B, int w, cudaStream_t st);
int lmc_ulpda_band_launch_nt6(int dtype, void* A, const void* B, int w, cudaStream_t st);
int lmc_ulpda_band_launch_nt7(int dtype, void* A, const void* B, int w, cudaStream_t st);

namespace {

// rows per band of the band kernel (img_ulpda_band.inl)
#ifndef LMC_ULPDA_BAND_RB
#define LMC_ULPDA_BAND_RB 16
#endif
constexpr int BAND_RB = LMC_ULPDA_BAND_RB;

template <typename real>
size_t ulpda_ws_bytes(int ny, int nx, long long C) {
  const size_t P = (size_t)ny * nx;
  size_t tiles = (size_t)((ny + TH - 1) / TH) * ((nx + TW - 1) / TW);
  const size_t bands = (size_t)(ny + BAND_RB - 1) / BAND_RB;   // the band kernel keeps one partial per band
  if (bands > tiles) tiles = bands;
  return 256 + align_up(C * sizeof(LsqrState)) + align_up((size_t)3 * C * tiles * sizeof(double)) +
         4 * align_up((size_t)C * P * sizeof(real)) /* optb, u, v, w */;
}

// extra workspace of an in-kernel non-convex data term: arg_w, arg2_w; ME-TV adds the state of the TV pipeline
// (n_active, per-chain TvState, per-tile objective partials, three rotating dual pairs)
template <typename real>
size_t ulpda_ncvx_ws_bytes(int ny, int nx, long long C, int kind) {
  if (kind == 0) return 0;
  const size_t P = (size_t)ny * nx, arr = align_up((size_t)C * P * sizeof(real));
  size_t n = 2 * arr;
  if (kind == 2) {
    const size_t tiles = (size_t)((ny + TH - 1) / TH) * ((nx + TW - 1) / TW);
    n += 256 + align_up(C * sizeof(TvState)) + align_up((size_t)C * tiles * sizeof(double)) + 6 * arr;
  }
  return n;
}

template <typename real>
bool band_supported(const lmc_img_problem* pb) {
  if (pb->ntap_y != pb->ntap_x || pb->ntap_y < 5 || pb->ntap_y > 7) return false;   // H5 / H6 / H7
  const size_t smem = Band<real, 7, BAND_RB>::smem_bytes(band_pitch(pb->nx));
  return smem <= 200 * 1024;
}

// two columns per thread: the tap windows of both passes must start on even columns and every array the
// band phases touch with vector accesses must be aligned to two elements
template <typename real>
bool band_pairs(const UlpdaArgs<real>& A) {
  const int lo = A.g.ntap_x == 7 ? BandWin<7>::LO : BandWin<5>::LO;   // the driver's k x k windows
  if (A.g.lo_x != lo || (A.g.nx & 1)) return false;
  const void* ptrs[] = {A.x, A.y, A.warm, A.optb, A.u, A.v, A.w, A.in, A.b};
  for (const void* p : ptrs)
    if (reinterpret_cast<size_t>(p) % (2 * sizeof(real)) != 0) return false;
  return true;
}

template <typename real>
int run_ulpda(const lmc_img_problem* pb, const lmc_img_ulpda_params* pr, long long C, long long N, uint64_t seed,
              uint64_t chain_offset, const lmc_img_io* io, int mode, const void* prox_in, void* prox_out,
              cudaStream_t st, bool band = false, long long group = 0) {
  UlpdaArgs<real> A;
  MyulaArgs<real> B;
  memset(&A, 0, sizeof(A));
  memset(&B, 0, sizeof(B));
  B.mode = -1;
  fill_geom(pb, A.g);
  if (band) {   // one partial per band of BAND_RB rows
    A.g.tiles_x = 1;
    A.g.tiles_y = (pb->ny + BAND_RB - 1) / BAND_RB;
    A.g.tiles = A.g.tiles_y;
    A.pitch = band_pitch(pb->nx);
    // group <= 0: all chains in lockstep.  Measured on B200 (512x512, 256 chains): L2-sized groups of 16-48
    // chains lose more to the extra grid barriers and to partial waves than they gain from L2 hits.
    if (group <= 0) group = C;
    A.group = group < C ? group : C;
  }
  const int nc_kind = (mode == 0 && io->ncvx != nullptr) ? io->ncvx->kind : 0;
  const size_t need = ulpda_ws_bytes<real>(pb->ny, pb->nx, C) + ulpda_ncvx_ws_bytes<real>(pb->ny, pb->nx, C, nc_kind);
  LMC_REQUIRE(io->d_workspace != nullptr && io->workspace_bytes >= need, "image workspace too small (%zu < %zu)",
              io->workspace_bytes, need);
  char* ws = (char*)io->d_workspace;
  A.n_active = (int*)ws;
  A.queue = (int*)(ws + 128);
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
  if (nc_kind != 0) {   // algs.L2_ncvx_tv.prox evaluated in-kernel (algs.py:201-223)
    const lmc_img_ncvx& nc = *io->ncvx;
    A.arg_w = (real*)ws; ws += arr;
    A.arg2_w = (real*)ws; ws += arr;
    A.nc.kind = nc_kind;
    A.nc.W.inv_gamma = (real)(1.0 / nc.gamma);
    A.nc.W.gamma = (real)nc.gamma;
    A.nc.W.iso = nc.isotropic;
    A.nc.post_div = real(1);
    // grouping of the scalars as the reference writes them: x += tau * lamda * m (isotropic),
    // x += tau * lamda / gamma * m (anisotropic MC-TV and ME-TV)
    A.nc_scale = (real)(nc.isotropic && nc_kind == 1 ? pr->tau * nc.lamda : pr->tau * nc.lamda / nc.gamma);
    if (nc_kind == 2) {   // TV(dims, 1., niter, rtol).prox(arg, gamma) by the tile TV pipeline
      fill_geom(pb, B.g);
      B.mode = 3;
      B.C = C;
      B.N = N;
      B.x = A.arg_w;
      B.out = A.arg2_w;
      B.me_scale = A.nc_scale;
      B.tv_gamma = (real)nc.gamma;
      B.tv_objw = (double)B.tv_gamma * 1.0;
      B.tv_passes = nc.niter + LMC_TV_EXTRA_PASS;
      B.tv_rtol = nc.rtol;
      B.n_active = (int*)ws;
      B.counters = (unsigned long long*)(ws + 64);
      ws += 256;
      B.st = (TvState*)ws; ws += align_up(C * sizeof(TvState));
      B.partial = (double*)ws; ws += align_up((size_t)C * B.g.tiles * sizeof(double));
      B.D = (real*)ws; ws += 6 * arr;
      A.group = C;   // the TV pipeline advances all chains together
    }
  }
  A.C = C;
  A.N = N;
  A.mode = mode;
  A.key = img_key(seed);
  A.chain_offset = chain_offset;
  A.iter_offset = io->iter_offset;
  A.burnin = io->burnin;
  // samples folded in by this call before global iteration `it` = it - max(burnin, iter_offset); the kernels evaluate
  // count = stat_base + (it - burnin) + 1, so stat_base absorbs the part of iter_offset beyond the burn-in
  A.stat_count0 = io->stat_count0 - ((long long)io->iter_offset > io->burnin ? (long long)io->iter_offset - io->burnin : 0);
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
    A.in = nc_kind != 0 ? A.arg2_w : (const real*)io->d_prox_arg;
  } else {
    A.in = (const real*)prox_in;
    A.out = (real*)prox_out;
  }
  LMC_CUDA_OK(cudaMemsetAsync(io->d_workspace, 0, 256, st));
  if (band) {
    const int dt = sizeof(real) == 8 ? LMC_F64 : LMC_F32, w = band_pairs(A) ? 2 : 1;
    switch (pb->ntap_y) {
      case 5: return lmc_ulpda_band_launch_nt5(dt, &A, &B, w, st);
      case 6: return lmc_ulpda_band_launch_nt6(dt, &A, &B, w, st);
      default: return lmc_ulpda_band_launch_nt7(dt, &A, &B, w, st);
    }
  }
  {   // the second kernel argument has its own type: launch by hand
    int dev = 0, sms = 0, occ = 0;
    const size_t smem = (size_t)(B.mode == 3 ? 4 : 3) * RS * RS * sizeof(real);
    auto kern = img_ulpda_kernel<real>;
    LMC_CUDA_OK(cudaGetDevice(&dev));
    LMC_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    LMC_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    LMC_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NTHREADS, smem));
    LMC_REQUIRE(occ >= 1, "ULPDA kernel does not fit on an SM (smem %zu)", smem);
    const long long cap = (long long)sms * occ, items = C * A.g.tiles;
    int grid = (int)(items < cap ? items : cap);
    if (grid < 1) grid = 1;
    void* kargs[] = {(void*)&A, (void*)&B};
    LMC_CUDA_OK(cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(NTHREADS), kargs, smem, st));
    return 0;
  }
}

}  // namespace

extern "C" size_t lmc_img_ulpda_workspace_bytes(int dtype, int32_t ny, int32_t nx, int64_t C) {
  if (ny < 1 || nx < 1 || C < 1) return 256;
  return dtype == LMC_F64 ? ulpda_ws_bytes<double>(ny, nx, C) : ulpda_ws_bytes<float>(ny, nx, C);
}

extern "C" size_t lmc_img_ulpda_workspace_bytes_ncvx(int dtype, int32_t ny, int32_t nx, int64_t C, int32_t kind) {
  if (ny < 1 || nx < 1 || C < 1) return 256;
  return lmc_img_ulpda_workspace_bytes(dtype, ny, nx, C) +
         (dtype == LMC_F64 ? ulpda_ncvx_ws_bytes<double>(ny, nx, C, kind) : ulpda_ncvx_ws_bytes<float>(ny, nx, C, kind));
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
  LMC_REQUIRE(io->d_prox_arg == nullptr || (N == 1 && !pr->gfirst),
              "lmc_img_ulpda_run: d_prox_arg needs N == 1 and gfirst == 0");
  LMC_REQUIRE(io->ncvx == nullptr || io->ncvx->kind == 0 || io->d_prox_arg == nullptr,
              "lmc_img_ulpda_run: ncvx and d_prox_arg are two ways to pass the same term");
  LMC_REQUIRE(io->ncvx == nullptr || (io->ncvx->kind >= 0 && io->ncvx->kind <= 2 && (io->ncvx->kind == 0 ||
              (io->ncvx->gamma > 0 && (io->ncvx->kind == 1 || (io->ncvx->isotropic && io->ncvx->niter >= 0))))),
              "lmc_img_ulpda_run: bad ncvx descriptor (ME-TV needs isotropic = 1)");
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == LMC_F64 ? run_ulpda<double>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st)
                          : run_ulpda<float>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st);
}

extern "C" int lmc_img_ulpda_band_supported(const lmc_img_problem* pb, int dtype) {
  if (pb == nullptr || pb->ny < 1 || pb->nx < 1) return 0;
  if (dtype == LMC_F32) return band_supported<float>(pb) ? 1 : 0;
  if (dtype == LMC_F64) return band_supported<double>(pb) ? 1 : 0;
  return 0;
}

extern "C" int lmc_img_ulpda_run_band(const lmc_img_problem* pb, const lmc_img_ulpda_params* pr, int dtype, int64_t C,
                                      int64_t N, uint64_t seed, uint64_t chain_offset, int64_t group,
                                      const lmc_img_io* io, void* stream) {
  if (check_problem(pb, "lmc_img_ulpda_run_band") != 0) return -1;
  LMC_REQUIRE(pr != nullptr && io != nullptr, "lmc_img_ulpda_run_band: null descriptor");
  LMC_REQUIRE(lmc_img_ulpda_band_supported(pb, dtype),
              "lmc_img_ulpda_run_band: needs 5/6/7 blur taps on both axes and a row that fits in shared memory "
              "(use lmc_img_ulpda_run)");
  LMC_REQUIRE(C >= 0 && N >= 0, "lmc_img_ulpda_run_band: negative C or N");
  LMC_REQUIRE(pr->tau > 0 && pr->mu > 0 && pr->g_sigma > 0, "lmc_img_ulpda_run_band: tau, mu, g_sigma must be positive");
  if (C == 0 || N == 0) return 0;
  LMC_REQUIRE(io->d_x && io->d_xhat && io->d_y && io->d_warm && pb->d_b,
              "lmc_img_ulpda_run_band: d_x, d_xhat, d_y, d_warm and d_b are required");
  LMC_REQUIRE(pb->b_chain_stride == 0 || pb->b_chain_stride == (int64_t)pb->ny * pb->nx,
              "lmc_img_ulpda_run_band: b_chain_stride must be 0 or ny*nx");
  LMC_REQUIRE(io->d_prox_arg == nullptr || (N == 1 && !pr->gfirst),
              "lmc_img_ulpda_run_band: d_prox_arg needs N == 1 and gfirst == 0");
  LMC_REQUIRE(io->ncvx == nullptr || io->ncvx->kind == 0 || io->d_prox_arg == nullptr,
              "lmc_img_ulpda_run_band: ncvx and d_prox_arg are two ways to pass the same term");
  LMC_REQUIRE(io->ncvx == nullptr || (io->ncvx->kind >= 0 && io->ncvx->kind <= 2 && (io->ncvx->kind == 0 ||
              (io->ncvx->gamma > 0 && (io->ncvx->kind == 1 || (io->ncvx->isotropic && io->ncvx->niter >= 0))))),
              "lmc_img_ulpda_run_band: bad ncvx descriptor (ME-TV needs isotropic = 1)");
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == LMC_F64
             ? run_ulpda<double>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st, true, group)
             : run_ulpda<float>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st, true, group);
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
